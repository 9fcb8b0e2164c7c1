#!/usr/bin/env python
"""Timing of the 20-state (AminoAcid) engines on one GPU: log-likelihood and marginal ASR of a Yule tree under a seeded
general 20-state GTR+G4, device-event timed, with the algorithmic bytes and flops per update beside it.

  python tools/aa_bench.py [--taxa 1000] [--patterns 20000] [--reps 3]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--taxa", type=int, default=1000)
    ap.add_argument("--patterns", type=int, default=20000)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--moves", type=int, default=1024)
    args = ap.parse_args()
    from phylo_rs_b200 import capi, synthetic

    tree = synthetic.yule_flat(args.taxa, 0x5EED00AA)
    rng = np.random.default_rng(1)
    # a seeded general 20-state GTR (pi, symmetric exchangeabilities) + discrete Gamma(0.6), 4 categories
    pi = rng.dirichlet(np.full(20, 8.0))
    w = rng.gamma(1.5, 1.0, (20, 20)) + 0.05
    w = (w + w.T) / 2
    np.fill_diagonal(w, 0.0)
    q = w * pi[None, :]
    np.fill_diagonal(q, -q.sum(1))
    q /= -(pi * np.diag(q)).sum()
    sp, isp = np.sqrt(pi), 1.0 / np.sqrt(pi)
    sym = sp[:, None] * q * isp[None, :]
    lam, vec = np.linalg.eigh(0.5 * (sym + sym.T))
    rates = synthetic._discrete_gamma(0.6, 4)
    K = len(rates)
    letters = np.frombuffer(b"ACDEFGHIKLMNPQRSTVWY", np.uint8)
    # columns: one ancestral letter per pattern, each taxon keeps it with probability 0.6 (site-pattern statistics do not
    # change the kernels' work; every value is checked against the oracle in tests/test_amino_acid.py instead)
    base = rng.integers(20, size=(args.patterns, 1))
    keep = rng.random((args.patterns, args.taxa)) < 0.6
    pats = np.ascontiguousarray(letters[np.where(keep, base, rng.integers(20, size=(args.patterns, args.taxa)))])
    with capi.Context(flags=capi.FLAG_AMINO_ACID) as ctx:
        ctx.set_alignment(pats)
        ctx.set_model(lam, vec, pi, sp, isp, rates, np.full(K, 1.0 / K))
        ctx.set_tree(tree["postorder"], tree["child_off"], tree["children"], tree["blen"], tree["leaf_row"], tree["root"])
        ctx.log_likelihood()
        best = {"lnl": 1e30, "asr": 1e30}
        for _ in range(args.reps):
            ctx.timer_start()
            tot = ctx.log_likelihood()
            best["lnl"] = min(best["lnl"], ctx.timer_stop())
            t0 = time.perf_counter()
            ctx.timer_start()
            tot2, states, _ = ctx.marginal_asr()
            best["asr"] = min(best["asr"], ctx.timer_stop())
            wall = (time.perf_counter() - t0) * 1e3
            ctx.timer_start()
            tot3, _jstates = ctx.joint_asr()
            best["joint"] = min(best.get("joint", 1e30), ctx.timer_stop())
        st = ctx.stats()
        # NNI neighbourhood (config 5's shape at 20 states): 1,024 (node, left_ch) moves, each the re-evaluation of the path
        # v -> root against the base tree's CLVs; a handful re-scored in full beside it (PHB_FLAG_NNI_FULL's loop)
        co = tree["child_off"]
        internal = [v for v in range(tree["n_nodes"]) if co[v + 1] > co[v] and v != tree["root"]]
        moves = [(v, left) for v in internal for left in (0, 1)][:args.moves]
        ids, left = [m[0] for m in moves], [m[1] for m in moves]
        ctx.score_nni(ids, left)
        for _ in range(args.reps):
            ctx.timer_start()
            nni = ctx.score_nni(ids, left)
            best["nni"] = min(best.get("nni", 1e30), ctx.timer_stop())
        st_nni = ctx.stats()
        os.environ["PHB_AA_NNI_PATH"] = "1"  # cross-check mode: whole paths v -> root, no upper partials
        ctx.timer_start()
        nni_walk = ctx.score_nni(ids, left)
        walk_ms = ctx.timer_stop()
        del os.environ["PHB_AA_NNI_PATH"]
    with capi.Context(flags=capi.FLAG_AMINO_ACID | capi.FLAG_NNI_FULL) as ctx:
        ctx.set_alignment(pats)
        ctx.set_model(lam, vec, pi, sp, isp, rates, np.full(K, 1.0 / K))
        ctx.set_tree(tree["postorder"], tree["child_off"], tree["children"], tree["blen"], tree["leaf_row"], tree["root"])
        ctx.timer_start()
        nni_full = ctx.score_nni(ids[:8], left[:8])
        full_ms = ctx.timer_stop() / 8
    n_int = args.taxa - 1
    updates = n_int * args.patterns
    # per update (internal node x pattern), K categories: internal children read 160 B each + 4 B exponent, the result
    # written 160 B + 4 B, tip children 1 B; flops: 2 * 400 per internal child mat-vec + 20 per child product
    int_children = n_int - 1
    tip_children = args.taxa
    bytes_lnl = args.patterns * (K * 164.0 * (n_int + int_children) + tip_children)
    flops_lnl = args.patterns * K * (820.0 * int_children + 20.0 * tip_children)
    print(json.dumps({
        "case": f"AminoAcid GTR20+G4 {args.taxa} x {args.patterns}", "lnl": tot, "asr_lnl": tot2,
        "lnl_ms": best["lnl"], "asr_ms": best["asr"], "joint_ms": best["joint"], "joint_value": tot3, "asr_wall_ms": wall, "tiles": st["n_tiles"],
        "updates_per_s": updates / (best["lnl"] * 1e-3), "algorithmic_GBps": bytes_lnl / (best["lnl"] * 1e-3) / 1e9,
        "fp64_TFLOPs": flops_lnl / (best["lnl"] * 1e-3) / 1e12, "states_hist_head": np.bincount(states.ravel(), minlength=20)[:5].tolist(),
        "nni_moves": len(moves), "nni_ms": best["nni"], "nni_trees_per_s": len(moves) / (best["nni"] * 1e-3), "nni_path_steps": st_nni["nni_path_steps"],
        "nni_cache_GB": st_nni["nni_cache_bytes"] / 1e9, "nni_tiles": st_nni["nni_tiles"], "nni_full_ms_per_tree": full_ms,
        "nni_path_mode_ms": walk_ms, "nni_path_mode_equals_full_rescoring": bool(np.array_equal(nni_walk[:8], nni_full)),
        "nni_max_rel_diff_to_path_mode": float(np.nanmax(np.abs(nni - nni_walk) / np.abs(nni_walk))), "nni_best_delta": float(np.nanmax(nni) - tot)}))


if __name__ == "__main__":
    main()
