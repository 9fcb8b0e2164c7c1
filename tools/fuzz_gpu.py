#!/usr/bin/env python
"""One-off parity hunt on the GPU (not part of the test suite): random trees through every engine against the oracle,
random NNI move sets, random 20-state cases, deep trees, compression and the host-streaming call.  Prints one line per failing seed.

  python tools/fuzz_gpu.py [--seeds 0:600] [--aa 0:200] [--nni 0:100]"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _cases  # noqa: E402
from oracle import workloads  # noqa: E402
from phylo_rs_b200 import capi  # noqa: E402


def rng_of(spec):
    a, b = spec.split(":")
    return range(int(a), int(b))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", default="0:600")
    ap.add_argument("--aa", default="0:200")
    ap.add_argument("--nni", default="0:100")
    ap.add_argument("--deep", default="0:0")
    ap.add_argument("--misc", default="0:0")
    ap.add_argument("--spr", default="0:0")
    args = ap.parse_args()
    lib = capi.default_library()
    bad = 0
    for seed in rng_of(args.seeds):
        try:
            _cases.check_random_trees(lib, [seed])
        except Exception as e:
            bad += 1
            print("FAIL tree", seed, str(e)[:400], flush=True)
    for seed in rng_of(args.aa):
        try:
            _cases.check_random_trees_amino_acid(lib, [seed])
        except Exception as e:
            bad += 1
            print("FAIL aa", seed, str(e)[:400], flush=True)
    for seed in rng_of(args.nni):
        rng = np.random.default_rng(seed)
        n = int(rng.integers(4, 120))
        model = workloads.MODELS[int(rng.integers(len(workloads.MODELS)))]
        case = workloads.make_case(n, int(rng.integers(5, 3000)), model=model, seed=seed, ambiguity=float(rng.choice([0, 0.1])))
        ft = case.ft
        k = int(rng.integers(1, 2 * ft["n_nodes"]))
        moves = [(int(rng.integers(-1, ft["n_nodes"] + 1)), int(rng.integers(2))) for _ in range(k)]
        try:
            _cases.check_nni(lib, case, moves=moves)
        except Exception as e:
            bad += 1
            print("FAIL nni", seed, n, model, str(e)[:400], flush=True)
    # deep trees: hundreds to thousands of taxa, long branches (CLVs shrink by orders of magnitude per level: the
    # host-planned underflow guards of the fused sweep carry the result), caterpillars and Yule shapes, a polytomy on top
    from oracle import parse_newick, refhost

    for seed in rng_of(args.deep):
        rng = np.random.default_rng(seed)
        n = int(rng.integers(150, 2500))
        mean = float(rng.choice([0.05, 0.5, 2.0]))
        model = workloads.MODELS[int(rng.integers(len(workloads.MODELS)))]
        shape = int(rng.integers(3))
        if shape == 0:
            tree = refhost.yule(n, rng)
        else:
            names = [f"t{i}" for i in range(n)]
            s_ = f"({names[0]}:0.1,{names[1]}:0.1)"
            for i in range(2, n):
                s_ = f"({s_}:0.1,{names[i]}:0.1)" if shape == 1 or i % 7 else f"({s_}:0.1,{names[i]}:0.1,x{i}:0.1)"
            tree = parse_newick(s_ + ";")
        for nid in tree.get_node_ids():
            if nid != tree.root:
                tree.set_edge_weight(nid, np.float32(min(max(rng.exponential(mean), 1e-6), 8.0)))
        leaf_order = [tree.nodes[i].taxa for i in tree.get_node_ids() if tree.nodes[i].taxa is not None]
        ft = workloads.flatten_fast(tree, leaf_order)
        m = workloads.make_model(model)
        pats = workloads.simulate_patterns(ft, m, 24, rng)
        case = workloads.Case(tree, ft, m, model, np.ascontiguousarray(pats), np.ones(24, np.uint64), leaf_order)
        try:
            _cases.check_lnl(lib, case)
            _cases.check_lnl(lib, case, flags=capi.FLAG_LEVEL_KERNELS)
            _cases.check_asr(lib, case)
        except Exception as e:
            bad += 1
            print("FAIL deep", seed, n, mean, model, shape, str(e)[:300], flush=True)
    # column compression on random alignments (any byte values, repeated columns, both entry points) against the
    # first-appearance numbering; the host-streaming call against the resident one and the oracle, over random sizes,
    # virtual ranks and CLV budgets
    import oracle

    for seed in rng_of(args.misc):
        rng = np.random.default_rng(seed)
        try:
            # ---- compression: random alignment with repeated columns, any byte values
            n_rows = int(rng.integers(1, 40)); width = int(rng.integers(0, 3000)); distinct = int(rng.integers(1, max(2, width // 3 + 2)))
            alphabet = np.frombuffer(b"ACGT-NRYKM?xz", np.uint8) if seed % 3 else np.arange(256, dtype=np.uint8)
            cols = alphabet[rng.integers(len(alphabet), size=(distinct, n_rows))]
            pick = rng.integers(distinct, size=width)
            mat = np.ascontiguousarray(cols[pick].T) if width else np.zeros((n_rows, 0), np.uint8)
            # reference numbering: first appearance
            idx, s2p_ref, mult_ref, pats_ref = {}, [], [], []
            for i in range(width):
                key = mat[:, i].tobytes()
                j = idx.setdefault(key, len(idx))
                if j == len(pats_ref):
                    pats_ref.append(mat[:, i]); mult_ref.append(0)
                mult_ref[j] += 1; s2p_ref.append(j)
            with capi.Context(library=lib) as ctx:
                if seed % 2:
                    n, s2p, mult = ctx.compress_columns(mat)
                else:
                    n, s2p, mult = ctx.compress_rows([mat[r].tobytes() for r in range(n_rows)])
                assert n == len(pats_ref), ("n_patterns", n, len(pats_ref))
                assert np.array_equal(s2p.astype(np.int64), np.asarray(s2p_ref, np.int64)), "s2p"
                assert np.array_equal(mult.astype(np.int64), np.asarray(mult_ref, np.int64)), "mult"
            # ---- host-streaming lnL == resident lnL == oracle, random sizes, virtual ranks, budgets
            case = workloads.make_case(int(rng.integers(3, 50)), int(rng.integers(1, 2500)), model=workloads.MODELS[int(rng.integers(6))], seed=seed, ambiguity=float(rng.choice([0, 0.1])))
            ref_tot, ref_per = oracle.log_likelihood_flat(case.ft, case.model, case.patterns, case.mult, mode=2, n_threads=8)
            devs = [0] * int(rng.integers(1, 4))
            with capi.Context(devices=devs, library=lib) as ctx:
                if seed % 4 == 0:
                    ctx.set_clv_budget(int(rng.integers(8, 64)) * 1_000_000)
                m, ft = case.model, case.ft
                ctx.set_model(m.eigenvalues, m.eigenvectors, m.pi, m.sqrt_pi, m.inv_sqrt_pi, m.rates, m.weights)
                ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], ft["root"])
                tot_h, per_h = ctx.log_likelihood_host(case.patterns, multiplicity=case.mult, per_pattern=True)
                _cases.assert_lnl_close(tot_h, per_h, ref_tot, ref_per)
                tot_r, per_r = ctx.log_likelihood(per_pattern=True)
                assert np.array_equal(per_r, per_h), "resident vs streamed per-pattern"
                tot_a, st, _ = ctx.marginal_asr()
                assert abs(tot_a - ref_tot) <= 1e-9 * abs(ref_tot)
        except Exception as e:
            bad += 1
            print("FAIL misc", seed, type(e).__name__, str(e)[:300], flush=True)
    # phb_score_trees on random SPR candidates (prune + graft with the reference's semantics: src/tree/ops.rs:34-41)
    from oracle.refhost import Node

    for seed in rng_of(args.spr):
        rng = np.random.default_rng(seed)
        model = workloads.MODELS[int(rng.integers(len(workloads.MODELS)))]
        case = workloads.make_case(int(rng.integers(5, 60)), int(rng.integers(10, 400)), model=model, seed=seed, ambiguity=float(rng.choice([0, 0.1])))
        base = case.tree
        ids = list(base.get_node_ids())
        cands, refs = [], []
        for _ in range(200):
            if len(cands) == 6:
                break
            t = base.clone()
            x, y = int(rng.choice(ids)), int(rng.choice(ids))
            px, py = t.nodes[x].parent, t.nodes[y].parent
            if px is None or py is None or y in set(t.preord_ids(x)):
                continue
            t.delete_edge(px, x)
            new = Node(t.next_id())
            t.split_edge((py, y), new)
            t.set_child(new.id, x)
            ft = workloads.flatten_fast(t, case.leaf_order)
            cands.append((ft["postorder"], ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], ft["root"]))
            refs.append(oracle.log_likelihood_flat(ft, case.model, case.patterns, case.mult, mode=0, n_threads=8)[0])
        try:
            with capi.Context(library=lib) as ctx:
                workloads.load_into(ctx, case)
                before = ctx.log_likelihood()
                got = ctx.score_trees(cands)
                assert ctx.log_likelihood() == before
            for g, r in zip(got, refs):
                assert (abs(g - r) <= 1e-9 * abs(r)) if np.isfinite(r) else (g == r), (g, r)
        except Exception as e:
            bad += 1
            print("FAIL spr", seed, model, str(e)[:300], flush=True)
    print("fuzz done, failures:", bad)


if __name__ == "__main__":
    main()
