#!/usr/bin/env python
"""Times BASELINE.json's other configurations (1, 2, 4, 5) and per-rank shards of config 3 on ONE GPU.

bench.py measures the headline configuration; this harness produces the numbers quoted in DESIGN.md for
the remaining rows of SURVEY.md §8 (same synthetic generators, same C ABI).  One JSON line per case.

  python tools/configs_bench.py [--cases cfg1,cfg2,cfg3_rank8,cfg4,cfg5] [--reps 3]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from phylo_rs_b200 import capi, synthetic  # noqa: E402


def model_for(name):
    if name == "jc69":
        return synthetic.gtr_model(pi=(0.25,) * 4, rates=(1.0,) * 6)
    if name == "gtr_g4":
        return synthetic.gtr_model(alpha=0.5, k=4)
    if name == "gtr_i_g4":
        return synthetic.gtr_model(alpha=0.5, k=4, p_inv=0.2)
    raise ValueError(name)


def timed(ctx, fn, reps):
    fn()  # warm-up (allocations, schedule upload)
    fn()
    ms = []
    for _ in range(reps):
        ctx.timer_start()
        t0 = time.perf_counter()
        out = fn()
        dev = ctx.timer_stop()
        ms.append((dev, (time.perf_counter() - t0) * 1e3))
    return out, min(m[0] for m in ms), min(m[1] for m in ms)


def run_lnl(tag, n_taxa, n_sites, model, seed, reps, flags=0):
    tree = synthetic.yule_flat(n_taxa, seed)
    with capi.Context(flags=flags) as ctx:
        synthetic.load_into(ctx, tree, model_for(model))
        ctx.simulate_alignment(n_sites, seed=seed ^ 0xA11C)
        lnl, dev_ms, wall_ms = timed(ctx, ctx.log_likelihood, reps)
        st = ctx.stats()
    upd = (n_taxa - 1) * n_sites
    return dict(case=tag, taxa=n_taxa, patterns=n_sites, model=model, lnl=lnl, device_ms=dev_ms, wall_ms=wall_ms,
                updates_per_s=upd / (wall_ms * 1e-3), algorithmic_GBps=st["algorithmic_bytes"] / (wall_ms * 1e-3) / 1e9,
                sweep_block_sites=st["sweep_block_sites"], sweep_levels=st["sweep_stack_levels"],
                sweep_smem_levels=st["sweep_smem_levels"], path="levels" if flags & capi.FLAG_LEVEL_KERNELS else "sweep")


def run_asr(tag, n_taxa, n_sites, model, seed, reps):
    tree = synthetic.yule_flat(n_taxa, seed)
    K = 4
    with capi.Context(flags=capi.FLAG_TIMING) as ctx:
        synthetic.load_into(ctx, tree, model_for(model))
        ctx.simulate_alignment(n_sites, seed=seed ^ 0xA11C)
        import torch

        states_t = torch.zeros((tree["n_nodes"], n_sites), dtype=torch.uint8, pin_memory=True)  # caller-owned, pinned
        states = states_t.numpy()
        tot = capi.C.c_double(0.0)

        def call():
            ctx._check(ctx.lib.L.phb_marginal_asr(ctx._h, states.ctypes.data_as(capi.C.POINTER(capi.C.c_uint8)), None,
                                                  capi.C.byref(tot)))
            return tot.value

        lnl, dev_ms, wall_ms = timed(ctx, call, reps)
        st = ctx.stats()
    n = n_taxa
    up_bytes = (32.0 * K * (2 * n - 3) + n) * n_sites
    down_bytes = (32.0 * K * (4 * n - 4) + 3 * n - 1) * n_sites
    node_sites = ((n - 1) + (2 * n - 1)) * n_sites
    return dict(case=tag, taxa=n_taxa, patterns=n_sites, model=model, lnl=lnl, device_ms=dev_ms, wall_ms=wall_ms,
                ms_up=st["ms_clv"], ms_down=st["ms_down"], ms_root=st["ms_root"], tiles=st["n_tiles"],
                tile_patterns=st["tile_patterns"], node_site_updates_per_s=node_sites / (wall_ms * 1e-3),
                algorithmic_GBps_device=(up_bytes + down_bytes) / (dev_ms * 1e-3) / 1e9,
                states_d2h_bytes=int(states.nbytes), state_histogram=np.bincount(states[:, :4096].ravel(), minlength=4).tolist())


def run_joint(tag, n_taxa, n_sites, model, seed, reps):
    tree = synthetic.yule_flat(n_taxa, seed)
    K = 4
    with capi.Context(flags=capi.FLAG_TIMING) as ctx:
        synthetic.load_into(ctx, tree, model_for(model))
        ctx.simulate_alignment(n_sites, seed=seed ^ 0xA11C)
        import torch

        states_t = torch.zeros((tree["n_nodes"], n_sites), dtype=torch.uint8, pin_memory=True)
        states = states_t.numpy()
        tot = capi.C.c_double(0.0)

        def call():
            ctx._check(ctx.lib.L.phb_joint_asr(ctx._h, states.ctypes.data_as(capi.C.POINTER(capi.C.c_uint8)), capi.C.byref(tot)))
            return tot.value

        val, dev_ms, wall_ms = timed(ctx, call, reps)
        st = ctx.stats()
        lnl = ctx.log_likelihood()
    n = n_taxa
    up_bytes = (32.0 * K * (2 * n - 3) + n) * n_sites          # as the marginal up pass
    down_bytes = (34.0 * (n - 2) + 3.0 * n + 1) * n_sites       # per child: parent state + category + 4 C values / code + state out
    return dict(case=tag, taxa=n_taxa, patterns=n_sites, model=model, joint_value=val, lnl=lnl, device_ms=dev_ms,
                wall_ms=wall_ms, ms_up=st["ms_clv"], ms_down=st["ms_down"], tiles=st["n_tiles"],
                node_site_updates_per_s=((n - 1) + (2 * n - 1)) * n_sites / (wall_ms * 1e-3),
                algorithmic_GBps_device=(up_bytes + down_bytes) / (dev_ms * 1e-3) / 1e9,
                state_histogram=np.bincount(states[:, :4096].ravel(), minlength=4).tolist())


def run_compress(tag, n_taxa, n_sites, distinct, seed, reps):
    """phb_compress_columns from pinned host rows [row][site]: H2D + hash + sort + verify + gather, end to end."""
    import torch

    tree = synthetic.yule_flat(n_taxa, seed)
    with capi.Context(flags=capi.FLAG_TIMING) as ctx:
        synthetic.load_into(ctx, tree, model_for("gtr_g4"))
        n_sim = distinct or n_sites
        ctx.simulate_alignment(n_sim, seed=seed ^ 0xA11C)
        host = torch.empty((n_taxa, n_sites), dtype=torch.uint8, pin_memory=True)
        h = host.numpy()
        chunk = max(1, (256 << 20) // n_taxa)
        rng = np.random.default_rng(seed)
        pick = rng.integers(0, n_sim, size=n_sites) if distinct else None
        if distinct:
            base = ctx.get_alignment(0, n_sim).T  # [row][distinct]
            for p0 in range(0, n_sites, chunk):
                h[:, p0:p0 + chunk] = base[:, pick[p0:p0 + chunk]]
        else:
            for p0 in range(0, n_sites, chunk):
                c = min(chunk, n_sites - p0)
                h[:, p0:p0 + c] = ctx.get_alignment(p0, c).T
        s2p = np.zeros(n_sites, np.uint64)
        mult = np.zeros(n_sites, np.uint64)
        n = capi.C.c_int64(0)
        u64p = capi.C.POINTER(capi.C.c_uint64)

        def call():
            ctx._check(ctx.lib.L.phb_compress_columns(ctx._h, n_taxa, n_sites, capi.C.c_void_p(host.data_ptr()), capi.C.byref(n),
                                                      s2p.ctypes.data_as(u64p), mult.ctypes.data_as(u64p)))
            return n.value

        P, dev_ms, wall_ms = timed(ctx, call, reps)
        ctx.n_rows, ctx.n_patterns = n_taxa, int(P)
        lnl = ctx.log_likelihood()
    nbytes = n_taxa * n_sites
    return dict(case=tag, rows=n_taxa, sites=n_sites, n_patterns=int(P), wall_ms=wall_ms, device_ms=dev_ms,
                input_GB=nbytes / 1e9, end_to_end_GBps=nbytes / (wall_ms * 1e-3) / 1e9, lnl_after=lnl,
                mult_sum=int(mult[:P].sum()), note="pinned host rows; includes the H2D copy of the whole alignment")


def run_nni(tag, n_taxa, n_sites, model, seed, n_moves, reps, flags=0):
    tree = synthetic.yule_flat(n_taxa, seed)
    co = tree["child_off"]
    internal = [v for v in range(tree["n_nodes"]) if co[v + 1] > co[v] and v != tree["root"]]
    moves = [(v, left) for v in internal for left in (0, 1)][:n_moves]
    with capi.Context(flags=flags | capi.FLAG_TIMING) as ctx:
        synthetic.load_into(ctx, tree, model_for(model))
        ctx.simulate_alignment(n_sites, seed=seed ^ 0xA11C)
        base = ctx.log_likelihood()
        ids, left = [m[0] for m in moves], [m[1] for m in moves]
        out, dev_ms, wall_ms = timed(ctx, lambda: ctx.score_nni(ids, left), 1 if flags & capi.FLAG_NNI_FULL else reps)
        st = ctx.stats()
        t0 = time.perf_counter()
        ctx.log_likelihood()
        full_ms = (time.perf_counter() - t0) * 1e3
    return dict(case=tag, taxa=n_taxa, patterns=n_sites, model=model, n_moves=len(moves), base_lnl=base,
                best_delta=float(np.nanmax(out) - base), wall_ms=wall_ms, device_ms=dev_ms, ms_per_neighbour=wall_ms / len(moves),
                full_evaluation_ms=full_ms, trees_per_s=len(moves) / (wall_ms * 1e-3),
                equivalent_updates_per_s=len(moves) * (n_taxa - 1) * n_sites / (wall_ms * 1e-3),
                path=("full rescoring" if flags & capi.FLAG_NNI_FULL else "walks v->root over cached CLVs" if flags & capi.FLAG_NNI_WALK
                      else "edge-local scoring with upper partials"),
                nni_path_moves=st["nni_path_moves"], nni_path_steps=st["nni_path_steps"], nni_tiles=st["nni_tiles"],
                nni_cache_GB=st["nni_cache_bytes"] / 1e9, nni_read_GB=st["nni_read_bytes"] / 1e9,
                walk_read_GBps=st["nni_read_bytes"] / (dev_ms * 1e-3) / 1e9,
                phases_ms=dict(cache_pass=st["ms_nni_cache"], upper_pass=st["ms_nni_upper"], edge_kernel=st["ms_nni_edge"]))


CASES = {
    "cfg1": lambda r: run_lnl("cfg1 JC69 64 x 10k", 64, 10_000, "jc69", 0x5EED0001, r),
    "cfg2": lambda r: run_lnl("cfg2 GTR+G4 1000 x 100k", 1000, 100_000, "gtr_g4", 0x5EED0002, r),
    "cfg2_levels": lambda r: run_lnl("cfg2 GTR+G4 1000 x 100k (level kernels)", 1000, 100_000, "gtr_g4", 0x5EED0002, r,
                                     capi.FLAG_LEVEL_KERNELS),
    "cfg3_rank2": lambda r: run_lnl("cfg3 per-rank shard at 2 GPUs", 10_000, 500_000, "gtr_i_g4", 0x5EED0003, r),
    "cfg3_rank4": lambda r: run_lnl("cfg3 per-rank shard at 4 GPUs", 10_000, 250_000, "gtr_i_g4", 0x5EED0003, r),
    "cfg3_rank8": lambda r: run_lnl("cfg3 per-rank shard at 8 GPUs", 10_000, 125_000, "gtr_i_g4", 0x5EED0003, r),
    "cfg4": lambda r: run_asr("cfg4 marginal ASR GTR+G4 5000 x 500k", 5000, 500_000, "gtr_g4", 0x5EED0004, r),
    "cfg4_small": lambda r: run_asr("cfg4/10 marginal ASR GTR+G4 5000 x 50k", 5000, 50_000, "gtr_g4", 0x5EED0004, r),
    "joint_small": lambda r: run_joint("joint ASR GTR+G4 5000 x 50k", 5000, 50_000, "gtr_g4", 0x5EED0004, r),
    "joint": lambda r: run_joint("joint ASR GTR+G4 5000 x 500k", 5000, 500_000, "gtr_g4", 0x5EED0004, r),
    "compress_dup": lambda r: run_compress("compress 1000 rows x 1M sites, 50k distinct columns", 1000, 1_000_000, 50_000, 0x5EED0002, r),
    "compress_cfg2": lambda r: run_compress("compress cfg2 1000 rows x 100k sites (all distinct)", 1000, 100_000, 0, 0x5EED0002, r),
    "compress_cfg3": lambda r: run_compress("compress cfg3 10000 rows x 1M sites (all distinct)", 10_000, 1_000_000, 0, 0x5EED0003, r),
    "cfg5": lambda r: run_nni("cfg5 1024 NNI neighbours 2000 x 200k", 2000, 200_000, "gtr_g4", 0x5EED0005, 1024, r),
    "cfg5_walk": lambda r: run_nni("cfg5 1024 NNI neighbours 2000 x 200k (walk kernel)", 2000, 200_000, "gtr_g4", 0x5EED0005, 1024, r,
                                   capi.FLAG_NNI_WALK),
    "cfg5_all": lambda r: run_nni("cfg5+ all 3,996 NNI neighbours 2000 x 200k", 2000, 200_000, "gtr_g4", 0x5EED0005, 4000, r),
    "cfg5_full_128": lambda r: run_nni("cfg5/8 128 NNI neighbours 2000 x 200k, full rescoring", 2000, 200_000, "gtr_g4",
                                       0x5EED0005, 128, r, capi.FLAG_NNI_FULL),
    "cfg5_small": lambda r: run_nni("cfg5/8 128 NNI neighbours 2000 x 200k", 2000, 200_000, "gtr_g4", 0x5EED0005, 128, r),
}


# --------------------------------------------------------------------------------------------------
# bench.py's `configs` object: BASELINE.json configs 2, 4, 5 + joint ASR + compression at full size on one GPU,
# each with its device time, its throughput, the fraction of the measured HBM copy rate its algorithmic bytes
# amount to, and parity deltas against the CPU oracle on a small pattern sample of the SAME data.
# (The oracle is the checker here, never the thing timed.)
# --------------------------------------------------------------------------------------------------
def _oracle_model(name):
    from oracle import workloads

    return workloads.make_model(name)


def _sample_patterns(ctx, n):
    return ctx.get_alignment(0, n)


def _oracle_tree(tree):
    """the flat tree as the oracle wants it (adds the pre-order)"""
    ft = dict(tree)
    pre, st = [], [int(ft["root"])]
    co, ch = ft["child_off"], ft["children"]
    while st:
        v = st.pop()
        pre.append(v)
        st.extend(reversed(ch[co[v]:co[v + 1]].tolist()))
    ft["preorder"] = np.asarray(pre, np.int32)
    return ft


def driver_configs(hbm_gbs, reps=3, which=("cfg2", "cfg4", "joint", "cfg5", "compress", "amino_acid")):
    import oracle

    oracle.build()
    out = {}

    def guard(name, fn):
        if name not in which:
            return
        try:
            out[name] = fn()
        except Exception as e:  # one failing case must not hide the others or the headline
            out[name] = {"error": f"{type(e).__name__}: {e}"}

    def cfg2():
        n, P, seed = 1000, 100_000, 0x5EED0002
        tree = synthetic.yule_flat(n, seed)
        with capi.Context() as ctx:
            synthetic.load_into(ctx, tree, model_for("gtr_g4"))
            ctx.simulate_alignment(P, seed=seed ^ 0xA11C)
            _, dev_ms, wall_ms = timed(ctx, ctx.log_likelihood, reps)
            tot, per = ctx.log_likelihood(per_pattern=True)
            st = ctx.stats()
            pats = _sample_patterns(ctx, 256)
        _r, ref_per = oracle.log_likelihood_flat(_oracle_tree(tree), _oracle_model("gtr_g4"), pats, mode=0, n_threads=8)
        return dict(workload="GTR+G4 log-likelihood, 1000 taxa x 100k patterns", device_ms=dev_ms, wall_ms=wall_ms,
                    updates_per_s=(n - 1) * P / (dev_ms * 1e-3),
                    algorithmic_frac_of_hbm=st["algorithmic_bytes"] / (dev_ms * 1e-3) / 1e9 / hbm_gbs,
                    parity=dict(sample_patterns=256, max_abs_per_pattern_lnl=float(np.max(np.abs(per[:256] - ref_per))),
                                total_equals_sum_rel=float(abs(tot - per.sum()) / abs(tot))))

    def cfg4():
        import torch

        n, P, seed, K = 5000, 500_000, 0x5EED0004, 4
        tree = synthetic.yule_flat(n, seed)
        with capi.Context(flags=capi.FLAG_TIMING) as ctx:
            synthetic.load_into(ctx, tree, model_for("gtr_g4"))
            ctx.simulate_alignment(P, seed=seed ^ 0xA11C)
            states = torch.zeros((tree["n_nodes"], P), dtype=torch.uint8, pin_memory=True).numpy()
            tot = capi.C.c_double(0.0)

            def call():
                ctx._check(ctx.lib.L.phb_marginal_asr(ctx._h, states.ctypes.data_as(capi.C.POINTER(capi.C.c_uint8)), None,
                                                      capi.C.byref(tot)))
                return tot.value

            _, dev_ms, wall_ms = timed(ctx, call, reps)
            st = ctx.stats()
            pats = _sample_patterns(ctx, 64)
        rt, rper, rstates, rpost = oracle.marginal_asr_flat(_oracle_tree(tree), _oracle_model("gtr_g4"), pats, want_posteriors=True, n_threads=8)
        diff = states[:, :64] != rstates
        srt = np.sort(rpost, axis=2)
        gap = srt[..., 3] - srt[..., 2]
        up_bytes = (32.0 * K * (2 * n - 3) + n) * P
        down_bytes = (32.0 * K * (4 * n - 4) + 3 * n - 1) * P
        return dict(workload="marginal ASR GTR+G4, 5000 taxa x 500k patterns (states to pinned host memory)", device_ms=dev_ms,
                    wall_ms=wall_ms, ms_up=st["ms_clv"], ms_down=st["ms_down"],
                    node_site_updates_per_s=((n - 1) + (2 * n - 1)) * P / (dev_ms * 1e-3),
                    algorithmic_frac_of_hbm=(up_bytes + down_bytes) / (dev_ms * 1e-3) / 1e9 / hbm_gbs,
                    d2h_bytes=int(states.nbytes),
                    parity=dict(sample_patterns=64, state_mismatches=int(diff.sum()),
                                state_mismatches_away_from_ties=int((diff & (gap >= 1e-10)).sum()), states_compared=int(diff.size)))

    def joint():
        import torch

        n, P, seed, K = 5000, 500_000, 0x5EED0004, 4
        tree = synthetic.yule_flat(n, seed)
        with capi.Context(flags=capi.FLAG_TIMING) as ctx:
            synthetic.load_into(ctx, tree, model_for("gtr_g4"))
            ctx.simulate_alignment(P, seed=seed ^ 0xA11C)
            states = torch.zeros((tree["n_nodes"], P), dtype=torch.uint8, pin_memory=True).numpy()
            tot = capi.C.c_double(0.0)

            def call():
                ctx._check(ctx.lib.L.phb_joint_asr(ctx._h, states.ctypes.data_as(capi.C.POINTER(capi.C.c_uint8)), capi.C.byref(tot)))
                return tot.value

            _, dev_ms, wall_ms = timed(ctx, call, reps)
            per = ctx.per_pattern()
            pats = _sample_patterns(ctx, 32)
        rt, rper, rstates = oracle.joint_asr_flat(_oracle_tree(tree), _oracle_model("gtr_g4"), pats, n_threads=8)
        up_bytes = (32.0 * K * (2 * n - 3) + n) * P
        down_bytes = (34.0 * (n - 2) + 3.0 * n + 1) * P
        return dict(workload="joint (Viterbi) ASR GTR+G4, 5000 taxa x 500k patterns", device_ms=dev_ms, wall_ms=wall_ms,
                    node_site_updates_per_s=((n - 1) + (2 * n - 1)) * P / (dev_ms * 1e-3),
                    algorithmic_frac_of_hbm=(up_bytes + down_bytes) / (dev_ms * 1e-3) / 1e9 / hbm_gbs,
                    parity=dict(sample_patterns=32, max_abs_per_pattern_value=float(np.max(np.abs(per[:32] - rper))),
                                patterns_with_state_mismatch=int((states[:, :32] != rstates).any(axis=0).sum())))

    def cfg5():
        n, P, seed, n_moves = 2000, 200_000, 0x5EED0005, 1024
        tree = synthetic.yule_flat(n, seed)
        co = tree["child_off"]
        internal = [v for v in range(tree["n_nodes"]) if co[v + 1] > co[v] and v != tree["root"]]
        moves = [(v, left) for v in internal for left in (0, 1)][:n_moves]
        ids, left = [m[0] for m in moves], [m[1] for m in moves]
        with capi.Context(flags=capi.FLAG_TIMING) as ctx:
            synthetic.load_into(ctx, tree, model_for("gtr_g4"))
            ctx.simulate_alignment(P, seed=seed ^ 0xA11C)
            base, base_per = ctx.log_likelihood(per_pattern=True)
            got, dev_ms, wall_ms = timed(ctx, lambda: ctx.score_nni(ids, left), reps)
            st = ctx.stats()
            pats = _sample_patterns(ctx, 64)
        with capi.Context(flags=capi.FLAG_NNI_FULL) as ctx:  # the reference's caller loop: rebuild + full evaluation
            synthetic.load_into(ctx, tree, model_for("gtr_g4"))
            ctx.simulate_alignment(P, seed=seed ^ 0xA11C)
            pick = [0, 1, len(moves) // 2, len(moves) - 1]
            full = ctx.score_nni([ids[i] for i in pick], [left[i] for i in pick])
        _r, ref_per = oracle.log_likelihood_flat(_oracle_tree(tree), _oracle_model("gtr_g4"), pats, mode=0, n_threads=8)
        return dict(workload="1,024 NNI neighbours of a 2000-taxon tree x 200k patterns (GTR+G4), path-only rescoring",
                    device_ms=dev_ms, wall_ms=wall_ms, trees_per_s=len(moves) / (dev_ms * 1e-3),
                    equivalent_updates_per_s=len(moves) * (n - 1) * P / (dev_ms * 1e-3),
                    read_frac_of_hbm=st["nni_read_bytes"] / (dev_ms * 1e-3) / 1e9 / hbm_gbs, cache_GB=st["nni_cache_bytes"] / 1e9,
                    phases_ms=dict(cache_pass=st["ms_nni_cache"], upper_pass=st["ms_nni_upper"], edge_kernel=st["ms_nni_edge"]),
                    parity=dict(base_tree_sample_patterns=64, base_max_abs_per_pattern_lnl=float(np.max(np.abs(base_per[:64] - ref_per))),
                                neighbours_rescored_in_full=len(pick),
                                max_rel_vs_full_rescoring=float(np.max(np.abs(got[pick] - full) / np.abs(full)))))

    def compress():
        import torch

        n, W, seed = 10_000, 1_000_000, 0x5EED0003
        tree = synthetic.yule_flat(n, seed)
        with capi.Context(flags=capi.FLAG_TIMING) as ctx:
            synthetic.load_into(ctx, tree, model_for("gtr_g4"))
            ctx.simulate_alignment(W, seed=seed ^ 0xA11C)
            host = torch.empty((n, W), dtype=torch.uint8, pin_memory=True)
            h = host.numpy()
            chunk = max(1, (256 << 20) // n)
            for p0 in range(0, W, chunk):
                c = min(chunk, W - p0)
                h[:, p0:p0 + c] = ctx.get_alignment(p0, c, layout=capi.LAYOUT_ROW_MAJOR)
            h[:, W - 1000:] = h[:, :1000]  # the last 1,000 columns repeat the first 1,000: known duplicates
            s2p = np.zeros(W, np.uint64)
            mult = np.zeros(W, np.uint64)
            npat = capi.C.c_int64(0)
            u64p = capi.C.POINTER(capi.C.c_uint64)

            def call():
                ctx._check(ctx.lib.L.phb_compress_columns(ctx._h, n, W, capi.C.c_void_p(host.data_ptr()), capi.C.byref(npat),
                                                          s2p.ctypes.data_as(u64p), mult.ctypes.data_as(u64p)))
                return npat.value

            Pn, dev_ms, wall_ms = timed(ctx, call, max(1, reps - 1))
        # first-appearance numbering (alignment.rs:104-120): distinct columns keep their order, the repeats map back
        ok_s2p = bool(np.array_equal(s2p[:W - 1000], np.arange(W - 1000, dtype=np.uint64)) and
                      np.array_equal(s2p[W - 1000:], np.arange(1000, dtype=np.uint64)))
        ok_mult = bool(np.all(mult[:1000] == 2) and np.all(mult[1000:Pn] == 1))
        return dict(workload="compress_columns, 10,000 rows x 1M sites from pinned host rows (999,000 distinct columns)",
                    wall_ms=wall_ms, device_ms=dev_ms, input_GB=n * W / 1e9, end_to_end_GBps=n * W / (wall_ms * 1e-3) / 1e9,
                    parity=dict(n_patterns=int(Pn), expected_patterns=W - 1000, site_to_pattern_exact=ok_s2p, multiplicity_exact=ok_mult))

    def amino_acid():
        """the 20-state engines (PHB_FLAG_AMINO_ACID) on a seeded general GTR20+G4; the oracle's NS=20 build
        checks 32 patterns of the same alignment."""
        from oracle import workloads

        n, P = 1000, 20_000
        case = workloads.make_case(n, 32, model="aa_g4", seed=0x5EED00AA, ambiguity=0.02)
        m, ft = case.model, case.ft
        rng = np.random.default_rng(7)
        pats = case.patterns[rng.integers(32, size=P)].copy()
        flip = rng.random(pats.shape) < 0.3  # decorrelate the resampled columns
        pats[flip] = np.frombuffer(workloads.AA_LETTERS, np.uint8)[rng.integers(20, size=int(flip.sum()))]
        pats[:32] = case.patterns
        with capi.Context(flags=capi.FLAG_AMINO_ACID) as ctx:
            ctx.set_alignment(pats)
            ctx.set_model(m.eigenvalues, m.eigenvectors, m.pi, m.sqrt_pi, m.inv_sqrt_pi, m.rates, m.weights)
            ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], ft["root"])
            _, dev_ms, wall_ms = timed(ctx, ctx.log_likelihood, reps)
            tot, per = ctx.log_likelihood(per_pattern=True)
            (_t, states, _p), asr_ms, _w = timed(ctx, ctx.marginal_asr, max(1, reps - 1))
            # 1,024 NNI neighbours at 20 states (two node updates + a dot product each against resident CLVs / upper partials)
            co = ft["child_off"]
            internal = [v for v in range(ft["n_nodes"]) if co[v + 1] > co[v] and v != ft["root"]]
            moves = [(v, lc) for v in internal for lc in (0, 1)][:1024]
            ids, left = [x[0] for x in moves], [x[1] for x in moves]
            _nn, nni_ms, _w2 = timed(ctx, lambda: ctx.score_nni(ids, left), max(1, reps - 1))
            # ... and eight of them on the 32 sample patterns against the oracle on each rebuilt tree
            ctx.set_alignment(case.patterns)
            sample = moves[::128]
            got_nni = ctx.score_nni([x[0] for x in sample], [x[1] for x in sample])
        nni_rel = 0.0
        for (v, lc), val in zip(sample, got_nni):
            t = case.tree.clone()
            assert not t.nni(int(ft["ids"][v]), bool(lc))
            ref, _ = oracle.log_likelihood_flat(workloads.flatten_fast(t, case.leaf_order), m, case.patterns, mode=2, n_threads=8)
            nni_rel = max(nni_rel, abs(val - ref) / abs(ref))
        _rt, ref_per, ref_states, _rp = oracle.marginal_asr_flat(ft, m, case.patterns, n_threads=8)
        K = len(m.rates)
        flops = P * K * (820.0 * (n - 2) + 20.0 * n)
        return dict(workload="AminoAcid GTR20+G4 log-likelihood, marginal ASR and 1,024 NNI neighbours, 1000 taxa x 20k patterns", device_ms=dev_ms,
                    wall_ms=wall_ms, asr_device_ms=asr_ms, updates_per_s=(n - 1) * P / (dev_ms * 1e-3),
                    fp64_tflops=flops / (dev_ms * 1e-3) / 1e12, frac_of_hbm=P * K * 164.0 * (2 * n - 3) / (dev_ms * 1e-3) / 1e9 / hbm_gbs,
                    nni_1024_device_ms=nni_ms, nni_trees_per_s=len(moves) / (nni_ms * 1e-3),
                    parity=dict(sample_patterns=32, max_abs_per_pattern_lnl=float(np.max(np.abs(per[:32] - ref_per))),
                                asr_state_mismatches=int((states[:, :32] != ref_states).sum()), nni_neighbours_checked=len(sample),
                                nni_max_rel_total=float(nni_rel)))

    guard("cfg2", cfg2)
    guard("amino_acid", amino_acid)
    guard("cfg4", cfg4)
    guard("joint", joint)
    guard("cfg5", cfg5)
    guard("compress", compress)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="cfg1,cfg2,cfg3_rank8,cfg4_small,cfg5")
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    for name in args.cases.split(","):
        try:
            line = CASES[name](args.reps)
        except Exception as e:  # keep going: one failing case must not hide the others
            line = dict(case=name, error=f"{type(e).__name__}: {e}")
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
