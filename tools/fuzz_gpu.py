#!/usr/bin/env python
"""One-off parity hunt on the GPU (not part of the test suite): random trees through every engine against the oracle,
random NNI move sets, random 20-state cases.  Prints one line per failing seed.

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
    print("fuzz done, failures:", bad)


if __name__ == "__main__":
    main()
