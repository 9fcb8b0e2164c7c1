#!/usr/bin/env python
"""Two phb_score_nni calls in a 20-state context (1,024 neighbours of 1,000 taxa x 20k patterns, GTR20+G4) and nothing
else: the subject of the ncu launch list profiles/r2_aa_nni_launches.csv (the second call is the warm one).

  python tools/aa_nni_probe.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from oracle import workloads
    from phylo_rs_b200 import capi

    n, P = 1000, 20_000
    case = workloads.make_case(n, 32, model="aa_g4", seed=0x5EED00AA, ambiguity=0.02)
    m, ft = case.model, case.ft
    rng = np.random.default_rng(7)
    pats = case.patterns[rng.integers(32, size=P)].copy()
    flip = rng.random(pats.shape) < 0.3
    pats[flip] = np.frombuffer(workloads.AA_LETTERS, np.uint8)[rng.integers(20, size=int(flip.sum()))]
    co = ft["child_off"]
    internal = [v for v in range(ft["n_nodes"]) if co[v + 1] > co[v] and v != ft["root"]]
    moves = [(v, lc) for v in internal for lc in (0, 1)][:1024]
    ids, left = [x[0] for x in moves], [x[1] for x in moves]
    with capi.Context(flags=capi.FLAG_AMINO_ACID | capi.FLAG_TIMING) as ctx:
        ctx.set_alignment(pats)
        ctx.set_model(m.eigenvalues, m.eigenvectors, m.pi, m.sqrt_pi, m.inv_sqrt_pi, m.rates, m.weights)
        ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], ft["root"])
        ms = []
        for _ in range(2):
            ctx.timer_start()
            tot = ctx.score_nni(ids, left)
            ms.append(ctx.timer_stop())
        st = ctx.stats()
    print(json.dumps({"case": "20-state NNI, 1,024 neighbours of 1,000 x 20k", "ms": ms, "path_steps": st["nni_path_steps"],
                      "kernel_launches": st["kernel_launches"], "best": float(np.nanmax(tot))}))


if __name__ == "__main__":
    main()
