"""Worker of tests/test_multirank_gloo.py: one rank of a world_size-N gloo job.  Each rank holds the full tree and
model and the pattern slice bench.shard() gives it, evaluates its partial log-likelihood through the C ABI (the
emulation build: no GPU in this container) and all-reduces the scalar — the data path of bench.py at N > 1."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import bench  # noqa: E402
from oracle import workloads  # noqa: E402
from phylo_rs_b200 import capi  # noqa: E402


def main():
    out_path = sys.argv[1]
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    lib = capi.Library(os.path.join(ROOT, "tests", "emu", "libphylo_b200_emu.so"))
    case = workloads.make_case(11, 1501, model="gtr_i_g4", seed=61, ambiguity=0.05, compress=True)
    P = len(case.patterns)
    lo, cnt = bench.shard(P, world, rank)
    with capi.Context(library=lib) as ctx:
        workloads.load_into(ctx, case)
        ctx.set_alignment(case.patterns[lo:lo + cnt], multiplicity=case.mult[lo:lo + cnt])
        part = ctx.log_likelihood() if cnt else 0.0
        # NNI batch: per-neighbour partial sums are all-reduced the same way (count = n_moves)
        moves = [(v, lc) for v in range(case.ft["n_nodes"]) for lc in (0, 1)]
        nni = ctx.score_nni([m[0] for m in moves], [m[1] for m in moves]) if cnt else np.zeros(len(moves))
    red = torch.tensor([part], dtype=torch.float64)
    dist.all_reduce(red)
    nni_t = torch.from_numpy(np.nan_to_num(np.asarray(nni, np.float64), nan=0.0))
    dist.all_reduce(nni_t)
    counts = torch.tensor([cnt], dtype=torch.int64)
    dist.all_reduce(counts)
    if rank == 0:
        with open(out_path, "w") as f:
            json.dump({"total": float(red.item()), "patterns": int(counts.item()), "world": world,
                       "nni": nni_t.tolist(), "nni_valid": [bool(np.isfinite(x)) for x in nni]}, f)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
