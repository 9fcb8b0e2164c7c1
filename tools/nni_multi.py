#!/usr/bin/env python
"""Config 5 on N GPUs: 1,024 NNI neighbours of a 2,000-taxon tree x 200k site patterns (GTR+G4), the patterns sharded
contiguously over the ranks (one process per GPU), every rank scoring all neighbours on its shard, the per-neighbour
totals all-reduced over NCCL (count = n_moves) -- the only exchange the path has.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29512 tools/nni_multi.py
  python tools/nni_multi.py            (N = 1)

Device-timed per rank (the context's event timer around phb_score_nni + CUDA events around the all-reduce), the maximum
over ranks is reported; rank 0 prints one JSON line."""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from phylo_rs_b200 import capi, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--taxa", type=int, default=2000)
    ap.add_argument("--patterns", type=int, default=200_000)
    ap.add_argument("--moves", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        sys.stdout.flush()
        real_stdout = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)  # NCCL's banner goes to stderr
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    else:
        real_stdout = sys.stdout
    seed = 0x5EED0005
    tree = synthetic.yule_flat(args.taxa, seed)
    co = tree["child_off"]
    internal = [v for v in range(tree["n_nodes"]) if co[v + 1] > co[v] and v != tree["root"]]
    moves = [(v, left) for v in internal for left in (0, 1)][:args.moves]
    ids, left = [m[0] for m in moves], [m[1] for m in moves]
    base_cnt, extra = divmod(args.patterns, world)
    cnt = base_cnt + (1 if rank < extra else 0)
    lo = rank * base_cnt + min(rank, extra)
    red = torch.zeros(len(moves), dtype=torch.float64, device="cuda")
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    with capi.Context(devices=[local_rank], flags=capi.FLAG_TIMING) as ctx:
        synthetic.load_into(ctx, tree, synthetic.gtr_model(alpha=0.5, k=4))
        ctx.simulate_alignment(cnt, global_offset=lo, seed=seed ^ 0xA11C)
        base = ctx.log_likelihood()

        def step():
            ctx.timer_start()
            part = ctx.score_nni(ids, left)
            ms = ctx.timer_stop()
            red.copy_(torch.from_numpy(np.ascontiguousarray(part)))
            ev[0].record()
            if world > 1:
                dist.all_reduce(red)
            ev[1].record()
            torch.cuda.synchronize()
            return ms + ev[0].elapsed_time(ev[1]), red.cpu().numpy()

        for _ in range(2):
            step()
        times = []
        for _ in range(args.reps):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            ms, totals = step()
            times.append(ms)
        st = ctx.stats()
    t = torch.tensor([min(times), st["nni_cache_bytes"] / 1e9, base], dtype=torch.float64, device="cuda")
    tmax, tsum = t.clone(), t.clone()
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    if rank == 0:
        ms = float(tmax[0])
        print(json.dumps({
            "case": f"cfg5 {len(moves)} NNI neighbours {args.taxa} x {args.patterns}, {world} GPU(s)", "n_gpus": world,
            "device_ms_max_over_ranks": ms, "trees_per_s": len(moves) / (ms * 1e-3), "cache_GB_per_rank": float(tmax[1]),
            "base_lnl": float(tsum[2]), "best_delta": float(np.nanmax(totals) - float(tsum[2])),
            "totals_head": [float(x) for x in totals[:3]], "exchange": "all-reduce of the per-neighbour totals (NCCL, f64 x n_moves)"}),
            file=real_stdout, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
