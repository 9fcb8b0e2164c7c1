#!/usr/bin/env python
"""Pinned host -> device copy rate of this box, alone and with N ranks copying at once (the ceiling of bench.py's
e2e leg, which ships 10 GB / N per rank and step).

  python tools/pcie_probe.py                                  # one GPU
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/pcie_probe.py [--bind]

--bind pins each rank's thread (and therefore the first-touch placement of its pinned buffer) to the CPUs that are
local to its GPU (/sys/bus/pci/devices/<bdf>/local_cpulist), as bench.py does.  One JSON line per rank-0 report.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def local_cpus(device_index):
    """CPUs local to the GPU's PCIe root (None when sysfs does not say)."""
    bdf = None
    try:
        pr = torch.cuda.get_device_properties(device_index)
        if isinstance(getattr(pr, "pci_bus_id", None), int):
            bdf = f"{getattr(pr, 'pci_domain_id', 0):04x}:{pr.pci_bus_id:02x}:{getattr(pr, 'pci_device_id', 0):02x}.0"
    except Exception:
        bdf = None
    if bdf is None:
        try:
            import pynvml

            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            bdf = pynvml.nvmlDeviceGetPciInfo(h).busId
            if isinstance(bdf, bytes):
                bdf = bdf.decode()
        except Exception:
            return None, None
    bdf = bdf.lower()
    if len(bdf.split(":")[0]) == 8:  # 00000000:1b:00.0 -> 0000:1b:00.0
        bdf = bdf[4:]
    base = f"/sys/bus/pci/devices/{bdf}"
    try:
        cpulist = open(os.path.join(base, "local_cpulist")).read().strip()
        node = int(open(os.path.join(base, "numa_node")).read().strip())
    except OSError:
        return None, None
    cpus = set()
    for part in cpulist.split(","):
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        elif part:
            cpus.add(int(part))
    return sorted(cpus), node


def bind_to_gpu(device_index, rank_in_node=0, ranks_on_node=1):
    """Restrict this process to its share of the CPUs local to its GPU.  Returns a description for the JSON line."""
    cpus, node = local_cpus(device_index)
    if not cpus:
        return {"bound": False, "why": "no local_cpulist in sysfs"}
    allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
    if not allowed:
        return {"bound": False, "why": "local CPUs are outside this process's cpuset"}
    # ranks whose GPUs share the same local CPUs take disjoint slices, so their copy threads do not stack up
    per = max(1, len(allowed) // max(ranks_on_node, 1))
    mine = allowed[rank_in_node * per:(rank_in_node + 1) * per] or allowed
    os.sched_setaffinity(0, mine)
    return {"bound": True, "numa_node": node, "cpus": f"{mine[0]}-{mine[-1]}", "n_cpus": len(mine)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bind", action="store_true")
    ap.add_argument("--gib", type=float, default=0.0, help="GiB per rank (default: 10 GB / world)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    binding = bind_to_gpu(local, local, world) if args.bind else {"bound": False}
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = int(args.gib * (1 << 30)) if args.gib > 0 else int(1e10 // world)
    h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    h.fill_(65)  # first touch by this (possibly bound) thread
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    a.record()
    for _ in range(reps):
        d.copy_(h, non_blocking=True)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({"probe": "pinned H2D, all ranks at once", "ranks": world, "bytes_per_rank": n, "ms_max_over_ranks": float(t.item()),
                          "GBps_per_rank": n / float(t.item()) / 1e6, "GBps_aggregate": world * n / float(t.item()) / 1e6,
                          "binding_rank0": binding}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
