#!/usr/bin/env python
"""bench.py — headline benchmark: CLV site-node updates/s for GTR+I+G4 log-likelihood.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[2], the configuration the metric is quoted on): GTR+I+G4
log-likelihood of a 10,000-taxon Yule tree x 1,000,000 site patterns (synthetic, simulated under the
model on the device).  The pattern axis is sharded contiguously over the N ranks (strong scaling: the
total stays 1M patterns); every rank holds the full tree and model; only the scalar partial
log-likelihood crosses GPUs (one NCCL all-reduce per step).

A "step" is one full evaluation through the C ABI (P(t) kernel, all CLV updates, root kernel, scalar
back on the host).  `value` times K steps with the packed alignment resident in HBM; `e2e` times the
same evaluation fed from HOST buffers (alignment bytes in pinned host memory, [pattern][row] as
CompressedColumns.patterns) with the host->device copy, packing, tree/model upload and the result
read-back inside the timed region.

--impl reference: the reference's own algorithm (oracle port, reference-faithful mode: single thread,
P(t) rebuilt per child, linear leaf search — `parallel`/rayon does not touch this path) timed on a
bounded pattern sample of the same workload on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

N_TAXA = int(os.environ.get("PHB_BENCH_TAXA", 10_000))
N_SITES = int(os.environ.get("PHB_BENCH_SITES", 1_000_000))
TREE_SEED = 0x5EED0003
ALN_SEED = TREE_SEED ^ 0xA11C
METRIC = "CLV site-node updates/s (GTR+I+G4)"
UNIT = "updates/s"
WORKLOAD = f"GTR+I+G4 log-likelihood, {N_TAXA}-taxon Yule tree x {N_SITES} site patterns"


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([time.perf_counter()] + [x.strip() for x in line.split(",")])

    def stop(self, t_begin=None, t_end=None):
        """Summary over the samples that arrived inside [t_begin, t_end] (the timed region); when the region is
        shorter than the sampling period, over the samples since the sampler started (warm-up steps run the same
        kernels) — `window` says which."""
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        rows = [r[1:] for r in self.rows if t_begin is None or t_begin <= r[0] <= t_end + 0.15]
        window = "timed region"
        if not rows:
            rows, window = [r[1:] for r in self.rows], "warm-up + timed region (timed region shorter than one sample)"
        for r in rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "window": window}


def shard(P, world, rank):
    per = (P + world - 1) // world
    per += per & 1
    lo = min(rank * per, P)
    return lo, min(per, P - lo)


# --------------------------------------------------------------------------------------------------
# CPU legs (the only places that may execute oracle/)
# --------------------------------------------------------------------------------------------------
def cpu_sample(n_patterns, mode, threads, seed=0):
    """Times the oracle on a bounded sample of the workload: the first `n_patterns` patterns of an
    alignment simulated under the same tree and model.  Returns (updates/s, seconds)."""
    import oracle
    from oracle import workloads
    from phylo_rs_b200 import synthetic

    oracle.build()
    ft = synthetic.yule_flat(N_TAXA, TREE_SEED)
    ft["preorder"] = np.zeros(ft["n_nodes"], np.int32)
    model = workloads.make_model("gtr_i_g4")
    rng = np.random.default_rng(ALN_SEED + seed)
    # pre-order for the simulator
    pre, st = [], [ft["root"]]
    co, ch = ft["child_off"], ft["children"]
    while st:
        v = st.pop()
        pre.append(v)
        st.extend(reversed(ch[co[v]:co[v + 1]].tolist()))
    ft["preorder"] = np.asarray(pre, np.int32)
    pats = workloads.simulate_patterns(ft, model, n_patterns, rng)
    t0 = time.perf_counter()
    tot, _ = oracle.log_likelihood_flat(ft, model, pats, mode=mode, n_threads=threads)
    dt = time.perf_counter() - t0
    assert np.isfinite(tot)
    n_internal = int(np.sum(co[1:] != co[:-1]))
    return n_internal * n_patterns / dt, dt


def run_reference(args, rank):
    if rank != 0:
        return
    sample = int(os.environ.get("PHB_REF_SAMPLE", 32 if N_TAXA >= 5000 else 2048))
    for i in range(args.warmup):
        cpu_sample(max(sample // 8, 1), 1, 1, seed=i)
    times, vals = [], []
    for i in range(args.steps):
        v, dt = cpu_sample(sample, 1, 1, seed=100 + i)
        vals.append(v)
        times.append(dt)
    total_updates = sum(v * t for v, t in zip(vals, times))
    value = total_updates / sum(times)
    cores = os.cpu_count() or 1
    opt_v, opt_dt = cpu_sample(sample * 64, 2, cores, seed=999)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": f"{sample} patterns per step (patterns are independent; throughput per update is size-independent)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "port",
                         "sample": f"oracle reference-faithful mode (likelihood.rs:83-143 loop nest, P(t) per child, linear leaf search), "
                                   f"{sample} patterns x {args.steps} steps, 1 thread: the reference's `parallel` feature does not reach this path",
                         "cpu_opt": {"value": opt_v, "unit": UNIT, "cores": cores,
                                     "sample": f"oracle cpu_opt mode (P cached per edge x category, SoA, pthreads), {sample * 64} patterns, {opt_dt:.2f} s"}},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------------
def run_gpu(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    from phylo_rs_b200 import capi, synthetic

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    # each rank's thread -- and with it the first-touch placement of its pinned host buffer -- goes to the CPUs that
    # are local to its GPU (tools/pcie_probe.py); PHB_BENCH_BIND=0 switches it off
    binding = {"bound": False}
    if os.environ.get("PHB_BENCH_BIND", "1") == "1":
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import pcie_probe

            binding = pcie_probe.bind_to_gpu(local_rank, local_rank, world)
        except Exception as e:  # binding is an optimisation, never a reason to fail
            binding = {"bound": False, "why": f"{type(e).__name__}: {e}"}
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL prints its version banner on fd 1 at communicator creation: point fd 1 at stderr for the run
        # and keep the real stdout for the one JSON line
        sys.stdout.flush()
        _real_stdout = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    tree = synthetic.yule_flat(N_TAXA, TREE_SEED)
    model = synthetic.gtr_model(alpha=0.5, k=4, p_inv=0.2)
    n_internal = N_TAXA - 1
    lo, cnt = shard(N_SITES, world, rank)
    level_path = os.environ.get("PHB_BENCH_LEVELS") == "1"  # the HBM-resident level kernels, for comparison
    ctx = capi.Context(devices=[local_rank], flags=capi.FLAG_TIMING | (capi.FLAG_LEVEL_KERNELS if level_path else 0))
    fp64_peak = ctx.measure_fp64_tflops()
    synthetic.load_into(ctx, tree, model)
    ctx.simulate_alignment(cnt, global_offset=lo, seed=ALN_SEED)
    red = torch.zeros(1, dtype=torch.float64, device="cuda")

    def step():
        v = ctx.log_likelihood()
        if world > 1:
            red[0] = v
            dist.all_reduce(red)
            return float(red.item())
        return v

    sampler = ClockSampler(local_rank)
    sampler.start()
    t_w = time.perf_counter()
    for _ in range(max(args.warmup, 3)):
        lnl = step()
    while time.perf_counter() - t_w < 0.6 and not sampler.rows:  # let nvidia-smi deliver its first sample under load
        ctx.log_likelihood()  # local evaluation only: ranks may run different numbers of these
    barrier()
    launches0 = ctx.stats()["kernel_launches"]
    ctx.timer_start()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        lnl = step()
    dev_ms = ctx.timer_stop()
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop(t0, t0 + wall_ms * 1e-3)
    st = ctx.stats()
    launches = st["kernel_launches"] - launches0
    # ---- end to end from host buffers ------------------------------------------------------------
    if os.environ.get("PHB_BENCH_KERNEL_ONLY") == "1":  # profiling runs (ncu): the device-timed part only, no JSON contract
        if rank == 0:
            print(json.dumps({"kernel_only": True, "ms_per_step": dev_ms / args.steps, "sweep": {k: st[k] for k in st if k.startswith("sweep_")}}), flush=True)
        ctx.close()
        if world > 1:
            dist.destroy_process_group()
        return
    e2e_steps = max(1, min(args.steps, 3))
    n_rows = tree["n_taxa"]
    host = torch.empty((cnt, n_rows), dtype=torch.uint8, pin_memory=True)
    chunk = max(1, (256 << 20) // n_rows)
    for p0 in range(0, cnt, chunk):
        c = min(chunk, cnt - p0)
        host.numpy()[p0:p0 + c] = ctx.get_alignment(p0, c)

    def e2e_step():
        # the reference-facing call shape: model + tree + host alignment bytes in, lnL out
        synthetic.load_into(ctx, tree, model)
        v = ctx.log_likelihood_host((host.data_ptr(), n_rows, cnt))
        if world > 1:
            red[0] = v
            dist.all_reduce(red)
            return float(red.item())
        return v

    lnl_e2e = e2e_step()
    barrier()
    t1 = time.perf_counter()
    for _ in range(e2e_steps):
        lnl_e2e = e2e_step()
    barrier()
    e2e_ms = (time.perf_counter() - t1) * 1e3 / e2e_steps
    # what the end-to-end step is made of, measured apart (all ranks at once): the H2D copy of this rank's bytes alone,
    # and the host-side preparation alone (model + tree upload, schedule, plans: one call on a 64-pattern alignment)
    dev_buf = torch.empty(min(cnt * n_rows, 1 << 31), dtype=torch.uint8, device="cuda")
    flat = host.view(-1)
    barrier()
    ev_a, ev_b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev_a.record()
    for off in range(0, flat.numel(), dev_buf.numel()):
        seg = flat[off:off + dev_buf.numel()]
        dev_buf[:seg.numel()].copy_(seg, non_blocking=True)
    ev_b.record()
    torch.cuda.synchronize()
    copy_only_ms = ev_a.elapsed_time(ev_b)
    del dev_buf
    small = np.ascontiguousarray(host.numpy()[:64])
    with capi.Context(devices=[local_rank]) as c2:
        synthetic.load_into(c2, tree, model)
        c2.log_likelihood_host(small)
        t2 = time.perf_counter()
        for _ in range(3):
            synthetic.load_into(c2, tree, model)
            c2.log_likelihood_host(small)
        host_prep_ms = (time.perf_counter() - t2) * 1e3 / 3
    # the same call from PAGEABLE host memory (a caller's plain Vec<u8>): one step, reported beside the pinned figure
    e2e_pageable_ms = None
    if world == 1 and os.environ.get("PHB_BENCH_PAGEABLE", "1") == "1":
        pageable = np.empty((cnt, n_rows), np.uint8)
        np.copyto(pageable, host.numpy())
        synthetic.load_into(ctx, tree, model)
        ctx.log_likelihood_host((pageable.ctypes.data, n_rows, cnt))  # first touch
        torch.cuda.synchronize()
        tp = time.perf_counter()
        synthetic.load_into(ctx, tree, model)
        lnl_pg = ctx.log_likelihood_host((pageable.ctypes.data, n_rows, cnt))
        e2e_pageable_ms = (time.perf_counter() - tp) * 1e3
        assert abs(lnl_pg - lnl_e2e) <= 1e-12 * abs(lnl_e2e)
        del pageable
    # ---- reduce over ranks ------------------------------------------------------------------------
    tmax = torch.tensor([dev_ms, wall_ms, e2e_ms, copy_only_ms, host_prep_ms], dtype=torch.float64, device="cuda")
    tsum = torch.tensor([float(launches)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    dev_ms, wall_ms, e2e_ms, copy_only_ms, host_prep_ms = (float(x) for x in tmax.tolist())
    launches = int(tsum.item())
    if rank == 0:
        peak, peak_src = peaks()
        updates = n_internal * N_SITES
        ms_step = dev_ms / args.steps
        kern_ms = st["ms_clv"] / max(st["clv_launches"], 1)
        bytes_launch = st["algorithmic_bytes"] / max(st["clv_launches"], 1)
        achieved = bytes_launch / (kern_ms * 1e-3) / 1e9 if kern_ms > 0 else None
        traffic, traffic_note = None, None
        kernel = "clv_update_level_kernel" if level_path else "sweep_kernel"
        try:  # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu capture of THIS kernel source
            import hashlib

            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                tj = json.load(f).get(kernel, {})
            src = os.path.join(ROOT, "phylo-rs_b200", "csrc", tj.get("source_file", "sweep.cuh"))
            sha = hashlib.sha256(open(src, "rb").read()).hexdigest()
            if tj.get("patterns") and tj.get("source_sha256") == sha:
                per_eval = tj.get("dram_bytes_per_evaluation", tj["dram_bytes_per_launch"])  # all launches of one evaluation
                traffic = per_eval * (cnt / tj["patterns"]) * (N_TAXA / tj["taxa"]) / max(st["clv_launches"], 1)
                traffic_note = f"ncu capture {tj.get('capture')} of {tj.get('source_file')} at {sha[:12]}"
            else:
                traffic_note = "no ncu capture on file for the current kernel source (profiles/traffic.json is for another revision)"
        except Exception:
            pass
        # FP64 pipe: the instructions the schedule actually executes (tip operands are table lookups and cost none)
        fp64_instr = st.get("sweep_fp64_instr", 0.0) * st["n_clv_categories"] * cnt
        fp64_flops = st.get("sweep_flops", 0.0) * st["n_clv_categories"] * cnt
        t_clv = st["ms_clv"] * 1e-3
        fp64_tflops = fp64_flops / t_clv / 1e12 if t_clv > 0 else None
        fp64_pipe_frac = (fp64_instr / t_clv) / (fp64_peak * 1e12 / 2.0) if (t_clv > 0 and fp64_peak) else None
        line = {
            "metric": METRIC, "value": updates / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "model": "GTR+I+G4 (pi=.3,.2,.2,.3; rates 1,2.5,1,1,2.5,1; alpha=0.5, k=4; p_inv=0.2)",
                       "patterns_per_gpu": cnt, "sharding": f"site patterns over {world} rank(s); scalar all-reduce",
                       "l2": "inputs larger than L2 (tip codes + CLV tiles >> 126 MB)",
                       "path": "level kernels" if level_path else "fused sweep",
                       "tiles_per_step": st["n_tiles"], "tile_patterns": st["tile_patterns"], "levels": st["n_levels"],
                       "sweep": {k: st[k] for k in st if k.startswith("sweep_")}},
            "lnl": lnl, "lnl_e2e": lnl_e2e, "wall_ms_per_step": wall_ms / args.steps,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": (achieved / peak) if achieved else None, "traffic": traffic, "traffic_source": traffic_note,
                         "peak_source": peak_src,
                         "kernel": kernel, "launches_per_step": st["clv_launches"],
                         "avg_launch_ms": kern_ms, "algorithmic_bytes_per_launch": bytes_launch,
                         "whole_step_frac": st["algorithmic_bytes"] / (ms_step * 1e-3) / 1e9 / peak,
                         "note": ("level-batched kernels: every CLV makes one HBM round trip" if level_path else
                                  "fused sweep keeps CLVs in registers/shared memory, so the algorithmic HBM bytes of "
                                  "SURVEY 8d are NOT moved (traffic = tip codes + the P(t) stream): frac > 1 means the "
                                  "HBM roofline of the per-node formulation is beaten, the binding resources are the "
                                  "FP64 pipe and shared-memory bandwidth (see fp64)"),
                         "fp64": {"executed_flops_per_pattern": st.get("sweep_flops", 0.0) * st["n_clv_categories"],
                                  "executed_fp64_instructions_per_pattern": st.get("sweep_fp64_instr", 0.0) * st["n_clv_categories"],
                                  "achieved_tflops": fp64_tflops, "peak_tflops_measured": fp64_peak,
                                  "pipe_frac": fp64_pipe_frac,
                                  "note": "pipe_frac = executed DMUL+DFMA per second / (measured DFMA rate): the figure "
                                          "to hold against ncu's sm__inst_executed_pipe_fp64; flops count a DFMA as 2"}},
            "e2e": {"value": updates / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": int(cnt) * n_rows + 8 * 60 + 8 * 4 * tree["n_nodes"] + 4 * 4 * tree["n_nodes"],
                    "d2h_bytes_per_step": 8 + 4 * n_rows, "steps": e2e_steps,
                    "pageable_ms_per_step": e2e_pageable_ms,
                    "breakdown_ms": {"h2d_copy_alone": copy_only_ms, "device_compute_alone": ms_step, "host_prep_alone": host_prep_ms,
                                     "note": "max over ranks, each measured apart with all ranks active: the copy and the compute "
                                             "overlap inside the step, the host preparation runs while the first chunks are on the wire"},
                    "cpu_binding_rank0": binding,
                    "note": "per rank: phb_log_likelihood_host on pinned host [pattern][row] bytes: chunked H2D overlapped with pack + sweep, "
                            "tree/model upload, scalar back; pageable_ms_per_step: the same call on plain malloc'ed memory (N=1 only); "
                            "compress_columns is not in this figure (configs.compress times it)"},
            "gpu_launches": launches,
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu:
            sample = int(os.environ.get("PHB_REF_SAMPLE", 64 if N_TAXA >= 5000 else 4096))
            v, dt = cpu_sample(sample, 1, 1)
            cores = os.cpu_count() or 1
            ov, odt = cpu_sample(sample * 64, 2, cores)
            line["cpu_baseline"] = {
                "value": v, "unit": UNIT, "cores": 1, "kind": "port",
                "sample": f"oracle reference-faithful mode, first {sample} patterns of the same workload, {dt:.1f} s, 1 thread "
                          "(the reference's rayon feature does not reach this path)",
                "cpu_opt": {"value": ov, "unit": UNIT, "cores": cores, "sample": f"oracle cpu_opt mode, {sample * 64} patterns, {odt:.1f} s"}}
        if world == 1 and not args.no_configs:
            # BASELINE.json's other configurations on this GPU, each with parity deltas against the oracle (outside `value`)
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import configs_bench

            del host
            ctx.close()
            torch.cuda.empty_cache()
            line["configs"] = configs_bench.driver_configs(peak, reps=3)
        print(json.dumps(line), flush=True, file=_real_stdout if world > 1 else sys.stdout)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-configs", action="store_true", help="skip the `configs` object (BASELINE configs 2, 4, 5, joint ASR, compression)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29531"), os.path.abspath(__file__),
               "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup)] + (["--no-cpu"] if args.no_cpu else []) + (["--no-configs"] if args.no_configs else [])
        raise SystemExit(subprocess.call(cmd))
    run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
