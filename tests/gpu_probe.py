"""Not a test: a quick timing probe used during development (python tests/gpu_probe.py)."""
import json
import sys
import time

sys.path.insert(0, "/root/repo")
import numpy as np

from oracle import workloads
from phylo_rs_b200 import capi


def main():
    n_taxa = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    P = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
    model = sys.argv[3] if len(sys.argv) > 3 else "gtr_g4"
    t0 = time.time()
    case = workloads.make_case(n_taxa, 4, model=model, seed=1)
    print("case built", time.time() - t0, flush=True)
    flags = capi.FLAG_TIMING | (capi.FLAG_LEVEL_KERNELS if len(sys.argv) > 4 and sys.argv[4] == 'levels' else 0)
    with capi.Context(flags=flags) as ctx:
        print("fp64 TFLOP/s:", ctx.measure_fp64_tflops(), flush=True)
        workloads.load_into(ctx, case)
        ctx.simulate_alignment(P, seed=5)
        for i in range(4):
            t0 = time.time()
            tot = ctx.log_likelihood()
            dt = time.time() - t0
            st = ctx.stats()
            print(json.dumps(dict(lnl=tot, wall_ms=dt * 1e3, **st)), flush=True)
            print("  GB/s(clv kernels) =", st["algorithmic_bytes"] / (st["ms_clv"] * 1e-3) / 1e9,
                  " GB/s(total) =", st["algorithmic_bytes"] / (st["ms_total"] * 1e-3) / 1e9, flush=True)


if __name__ == "__main__":
    main()
