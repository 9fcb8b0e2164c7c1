"""N > 1 data path on CPU: world_size-2 and -3 gloo jobs (torch.distributed.run, 127.0.0.1) where every rank
evaluates its pattern shard through the C ABI (emulation build) and only scalars cross ranks — what bench.py does
with NCCL on GPUs.  The all-reduced totals must equal the single-context evaluation and the oracle."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle
from oracle import workloads
from phylo_rs_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_ranks_allreduce_equals_single_rank(emu_lib, tmp_path, world):
    out = tmp_path / "result.json"
    port = 29600 + world + (os.getpid() % 200)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "_gloo_worker.py"), str(out)]
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    got = json.load(open(out))
    case = workloads.make_case(11, 1501, model="gtr_i_g4", seed=61, ambiguity=0.05, compress=True)
    assert got["world"] == world and got["patterns"] == len(case.patterns)
    ref, _ = oracle.log_likelihood_flat(case.ft, case.model, case.patterns, case.mult)
    assert abs(got["total"] - ref) <= 1e-9 * abs(ref)
    with capi.Context(library=emu_lib) as ctx:
        workloads.load_into(ctx, case)
        single = ctx.log_likelihood()
        moves = [(v, lc) for v in range(case.ft["n_nodes"]) for lc in (0, 1)]
        nni = ctx.score_nni([m[0] for m in moves], [m[1] for m in moves])
    assert abs(got["total"] - single) <= 1e-12 * abs(single)
    ok = np.isfinite(nni)
    assert got["nni_valid"] == ok.tolist()
    assert np.max(np.abs(np.asarray(got["nni"])[ok] - nni[ok]) / np.abs(nni[ok])) <= 1e-12
