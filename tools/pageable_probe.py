#!/usr/bin/env python
"""What a caller with plain (pageable) host memory pays: phb_compress_rows on per-row buffers (the Rust shim's `load_dev`:
one Vec<u8> per sequence) and phb_log_likelihood_host on a pageable [pattern][row] block, with the pinned bounce path
(default) and without it (PHB_FORCE_BOUNCE=0: the driver stages the copies itself).

  python tools/pageable_probe.py [--taxa 10000] [--sites 1000000]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from phylo_rs_b200 import capi, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--taxa", type=int, default=10_000)
    ap.add_argument("--sites", type=int, default=1_000_000)
    ap.add_argument("--asr", action="store_true", help="also time marginal ASR (states: taxa x sites bytes of pageable memory)")
    args = ap.parse_args()
    n, W = args.taxa, args.sites
    tree = synthetic.yule_flat(n, 0x5EED0003)
    with capi.Context() as ctx:
        synthetic.load_into(ctx, tree, synthetic.gtr_model(alpha=0.5, k=4, p_inv=0.2))
        ctx.simulate_alignment(W, seed=0xA11C)
        block = ctx.get_alignment(0, W)  # [pattern][row], pageable numpy memory
        per_row = [np.ascontiguousarray(block[:, r]) for r in range(n)]  # one buffer per sequence, as Alignment::seqs holds them
        out = {"taxa": n, "sites": W, "bounce": os.environ.get("PHB_FORCE_BOUNCE", "auto"), "copy_threads": os.environ.get("PHB_COPY_THREADS", "8")}
        for name, fn in (("compress_rows_ms", lambda: ctx.compress_rows(per_row)),
                         ("log_likelihood_host_ms", lambda: ctx.log_likelihood_host(block)),
                         # marginal ASR with its states written to a pageable numpy array (capi allocates one per call)
                         ("marginal_asr_pageable_states_ms", (lambda: ctx.marginal_asr()) if args.asr else None)):
            if fn is None:
                continue
            fn()
            best = 1e30
            for _ in range(2):
                t0 = time.perf_counter()
                fn()
                best = min(best, (time.perf_counter() - t0) * 1e3)
            out[name] = best
        print(json.dumps(out))


if __name__ == "__main__":
    main()
