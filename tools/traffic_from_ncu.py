#!/usr/bin/env python
"""Writes profiles/traffic.json from an `ncu --set full` report of bench.py's top kernel: DRAM bytes per launch
(dram__bytes_read.sum + dram__bytes_write.sum) together with the SHA-256 of the kernel source it was taken from.
bench.py quotes `roofline.traffic` only while that hash still matches the source in the tree.

  python tools/traffic_from_ncu.py gpurun_out/<report>.ncu-rep [--kernel sweep_kernel] [--source sweep.cuh]
"""
import argparse
import csv
import hashlib
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report", nargs="+", help="one or more reports; launches whose DRAM counters ncu left as NaN are skipped")
    ap.add_argument("--kernel", default="sweep_kernel")
    ap.add_argument("--source", default="sweep.cuh")
    ap.add_argument("--taxa", type=int, default=10_000)
    ap.add_argument("--patterns", type=int, default=1_000_000)
    args = ap.parse_args()
    import math

    scale = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12,
             "ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1, "msecond": 1, "s": 1e3, "second": 1e3}
    launches = []  # (read bytes, written bytes, ms, grid) of every launch with valid DRAM counters
    for rep in args.report:
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
        rows = list(csv.reader(out.splitlines()))
        hdr, units = rows[0], rows[1]
        col = {h: i for i, h in enumerate(hdr)}

        def val(name, data):
            return float(data[col[name]].replace(",", "")) * scale.get(units[col[name]].lower(), 1)

        for r in rows[2:]:
            if len(r) <= col["dram__bytes_read.sum"]:
                continue
            try:
                rd_ = val("dram__bytes_read.sum", r)
            except ValueError:
                continue
            if math.isnan(rd_):
                continue  # ncu left the counters of this launch empty
            launches.append((rd_, val("dram__bytes_write.sum", r), val("gpu__time_duration.sum", r), r[col["Grid Size"]]))
    # one evaluation = every captured launch (the sweep goes out as the full passes + a narrower launch for the rest)
    rd = sum(l[0] for l in launches)
    wr = sum(l[1] for l in launches)
    dur = sum(l[2] for l in launches)
    src = os.path.join(ROOT, "phylo-rs_b200", "csrc", args.source)
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        tj = json.load(open(path))
    except Exception:
        tj = {}
    tj[args.kernel] = {
        "dram_bytes_per_evaluation": rd + wr, "launches_per_evaluation": len(launches),
        "dram_bytes_per_launch": (rd + wr) / max(len(launches), 1), "dram_read": rd, "dram_write": wr,
        "patterns": args.patterns, "taxa": args.taxa, "duration_ms_under_ncu": dur,
        "capture": " + ".join(os.path.basename(r) for r in args.report), "grids": [l[3] for l in launches], "source_file": args.source,
        "source_sha256": hashlib.sha256(open(src, "rb").read()).hexdigest(),
    }
    json.dump(tj, open(path, "w"), indent=1)
    print(json.dumps(tj[args.kernel], indent=1))


if __name__ == "__main__":
    main()
