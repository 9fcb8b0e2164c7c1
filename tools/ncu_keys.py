#!/usr/bin/env python
"""Prints the metrics of an `ncu --page raw --csv` export (or of a .ncu-rep) whose names contain any of the given keys.
  python tools/ncu_keys.py profiles/r2_sweep_ncu_raw.csv gpu__time_duration sm__inst_executed_pipe_fp64"""
import csv
import subprocess
import sys

path, keys = sys.argv[1], sys.argv[2:]
if path.endswith(".ncu-rep"):
    text = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(text.splitlines()))
else:
    rows = list(csv.reader(open(path)))
hdr, units = rows[0], rows[1]
for data in rows[2:]:
    for i, h in enumerate(hdr):
        if any(k in h for k in keys):
            print(f"{h:90s} {data[i]:>18s} {units[i]}")
