"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list (profiles/*_launches.csv)."""
import csv
import sys
from collections import defaultdict

path, title = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
rows = [r for r in csv.reader(l for l in open(path, errors="replace") if l.startswith('"'))]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows[1:]:
    if len(r) <= vi:
        continue
    v = float(r[vi].replace(",", ""))
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui].replace("second", "s").replace("usecond", "us"), 1e-6)
    name = r[ki].split("(")[0]
    tot[name] += v * scale
    cnt[name] += 1
allt = sum(tot.values())
if title:
    print(title)
for k in sorted(tot, key=tot.get, reverse=True):
    print(f"{k:60s} n={cnt[k]:5d} total_ms={tot[k]:10.3f} share={tot[k] / allt:.4f}")
