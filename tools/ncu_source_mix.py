#!/usr/bin/env python
"""Instruction mix and stall picture of one kernel from an `ncu --set full --import-source on` report: per SASS region
(runs of instructions executed equally often) the instructions per unit of work, the share of warp samples, the shared-
memory wavefronts, and the top stall reasons overall.

  python tools/ncu_source_mix.py REPORT.ncu-rep UNITS [LO HI]
UNITS = what to normalise by (e.g. warp-level op executions = consumer warps x ops x blocks); LO HI (hex offsets)
dump that address range instruction by instruction with its two top stall reasons."""
import csv
import subprocess
import sys

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
data = [r for r in rows[2:] if len(r) >= len(hdr)]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]


def f(x):
    try:
        return float(x)
    except ValueError:
        return 0.0


N = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
tots = sum(f(r[ix["# Samples"]]) for r in data)
tot = {s: sum(f(r[ix[s]]) for r in data) for s in stalls}
print("samples", tots, "inst/unit", sum(f(r[ix["Instructions Executed"]]) for r in data) / N, "smem wavefronts/unit",
      sum(f(r[ix["L1 Wavefronts Shared"]]) for r in data) / N)
print(" ".join(f"{s[6:]}={100 * v / tots:.1f}%" for s, v in sorted(tot.items(), key=lambda kv: -kv[1])[:9]))
base = int(data[0][ix["Address"]], 16)
cur, agg = None, []
for r in data:
    a = int(r[ix["Address"]], 16) - base
    rat = f(r[ix["Instructions Executed"]]) / N
    s = f(r[ix["# Samples"]])
    wf = f(r[ix["L1 Wavefronts Shared"]]) / N
    key = round(rat, 3)
    if cur and cur["key"] == key:
        cur["n"] += 1
        cur["s"] += s
        cur["wf"] += wf
        cur["end"] = a
    else:
        cur = {"key": key, "n": 1, "s": s, "wf": wf, "start": a, "end": a, "first": r[ix["Source"]].strip()}
        agg.append(cur)
for g in agg:
    if g["key"] * g["n"] >= 0.5 or g["s"] / tots > 0.004:
        print(f"{g['start']:6x}-{g['end']:6x} n={g['n']:4d} ratio={g['key']:.3f} inst/unit={g['key'] * g['n']:6.1f} "
              f"samples={100 * g['s'] / tots:5.2f}% wf/unit={g['wf']:.2f}  {g['first'][:50]}")
if len(sys.argv) > 4:
    lo, hi = int(sys.argv[3], 16), int(sys.argv[4], 16)
    for r in data:
        a = int(r[ix["Address"]], 16) - base
        if lo <= a <= hi:
            st = sorted([(f(r[ix[s]]), s[6:]) for s in stalls], reverse=True)[:2]
            print(f"{a:6x} {f(r[ix['Instructions Executed']]) / N:.3f} {100 * f(r[ix['# Samples']]) / tots:5.2f}% "
                  f"{r[ix['Source']].strip()[:64]:64s} {st[0][1]}:{st[0][0]:.0f} {st[1][1]}:{st[1][0]:.0f}")
