#!/usr/bin/env python
"""Hot SASS instructions of one captured launch (run here, no GPU): samples, dominant stall reasons, lanes per instruction.

  python tools/ncu_hot.py <file.ncu-rep> <launch index> [min share in %]
"""
import csv
import subprocess
import sys

rep, idx = sys.argv[1], int(sys.argv[2])
cut = float(sys.argv[3]) if len(sys.argv) > 3 else 0.7
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--launch-skip", str(idx), "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
h = next(i for i, r in enumerate(rows) if "# Samples" in r)
hdr = rows[h]
data = [r for r in rows[h + 1:] if len(r) == len(hdr) and r[0].startswith("0x")]
ix = {n: i for i, n in enumerate(hdr)}
S = ix["# Samples"]
tot = sum(int(r[S]) for r in data)
stalls = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
agg = {n: sum(int(r[ix[n]]) for r in data) for n in stalls}
print(rows[0][1][:120] if rows and len(rows[0]) > 1 else "")
print("samples", tot, "instructions", len(data))
print("stalls:", ", ".join(f"{n[6:]} {100 * v / tot:.1f}%" for n, v in sorted(agg.items(), key=lambda x: -x[1])[:9]))
for k, r in enumerate(data):
    s = int(r[S])
    if s > tot * cut / 100:
        top = sorted(((int(r[ix[n]]), n[6:]) for n in stalls), reverse=True)[:2]
        print(f"{k:5d} {r[1].strip()[:64]:64s} {100 * s / tot:5.1f}%  lanes {r[ix['Avg. Threads Executed']]:>5s}  exec {r[ix['Instructions Executed']]:>9s}  {top[0][1]} {top[0][0]}, {top[1][1]} {top[1][0]}")
