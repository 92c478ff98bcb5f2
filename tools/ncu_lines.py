#!/usr/bin/env python
"""Where one captured launch spends its issue slots, by SOURCE line: joins the per-instruction counters of an ncu report
(--page source, SASS view) with the line table of the built library (nvdisasm -g).  Needs the report to come from the same build.

  python tools/ncu_lines.py <file.ncu-rep> <launch index> <substring of the mangled kernel name> [top N]
"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, idx, pat = sys.argv[1], int(sys.argv[2]), sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--launch-skip", str(idx), "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
h = next(i for i, r in enumerate(rows) if "# Samples" in r)
hdr = rows[h]
data = [r for r in rows[h + 1:] if len(r) == len(hdr) and r[0].startswith("0x")]
n_half = len(data) // 2 if len(data) > 1 and data[0][1] == data[len(data) // 2][1] and len(data) % 2 == 0 else len(data)
data = data[:n_half]  # some ncu versions list the function twice
ix = {n: i for i, n in enumerate(hdr)}
so = os.path.join(ROOT, "cpu-ray-tracing-implementation_b200", "libb200rt.so")
with tempfile.TemporaryDirectory() as td:
    subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=td, capture_output=True)
    cubin = [f for f in os.listdir(td) if f.endswith(".cubin")][0]
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], cwd=td, capture_output=True, text=True).stdout.split("\n")
starts = [i for i, l in enumerate(dis) if l.startswith("//--------------------- .text.")]
k = next(k for k, i in enumerate(starts) if pat in dis[i])
seg = dis[starts[k]:starts[k + 1] if k + 1 < len(starts) else len(dis)]
lines, cur = [], None
for l in seg:
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
    elif re.match(r"\s+/\*[0-9a-f]{4,6}\*/", l):
        lines.append(cur)
if len(lines) != len(data):
    sys.exit(f"instruction counts differ: report {len(data)}, library {len(lines)} (not the same build?)")
ex, th, smp = collections.Counter(), collections.Counter(), collections.Counter()
for r, ln in zip(data, lines):
    ex[ln] += int(r[ix["Instructions Executed"]])
    th[ln] += int(r[ix["Thread Instructions Executed"]])
    smp[ln] += int(r[ix["# Samples"]])
tot_ex, tot_s = sum(ex.values()), sum(smp.values())
per_file = collections.Counter()
for (f, _), v in ((k, v) for k, v in ex.items() if k):
    per_file[f] += v
print(f"warp instructions executed {tot_ex}, samples {tot_s}")
print("per file:", ", ".join(f"{f} {100 * v / tot_ex:.1f}%" for f, v in per_file.most_common(8)))
print(f"{'line':28s} {'% instr':>8s} {'% samples':>10s} {'lanes':>6s}")
for ln, v in ex.most_common(top):
    name = f"{ln[0]}:{ln[1]}" if ln else "?"
    print(f"{name:28s} {100 * v / tot_ex:8.2f} {100 * smp[ln] / max(1, tot_s):10.2f} {th[ln] / max(1, v):6.1f}")
