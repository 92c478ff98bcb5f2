#!/usr/bin/env python
"""Instruction count of one kernel per source file and per source line, from `nvdisasm -g -c` of the built library
(no GPU needed).  Usage: python tools/codesize.py <substring of the mangled kernel name> [top N]"""
import collections
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "cpu-ray-tracing-implementation_b200", "libb200rt.so")
pat = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
with tempfile.TemporaryDirectory() as td:
    subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=td, capture_output=True)
    cubin = [f for f in os.listdir(td) if f.endswith(".cubin")][0]
    text = subprocess.run(["nvdisasm", "-g", "-c", cubin], cwd=td, capture_output=True, text=True).stdout.split("\n")
starts = [i for i, l in enumerate(text) if l.startswith("//--------------------- .text.")]
for k, i in enumerate(starts):
    if pat not in text[i]:
        continue
    end = starts[k + 1] if k + 1 < len(starts) else len(text)
    cur, per_line, per_file, total = None, collections.Counter(), collections.Counter(), 0
    for l in text[i:end]:
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
        elif re.match(r"\s+/\*[0-9a-f]{4,6}\*/", l):
            total += 1
            if cur:
                per_line[cur] += 1
                per_file[cur[0]] += 1
    print(text[i].strip("/- ").replace(".text.", "")[:150])
    print(f"  {total} instructions = {total * 16 / 1024:.0f} KB")
    print("  per file:", ", ".join(f"{f} {n}" for f, n in per_file.most_common(8)))
    print("  per line:", ", ".join(f"{f}:{ln} {n}" for (f, ln), n in per_line.most_common(top)))
