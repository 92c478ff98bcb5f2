#!/usr/bin/env python
"""Summaries of ncu output for profiles/ (run here, no GPU needed).

  python tools/ncu_summary.py launches <launches.csv>             per-kernel count / total time / share
  python tools/ncu_summary.py kernel <file.ncu-rep> [index]       key metrics + executed-instruction mix of one captured launch
"""
import collections
import csv
import io
import re
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "launch__occupancy_limit_registers",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum"]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        k = re.sub(r"\(.*", "", row["Kernel Name"])
        v = float(row["Metric Value"].replace(",", ""))
        if row["Metric Unit"] in ("ns", "nsecond"):
            v /= 1e3
        elif row["Metric Unit"] in ("ms", "msecond"):
            v *= 1e3
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:64s} n={v[0]:5d} us={v[1]:12.1f} share={v[1] / tot:.3f}")
    print(f"{'TOTAL':64s} n={sum(v[0] for v in agg.values()):5d} us={tot:12.1f}")


def kernel(path, index=0):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[2 + index]
    name = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
    print("kernel:", name[:120])
    d = dict(zip(hdr, zip(vals, units)))
    for k in KEYS:
        if k in d:
            print(f"  {k:88s} {d[k][0]:>18s} {d[k][1]}")
    src = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    # the source page concatenates the captured launches; take the index-th block
    blocks = src.split('"Kernel Name"')
    n_launches = max(1, len(rows) - 2)
    views = max(1, (len(blocks) - 1) // n_launches)  # ncu prints every launch once per source view (e.g. SASS, and source + SASS)
    if len(blocks) <= index * views + 1:
        return
    rows = list(csv.reader(io.StringIO('"Kernel Name"' + blocks[index * views + 1])))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    agg = collections.defaultdict(lambda: [0, 0])
    tot = tt = 0
    for r in rows[2:]:
        if len(r) < len(hdr) - 2:
            continue
        op = re.sub(r"^@!?U?P\d+\s+", "", r[ix["Source"]].strip()).split()[0].split(".")[0]
        ie, te = int(r[ix["Instructions Executed"]]), int(r[ix["Thread Instructions Executed"]])
        agg[op][0] += ie
        agg[op][1] += te
        tot += ie
        tt += te
    print(f"  executed warp instructions {tot}, threads per instruction {tt / max(tot, 1):.2f}; mix:")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:18]:
        print(f"    {k:10s} {v[0]:12d} {v[0] / tot:6.3f}  threads/inst {v[1] / max(v[0], 1):5.1f}")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        kernel(sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 0)
