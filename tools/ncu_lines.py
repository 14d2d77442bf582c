#!/usr/bin/env python3
"""Per-source-line share of executed instructions and stall samples from an .ncu-rep
(ncu --set full --import-source on).  usage: tools/ncu_lines.py report.ncu-rep [min_pct]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw)))
h, v = r[0], r[2]
for k in ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
          "sm__warps_active.avg.pct_of_peak_sustained_active", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
          "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
          "launch__registers_per_thread",
          "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
          "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
          "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
          "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
          "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
          "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
          "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
          "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"]:
    if k in h:
        print(f"{k} = {v[h.index(k)]} {r[1][h.index(k)]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
cur, out, ti, ts = None, [], 0, 0
for x in rows:
    if len(x) >= 2 and x[0] == "File Path":
        cur = x[1].split("/")[-1]
        continue
    if len(x) < 8 or x[0] in ("Line No", "Function Name"):
        continue
    if x[0] != "" and x[2] == "-":
        try:
            out.append((cur, int(x[0]), x[1].strip()[:100], int(x[6]), int(x[7])))
            ti += int(x[7])
            ts += int(x[6])
        except ValueError:
            pass
print("warp instructions", ti, "samples", ts)
for f, l, s, sm, ins in out:
    if ins > thr / 100 * ti or sm > thr / 100 * ts:
        print(f"{f[:20]:20s} {l:4d} inst {100 * ins / ti:5.1f}% samp {100 * sm / ts:5.1f}%  {s}")
