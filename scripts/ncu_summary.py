"""Summarise an .ncu-rep for profiles/: judged raw metrics per launch plus the
hottest source lines (needs -lineinfo).  Run where ncu is installed:

    python scripts/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_x.txt
"""
import collections
import csv
import re
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40


def ncu(*args):
    return subprocess.run(["ncu", "-i", rep, *args], capture_output=True, text=True).stdout


rows = list(csv.reader(ncu("--page", "raw", "--csv").splitlines()))
h, units = rows[0], rows[1]
pat = re.compile(
    r"^(gpu__time_duration.sum|dram__bytes_(read|write)\.sum|"
    r"dram__throughput.avg.pct_of_peak_sustained_elapsed|sm__throughput.avg.pct_of_peak_sustained_elapsed|"
    r"smsp__issue_active.avg.pct_of_peak_sustained_active|smsp__inst_executed.sum$|"
    r"sm__warps_active.avg.pct_of_peak_sustained_active|launch__registers_per_thread$|"
    r"lts__throughput.avg.pct_of_peak_sustained_elapsed|l1tex__throughput.avg.pct_of_peak_sustained_elapsed|"
    r"lts__t_sectors.sum$|lts__t_bytes.sum$|"
    r"sm__pipe_(alu|fma|fp64|xu)_cycles_active.avg.pct_of_peak_sustained_active|"
    r"sm__inst_executed_pipe_(xu|lsu|tensor.*)\.avg.pct_of_peak_sustained_active|"
    r"sm__pipe_tensor.*cycles_active.avg.pct_of_peak_sustained_active|"
    r"smsp__average_warps_issue_stalled_.*_per_issue_active.ratio|"
    r"smsp__thread_inst_executed_per_inst_executed.ratio|launch__occupancy_limit.*)")
for r in rows[2:]:
    print("== kernel:", r[h.index("Kernel Name")][:70], "grid", r[h.index("Grid Size")],
          "block", r[h.index("Block Size")])
    for a, u, c in zip(h, units, r):
        if pat.search(a) and c not in ("", "0"):
            print(f"  {a} [{u}] = {c}")

# hottest source lines (all profiled launches summed)
src = list(csv.reader(ncu("--page", "source", "--csv", "--print-source", "cuda,sass").splitlines()))
cur, hdr, tot = None, None, 0
agg = collections.defaultdict(lambda: [0, 0, ""])
for r in src:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or cur is None:
        continue
    try:
        ln = int(r[0])
        d = dict(zip(hdr, r))
        ie = int(d.get("Instructions Executed") or 0)
        sm = int(d.get("# Samples") or 0)
    except ValueError:
        continue
    a = agg[(cur, ln)]
    a[0] += ie
    a[1] += sm
    a[2] = r[1][:100]
    tot += ie
if tot:
    print(f"\n== hottest source lines (warp instructions executed, all profiled launches: {tot})")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"  {k[0]:18s}:{k[1]:<5d} inst={v[0]:>11d} ({100 * v[0] / tot:4.1f}%) "
              f"stall-samples={v[1]:>6d} | {v[2]}")
