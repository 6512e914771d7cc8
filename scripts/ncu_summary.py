"""Print the judged metrics of an .ncu-rep (run where ncu is installed)."""
import csv, re, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, units = rows[0], rows[1]
pat = re.compile(r"^(gpu__time_duration.sum|dram__bytes_(read|write)\.sum|dram__throughput.avg.pct_of_peak_sustained_elapsed|"
                 r"sm__throughput.avg.pct|smsp__issue_active.avg.pct|smsp__inst_executed.sum$|sm__warps_active.avg.pct|"
                 r"launch__registers_per_thread$|launch__grid_size|launch__block_size|lts__throughput.avg.pct|"
                 r"lts__d_atomic_input_cycles_active.(avg|max).pct|lts__t_requests_srcunit_tex_op_red.sum|"
                 r"lts__t_sectors.sum$|sm__inst_executed_pipe_(xu|alu|fma|lsu|tensor.*|tc|tma)\.avg.pct_of_peak_sustained_active|"
                 r"sm__pipe_tensor_cycles_active.*|smsp__average_warps_issue_stalled_.*_per_issue_active.ratio|"
                 r"l1tex__m_l1tex2xbar_write_sectors_mem_global_op_red.sum$|launch__occupancy_limit.*|smsp__warps_eligible.avg.per_cycle_active)")
for r in rows[2:]:
    print("== kernel:", r[h.index("Kernel Name")][:60], "grid", r[h.index("Grid Size")], "block", r[h.index("Block Size")])
    for a, u, c in zip(h, units, r):
        if pat.search(a):
            print(f"  {a} [{u}] = {c}")
