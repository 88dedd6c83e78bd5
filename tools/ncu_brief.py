#!/usr/bin/env python
"""Print the handful of ncu raw metrics the design notes cite, from an .ncu-rep (read with `ncu -i`)."""
import csv, re, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
pat = re.compile(r"^(gpu__time_duration.sum|dram__bytes_(read|write).sum|sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed|"
                 r"sm__inst_executed_pipe_(xu|alu|fma|fmaheavy|fmalite|lsu|uniform|tensor.*)\.avg\.pct_of_peak_sustained_active|sm__pipe_(fma|fmaheavy|fmalite|alu|shared)_cycles_active.avg.pct_of_peak_sustained_active|"
                 r"sm__issue_active.avg.pct_of_peak_sustained_elapsed|smsp__issue_active.avg.per_cycle_active|launch__registers_per_thread|"
                 r"smsp__average_warps_issue_stalled_.*_per_issue_active.ratio|smsp__inst_executed.sum|sm__cycles_elapsed.avg|"
                 r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum|lts__t_bytes.sum|lts__throughput.avg.pct_of_peak_sustained_elapsed|"
                 r"l1tex__throughput.avg.pct_of_peak_sustained_elapsed|smsp__warps_eligible.avg.per_cycle_active|sm__warps_active.avg.per_cycle_active)$")
for v in rows[2:]:
    for a, b, c in zip(h, u, v):
        if pat.search(a):
            print(f"{a:95s} {b:10s} {c}")
