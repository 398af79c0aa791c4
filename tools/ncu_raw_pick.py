#!/usr/bin/env python
"""Print selected metrics of the first kernel in an .ncu-rep (ncu -i X --page raw --csv piped through this)."""
import csv
import subprocess
import sys

rep = sys.argv[1]
want = sys.argv[2:] or ["gpu__time_duration.sum", "smsp__issue_active.avg.pct", "sm__warps_active.avg.pct_of_peak",
                        "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__occupancy_limit",
                        "smsp__average_warp", "smsp__thread_inst_executed_per_inst_executed.ratio",
                        "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__data_bank_conflicts_pipe_lsu.sum",
                        "smsp__warps_eligible.avg.per_cycle_active", "sm__cycles_elapsed.max"]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
head, units = rows[0], rows[1]
for r in rows[2:]:
    for k, u, v in zip(head, units, r):
        if any(w in k for w in want):
            print("%-90s %-12s %s" % (k, u, v))
    print("----")
