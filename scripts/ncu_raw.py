#!/usr/bin/env python
"""Key raw metrics of every kernel in an .ncu-rep.  Usage: python scripts/ncu_raw.py <rep> [substr ...]"""
import csv, subprocess, sys
rep = sys.argv[1]
extra = sys.argv[2:]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_warps", "launch__occupancy_limit_blocks"]
want += [h for h in hdr if "pcsamp_warps_issue_stalled" in h and "not_issued" not in h]
want += [h for h in hdr if any(e in h for e in extra)]
for r in rows[2:]:
    print("----", r[hdr.index("Kernel Name")][:60])
    for w in want:
        if w in hdr:
            i = hdr.index(w)
            print("  %-75s %-12s %s" % (w, units[i], r[i]))
