#!/usr/bin/env python
"""shared-memory pipe metrics of one ncu report: usage tools/ncu_shared.py <report.ncu-rep>"""
import csv, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); hdr, vals = rows[0], rows[2]
d = dict(zip(hdr, vals))
keys = ["gpu__time_duration.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
        "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "launch__block_size",
        "launch__registers_per_thread"]
for k in keys:
    print("%-75s %s" % (k, d.get(k)))
for h, v in zip(hdr, vals):
    if "issue_stalled" in h and "per_issue_active" in h:
        try:
            if float(v) > 0.15: print("   stall %-40s %.2f" % (h.split("stalled_")[1].split("_per")[0], float(v)))
        except ValueError:
            pass
