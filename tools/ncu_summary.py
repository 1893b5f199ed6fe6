#!/usr/bin/env python
"""Summarise an ncu report (raw page + source page) the way DESIGN.md / profiles/ quote it.
usage: tools/ncu_summary.py <report.ncu-rep> <points-in-launch> [out.json]"""
import collections, csv, io, json, re, subprocess, sys

rep, npts = sys.argv[1], float(sys.argv[2])
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, vals = rows[0], rows[2]
d = dict(zip(hdr, vals))
units = dict(zip(hdr, rows[1]))
f = lambda k: float(d[k])
# ncu picks the unit per report (us for short launches, ms for long ones): normalise to ms
t_scale = {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}.get(units.get("gpu__time_duration.sum", "ms"), 1.0)
out = {"kernel": d.get("Kernel Name"), "points": npts, "time_ms": f("gpu__time_duration.sum") * t_scale,
       "regs": f("launch__registers_per_thread"), "grid": f("launch__grid_size"), "block": f("launch__block_size"),
       "warps_active_pct": f("sm__warps_active.avg.pct_of_peak_sustained_active"),
       "fp64_pipe_active_pct": f("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
       "issue_active_pct": f("smsp__issue_active.avg.pct_of_peak_sustained_active"),
       "dram_read_MB": f("dram__bytes_read.sum"), "dram_write_MB": f("dram__bytes_write.sum"),
       "inst_per_point": f("smsp__inst_executed.sum") * 32 / npts,
       "fp64_inst_per_point": f("sm__inst_executed_pipe_fp64.sum") * 32 / npts,
       "dfma_per_point": f("smsp__sass_thread_inst_executed_op_dfma_pred_on.sum") / npts,
       "dadd_per_point": f("smsp__sass_thread_inst_executed_op_dadd_pred_on.sum") / npts,
       "dmul_per_point": f("smsp__sass_thread_inst_executed_op_dmul_pred_on.sum") / npts}
out["flops_per_point"] = 2 * out["dfma_per_point"] + out["dadd_per_point"] + out["dmul_per_point"]
out["Gpts_per_s_in_this_launch"] = npts / out["time_ms"] / 1e6
out["stalls_per_issue"] = {h.split("stalled_")[1].split("_per")[0]: round(float(v), 3) for h, v in zip(hdr, vals)
                           if "issue_stalled" in h and "per_issue_active" in h and float(v) > 0.05}
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h2 = rows[1]
iS, iE = h2.index("Source"), h2.index("Instructions Executed")
ops = collections.Counter()
for r in rows[2:]:
    if len(r) <= iE:
        continue
    s = re.sub(r"^@!?U?P\d+\s+", "", r[iS].strip())
    if s:
        ops[s.split()[0].split(".")[0]] += int(r[iE])
out["opcode_mix_per_point"] = {k: round(v * 32 / npts, 1) for k, v in ops.most_common(16)}
js = json.dumps(out, indent=1)
print(js)
if len(sys.argv) > 3:
    open(sys.argv[3], "w").write(js + "\n")
