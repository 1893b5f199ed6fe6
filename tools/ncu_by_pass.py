#!/usr/bin/env python
"""Aggregate the source page of an ncu report (captured with --import-source on, built with -lineinfo) by source line and by the passes of the
lane-cooperative GERG isochoric kernel.  usage: tools/ncu_by_pass.py <report.ncu-rep>"""
import collections, csv, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
cur = None; hdr = None
agg = collections.defaultdict(lambda: [0.0, 0.0, 0.0, 0.0, ''])
tot = toti = 0.0
last_line = {}
for r in rows:
    if not r: continue
    if r[0] in ('File Name', 'File Path'): cur = r[1].split('/')[-1]; hdr = None; continue
    if r[0] == 'Function Name': continue
    if r[0] == 'Line No': hdr = r; continue
    if hdr and len(r) >= len(hdr) - 2:
        d = dict(zip(hdr, r))
        def f(k):
            try: return float(d.get(k, 0) or 0)
            except ValueError: return 0.0
        line = r[0].strip() or last_line.get(cur, '0')
        last_line[cur] = line
        a = agg[(cur, line)]
        a[0] += f('# Samples'); a[1] += f('Instructions Executed'); a[2] += f('L1 Wavefronts Shared Excessive'); a[3] += f('L1 Wavefronts Shared')
        if r[1].strip(): a[4] = r[1].strip()
        tot += f('# Samples'); toti += f('Instructions Executed')
src = open('teqp_b200/csrc/tq_gerg_iso_coop.cuh').read().split('\n')
def lineno(marker):
    for i, l in enumerate(src):
        if marker in l: return i + 1
marks = [('pass A', lineno('// ---- pass A')), ('pass B0', lineno('// ---- pass B0')), ('pass B1', lineno('// ---- pass B1')), ('pass B2', lineno('// ---- pass B2')),
         ('gradient', lineno('const double S0r = S0 * invr')), ('pass C', lineno('// ---- pass C')), ('end', len(src) + 1)]
byp = collections.defaultdict(lambda: [0.0, 0.0, 0.0, 0.0])
for (fn, line), a in agg.items():
    name = 'other: ' + str(fn)
    if fn == 'tq_gerg_iso_coop.cuh':
        ln = int(line); name = 'prologue'
        for i in range(len(marks) - 1):
            if marks[i][1] <= ln < marks[i + 1][1]: name = marks[i][0]
    b = byp[name]
    for i in range(4): b[i] += a[i]
wf = sum(x[3] for x in byp.values())
for k, b in sorted(byp.items(), key=lambda kv: -kv[1][0]):
    print("%-30s samples %5.1f%%  instructions %5.1f%%  shared wavefronts %5.1f%% (excess %4.1f%% of its own)" % (k, 100 * b[0] / tot, 100 * b[1] / toti, 100 * b[3] / max(wf, 1), 100 * b[2] / max(b[3], 1)))
for (fn, line), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[2]) if len(sys.argv) > 2 else 20]:
    print("%s:%s smp %4.1f%% inst %4.1f%% wf %6.0fk exc %6.0fk | %s" % (str(fn)[:22], line, 100 * a[0] / tot, 100 * a[1] / toti, a[3] / 1e3, a[2] / 1e3, a[4][:110]))
