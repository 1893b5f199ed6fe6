#!/usr/bin/env python
"""Build-container tool: generates tests/golden/aga8_gerg2008.json.

Runs the NIST AGA8 GERG-2008 code that the reference keeps as an INDEPENDENT analytic implementation for its own tests
(/root/reference/src/tests/GERG2008.cpp, AlpharGERG :481-612) on the reference's validation set
(/root/reference/src/tests/catch_test_GERG.cxx: compositions :119-320, state points :322-510, component order :512) and stores
ar[i][j] = tau^i delta^j d^(i+j) alphar / dtau^i ddelta^j for the entries that code derives by hand (ar00, ar01, ar02, ar10, ar11,
ar20; ar03 is skipped because AGA8 omits the exponential-departure part there, GERG2008.cpp:604).  These pin Lambda_nm independently
of any automatic differentiation.  The tabulated pressures of the validation table are stored too (published, 1e-5 level).
"""
import json, os, re, subprocess, sys, tempfile
REF = "/root/reference"
src = open(REF + "/src/tests/catch_test_GERG.cxx").read()
m = re.search(r"mixture_comps\s*=\s*\{(.*?)\n\};", src, re.S)
comps = [[float(v) for v in row.split(",")] for row in re.findall(r"\{([^{}]*)\}", m.group(1))]
m = re.search(r"validation_data\s*=\s*\{(.*?)\n\};", src, re.S)
vals = [[float(v) for v in row.split(",")] for row in re.findall(r"\{([^{}]*)\}", m.group(1))]
m = re.search(r"components\s*=\s*\{([^}]*)\}", src)
names = re.findall(r'"([^"]+)"', m.group(1))
assert len(names) == 21 and all(len(c) == 21 for c in comps), (len(names), len(comps))
drv = r'''
#include <vector>
#include <cstdio>
#include "GERG2008.h"
void AlpharGERG(const int iprop, const double T, const double D, const std::vector<double> &x, double ar[4][4]);
int main(){ SetupGERG(); int n; if (scanf("%d",&n)!=1) return 1;
 for (int k=0;k<n;++k){ double T,D; std::vector<double> x(22,0.0); if (scanf("%lf %lf",&T,&D)!=2) return 1; for(int i=1;i<=21;++i) if (scanf("%lf",&x[i])!=1) return 1;
  double ar[4][4]; AlpharGERG(1,T,D,x,ar);
  printf("%.17g %.17g %.17g %.17g %.17g %.17g\n", ar[0][0],ar[0][1],ar[0][2],ar[1][0],ar[1][1],ar[2][0]); }
 return 0; }
'''
with tempfile.TemporaryDirectory() as td:
    open(td + "/drv.cpp", "w").write(drv)
    subprocess.check_call(["g++", "-O2", "-w", "-I" + REF + "/src/tests", td + "/drv.cpp", REF + "/src/tests/GERG2008.cpp", "-o", td + "/drv"])
    inp = [str(len(vals))]
    pts = []
    for v in vals:
        gas = int(v[0]); T = v[1]; D = v[2]
        x = [c / 100.0 for c in comps[gas - 2]]
        pts.append(dict(GasNo=gas, T=T, rho=D * 1e3, x=x, P_MPa=v[3]))
        inp.append("%r %r %s" % (T, D, " ".join(repr(c) for c in x)))
    out = subprocess.run([td + "/drv"], input="\n".join(inp), text=True, capture_output=True, check=True).stdout.strip().split("\n")
for p, line in zip(pts, out):
    a = [float(s) for s in line.split()]
    p.update(ar00=a[0], ar01=a[1], ar02=a[2], ar10=a[3], ar11=a[4], ar20=a[5])
json.dump(dict(source="NIST AGA8 GERG-2008 code as vendored in the reference's tests (src/tests/GERG2008.cpp), run by tools/make_golden_aga8.py",
               components=names, R=8.314472, points=pts), open(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "aga8_gerg2008.json"), "w"))
print("wrote", len(pts), "points")
