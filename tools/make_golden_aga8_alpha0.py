#!/usr/bin/env python
"""Build-container tool: generates tests/golden/aga8_alpha0.json.

Runs Alpha0GERG of the NIST AGA8 GERG-2008 code the reference keeps for its own tests (/root/reference/src/tests/GERG2008.cpp:423-480) on the state
points already stored in tests/golden/aga8_gerg2008.json and stores a0[0..2] = alpha0, tau d(alpha0)/dtau, tau^2 d2(alpha0)/dtau2.  The reference
checks its ideal-gas model against this code through c_v (catch_test_GERG.cxx:597-620, 1e-9 relative), i.e. through a0[2] = Lambda_20, which does
not depend on the integration constants (AGA8 carries its own; the reference recomputes them, GERG.hpp:371-384) -- so a0[2] pins the second
temperature derivative of the GERG2008idealgas kind independently of any automatic differentiation; a0[0], a0[1] agree to ~1e-6 only
(catch_test_GERG.cxx:539-540) and are stored for that looser check."""
import json, os, subprocess, tempfile
REF = "/root/reference"
here = os.path.dirname(os.path.abspath(__file__))
g = json.load(open(os.path.join(here, "..", "tests", "golden", "aga8_gerg2008.json")))
drv = r'''
#include <vector>
#include <cstdio>
#include "GERG2008.h"
void Alpha0GERG(const double T, const double D, const std::vector<double> &x, double a0[3]);
int main(){ SetupGERG(); int n; if (scanf("%d",&n)!=1) return 1;
 for (int k=0;k<n;++k){ double T,D; std::vector<double> x(22,0.0); if (scanf("%lf %lf",&T,&D)!=2) return 1; for(int i=1;i<=21;++i) if (scanf("%lf",&x[i])!=1) return 1;
  double a0[3]; Alpha0GERG(T,D,x,a0); printf("%.17g %.17g %.17g\n", a0[0],a0[1],a0[2]); }
 return 0; }
'''
with tempfile.TemporaryDirectory() as td:
    open(td + "/drv.cpp", "w").write(drv)
    subprocess.check_call(["g++", "-O2", "-w", "-I" + REF + "/src/tests", td + "/drv.cpp", REF + "/src/tests/GERG2008.cpp", "-o", td + "/drv"])
    inp = [str(len(g["points"]))]
    for p in g["points"]:
        inp.append("%r %r %s" % (p["T"], p["rho"] / 1e3, " ".join(repr(c) for c in p["x"])))
    out = subprocess.run([td + "/drv"], input="\n".join(inp), text=True, capture_output=True, check=True).stdout.strip().split("\n")
pts = []
for p, line in zip(g["points"], out):
    a = [float(s) for s in line.split()]
    pts.append(dict(T=p["T"], rho=p["rho"], x=p["x"], a0_0=a[0], a0_1=a[1], a0_2=a[2]))
json.dump(dict(source="Alpha0GERG of the NIST AGA8 GERG-2008 code vendored in the reference's tests (src/tests/GERG2008.cpp:423-480), run by tools/make_golden_aga8_alpha0.py",
               components=g["components"], points=pts), open(os.path.join(here, "..", "tests", "golden", "aga8_alpha0.json"), "w"))
print("wrote", len(pts), "points")
