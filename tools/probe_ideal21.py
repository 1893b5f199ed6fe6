"""GPU probe: the 21-component GERG-2008 ideal-gas kind through the generic jet-sweep kernels with 21-wide per-thread arrays (isoprops incl.
d ln phi / dx, cross-composition tables, sigma derivatives), checked against the oracle.  Run under compute-sanitizer by tools/gpu_round_r02p.sh."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_ref
import parity_suite as ps

be = ps.GpuBackend()
name = "IG_GERG2008_21"
ps.isoprops(be, oracle_ref, name, n=40)
print("isoprops ok")
ps.cross_composition(be, oracle_ref, name, n=40)
print("cross ok")
ps.sigma_derivs(be, oracle_ref, name, n=40)
print("sigma ok")
