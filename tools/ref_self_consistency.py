#!/usr/bin/env python
"""How far the reference disagrees with ITSELF: Ar00 from get_Arn0 (Taylor sweep in 1/T) against Ar00 from get_Ar0n (sweep in rho).
This is the rounding noise of the reference formulation, the lower bound for any cross-implementation tolerance (tests/parity_suite.py)."""
import sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import oracle_ref, cases
for name in ["GSAFT_WATER","GSAFT_C4","GSAFT_CR1","MIE3","PCSAFT2","GERG10"]:
    spec,T,rho,x=cases.points(name,4000)
    m=oracle_ref.RefModel(spec)
    a=m.Arn0(2,T,rho,x)[0]; b=m.Ar0n(2,T,rho,x)[0]
    d=np.abs(a-b)
    k=np.argmax(d)
    print(name,"reference self-inconsistency Ar00 (T-sweep vs rho-sweep): max abs %.2e (value %.3e, rho %.1f), max rel %.2e"%(d[k],b[k],rho[k],(d/np.abs(b)).max()))
    for lo,hi in [(0,100),(100,1000),(1000,1e9)]:
        s=(rho>=lo)&(rho<hi)
        if s.any(): print("    rho in [%g,%g): max abs %.2e max rel %.2e"%(lo,hi,d[s].max(),(d[s]/np.abs(b[s])).max()))
