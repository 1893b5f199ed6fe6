"""60-digit mpmath evaluation of the quantum-corrected Peng-Robinson formulas (simple_cubics.hpp:486-541) against the oracle and the host replay of the
kernels: Lambda_20 / 30 / 40 and the X_i^2 / X_i^3 entries of get_ATrhoXi, in units of the family scale S.  Run from the repo root:
    PYTHONPATH=. python tools/adjudicate_qcpr.py QCPR_H2NE
The numbers behind REF_NOISE_S / floor_limit for the QCPR cases in tests/parity_suite.py."""
import sys; sys.path.insert(0,'tests')
import numpy as np, mpmath as mp, cases, oracle_ref, emul_lib, parity_suite as ps
mp.mp.dps = 60
name = sys.argv[1] if len(sys.argv) > 1 else "QCPR_FUNNY_R"
spec, T, rho, x = cases.points(name, 100)
M = spec["model"]; R = mp.mpf(M.get("R / J/mol/K", 8.31446261815324))
def alphar(T, rho, xs):
    N = len(xs)
    ai=[];bi=[]
    for i in range(N):
        Tc=mp.mpf(M["Tcrit / K"][i]); pc=mp.mpf(M["pcrit / Pa"][i]); L,Mm,Nn=[mp.mpf(M[k][i]) for k in ("Ls","Ms","Ns")]
        A=mp.mpf(M["As"][i]); B=mp.mpf(M["Bs"][i])
        Tr=T/Tc
        al=Tr**(Nn*(Mm-1))*mp.exp(L*(1-Tr**(Mm*Nn)))
        ai.append(mp.mpf(0.45723552892138218938)*(Tc*R)**2/pc*al)
        bi.append(mp.mpf(0.07780)*Tc*R/pc*(1+A/(T+B))**3/(1+A/(Tc+B))**3)
    a=b=c=mp.mpf(0)
    for i in range(N):
        c+=xs[i]*mp.mpf(M["cs / m^3/mol"][i])
        for j in range(N):
            b+=xs[i]*xs[j]*(bi[i]+bi[j])/2*(1-mp.mpf(M["lmat"][i][j]))
            a+=xs[i]*xs[j]*mp.sqrt(ai[i]*ai[j])*(1-mp.mpf(M["kmat"][i][j]))
    r=1/(1/rho+c)
    D1=mp.mpf(1.0+1.4142135623730951); D2=mp.mpf(1.0-1.4142135623730951)
    return -mp.log(1-b*r) - a/(R*T)*mp.log((D1*b*r+1)/(D2*b*r+1))/(b*(D1-D2))
m = oracle_ref.RefModel(spec); be = ps.EmulBackend(emul_lib)
if x.shape[0] == 2:
    got = be.ATrhoXi_table(spec, T, rho, x)
    for comp in (0,1):
      for NX in (2,3):
        ref = m.ATrhoXi(0,0,comp,NX,T,rho,x)
        ref0 = np.abs(np.stack([m.ATrhoXi(0,0,c2,k,T,rho,x) for c2 in (0,1) for k in (1,2,3)]+[m.Ar0n(1,T,rho,x)[0]])).max(axis=0)
        eo=[];ek=[]
        for k in range(T.size):
            Tk=mp.mpf(float(T[k])); rk=mp.mpf(float(rho[k])); xs=[mp.mpf(float(x[0,k])),mp.mpf(float(x[1,k]))]
            f=lambda s: alphar(Tk,rk,[xs[0]+(s if comp==0 else 0), xs[1]+(s if comp==1 else 0)])
            tr=mp.taylor(f,0,NX)[NX]*mp.factorial(NX)
            S=ref0[k]
            eo.append(abs(float(ref[k]-tr))/S); ek.append(abs(float(got[comp,0,0,NX][k]-tr))/S)
        print(comp,NX,"oracle max %.2e mean %.2e | kernel max %.2e mean %.2e"%(max(eo),np.mean(eo),max(ek),np.mean(ek)))

ref = m.Arn0(4,T,rho,x); got = be.ar_table(spec,4,0,T,rho,x)[:,0]
S = np.abs(ref/np.array([1,1,2,6,24])[:,None]).max(axis=0)
for n in (2,3,4):
    eo=[];ek=[]
    for k in range(0,T.size,3):
        Tk=mp.mpf(float(T[k])); rk=mp.mpf(float(rho[k])); xs=[mp.mpf(float(v)) for v in x[:,k]]
        f=lambda s: alphar(1/(1/Tk+s),rk,xs)
        tr=mp.taylor(f,0,n)[n]*mp.factorial(n)/Tk**n
        sc=S[k]*mp.factorial(n)
        eo.append(abs(float(ref[n][k]-tr))/sc); ek.append(abs(float(got[n][k]-tr))/sc)
    print(n,"oracle max %.2e mean %.2e | kernel max %.2e mean %.2e"%(max(eo),np.mean(eo),max(ek),np.mean(ek)))
