"""Fits the polynomial used by tq::fast_exp (csrc/tq_math.cuh): exp(r) ~ 1 + r + r^2 q(r), |r| <= ln2/2, q of degree DEG-2.
Remez exchange on the relative error in 60-digit arithmetic (mpmath); prints the coefficients as C hex-float literals and
the achieved error, then checks the double-precision evaluation against mpmath on random arguments."""
import sys
import mpmath as mp
import numpy as np

mp.mp.dps = 60
DEG = int(sys.argv[1]) if len(sys.argv) > 1 else 11
A = mp.log(2) / 2 * mp.mpf("1.0001")
nq = DEG - 1  # number of coefficients of q (degree DEG-2)


def f(r):  # target for q
    r = mp.mpf(r)
    if abs(r) < mp.mpf("1e-25"):
        return mp.mpf(1) / 2
    return (mp.exp(r) - 1 - r) / (r * r)


def w(r):  # weight so that the error of q maps to the RELATIVE error of exp: err_exp/exp = r^2 err_q / exp(r)
    return r * r / mp.exp(r)


# near-minimax: interpolate q at the Chebyshev nodes of [-A, A] (the exp error is r^2/exp(r) times the error of q)
n = nq
xs = [A * mp.cos(mp.pi * (2 * k + 1) / (2 * n)) for k in range(n)]
M = mp.matrix(n, n); b = mp.matrix(n, 1)
for i, x in enumerate(xs):
    for j in range(nq):
        M[i, j] = x ** j
    b[i] = f(x)
sol = mp.lu_solve(M, b)
c = [sol[j] for j in range(nq)]
grid = [-A + 2 * A * mp.mpf(k) / 2000 for k in range(2001)]
mx = max(abs((f(x) - sum(c[j] * x ** j for j in range(nq))) * w(x)) for x in grid)
print("degree", DEG, "max relative error (exact coefficients) %.3e" % float(mx))
coef = [1.0, 1.0] + [float(cj) for cj in c]
for k, v in enumerate(coef):
    print("c%d = %s  // %.17g" % (k, v.hex(), v))
# double-precision check of the full algorithm
rng = np.random.default_rng(1)
x = np.concatenate([rng.uniform(-700, 700, 200000), rng.uniform(-3, 3, 200000), rng.uniform(-0.4, 0.4, 100000)])
L2E = 1.4426950408889634; MAG = 6755399441055744.0
LN2H = float.fromhex("0x1.62e42fefa39efp-1"); LN2L = float.fromhex("0x1.abc9e3b39803fp-56")
t = x * L2E + MAG          # not fused here; close enough for an error survey
nf = t - MAG
r = (x - nf * LN2H) - nf * LN2L
p = np.full_like(r, coef[-1])
for v in coef[-2::-1]:
    p = p * r + v
got = np.ldexp(p, nf.astype(np.int64))
ref = np.array([float(mp.exp(mp.mpf(float(v)))) for v in x[::50]])
rel = np.abs(got[::50] - ref) / ref
print("double evaluation (unfused numpy): max rel err %.3e = %.2f ulp, mean %.2e" % (rel.max(), rel.max() / 2.2e-16, rel.mean()))
print("max |r|", np.abs(r).max())

# ---- the same for 2^r on |r| <= 1/2 (tq::fast_exp2): 2^r ~ 1 + r (ln2 + r q2(r)), coefficients of degree 1..DEG
B = mp.mpf("0.5") * mp.mpf("1.0001")
LN2 = mp.log(2)
nq2 = DEG - 1
f2 = lambda r: (mp.power(2, r) - 1 - r * LN2) / (r * r) if abs(r) > mp.mpf("1e-25") else LN2 ** 2 / 2
xs = [B * mp.cos(mp.pi * (2 * k + 1) / (2 * nq2)) for k in range(nq2)]
M = mp.matrix(nq2, nq2); b = mp.matrix(nq2, 1)
for i, x in enumerate(xs):
    for j in range(nq2):
        M[i, j] = x ** j
    b[i] = f2(x)
sol = mp.lu_solve(M, b)
c2 = [sol[j] for j in range(nq2)]
grid = [-B + 2 * B * mp.mpf(k) / 2000 for k in range(2001)]
mx2 = max(abs((f2(x) - sum(c2[j] * x ** j for j in range(nq2))) * x * x / mp.power(2, x)) for x in grid)
print("exp2 degree", DEG, "max relative error (exact coefficients) %.3e" % float(mx2))
coef2 = [1.0, float(LN2)] + [float(cj) for cj in c2]
print("exp2 coefficients c1..c%d:" % DEG)
print(", ".join(v.hex() for v in coef2[1:]))
