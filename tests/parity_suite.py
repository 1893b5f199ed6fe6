"""The parity checks proper, written once over a 'backend':
   * tests/test_emul_parity.py runs them on the host replay of the kernels' per-point functions (CPU tier, no GPU here);
   * tests/test_gpu_parity.py runs them through the C ABI of libteqpb200.so on the B200 (-m gpu).
The checker is always the oracle (reference templates + AD shim) or a committed golden file."""
import json
import os

import numpy as np

import cases
import parity

HERE = os.path.dirname(os.path.abspath(__file__))


class EmulBackend:
    name = "emul"

    def __init__(self, emul):
        self.e = emul

    def ar_table(self, spec, NT, ND, T, rho, x):
        if spec["kind"].startswith("GERG") and T.size:
            # the product evaluates GERG tables with the per-model GENERATED kernel: replay that source (compiled for the host) -- the
            # interpreter kernel's per-point function is replayed as well and must agree to the parity rule
            return self.e.jit_ar_table(spec, NT, ND, T, rho, x)
        return self.e.ar_table(spec, NT, ND, T, rho, x)

    def ATrhoXi_table(self, spec, T, rho, x):
        return self.e.ATrhoXi_table(spec, T, rho, x)

    def cross_table(self, spec, idx, T, rho, x):
        return self.e.cross_table(spec, idx, T, rho, x)

    def isoprops(self, spec, T, rv):
        return self.e.isoprops(spec, T, rv)

    def isochoric(self, spec, T, rv):
        f, g, H, lnphi, phi = self.e.isochoric(spec, T, rv)
        nc = rv.shape[0]
        return f, g, H.reshape(nc, nc, -1), lnphi, phi


class GpuBackend:
    """Everything goes through the C ABI (ctypes) with HOST buffers; device-pointer calls are covered separately."""
    name = "gpu"

    def __init__(self):
        import teqp_b200
        self.t = teqp_b200
        self.models = {}

    def model(self, spec):
        key = json.dumps(spec, sort_keys=True)
        if key not in self.models:
            self.models[key] = self.t.make_model(spec)
        return self.models[key]

    def ar_table(self, spec, NT, ND, T, rho, x):
        m = self.model(spec)
        pairs = [(i, j) for i in range(NT + 1) for j in range(ND + 1)]
        if NT == 0:
            return m.get_Ar0n_batch(ND, T, rho, x).reshape(1, ND + 1, -1)
        if ND == 0:
            return m.get_Arn0_batch(NT, T, rho, x).reshape(NT + 1, 1, -1)
        return m.get_Arxy_batch(pairs, T, rho, x).reshape(NT + 1, ND + 1, -1)

    def ATrhoXi_table(self, spec, T, rho, x):
        return self.model(spec).get_ATrhoXi_table_batch(T, rho, x)

    def cross_table(self, spec, idx, T, rho, x):
        m = self.model(spec)
        return m.get_ATrhoXiXj_table_batch(idx[0], idx[1], T, rho, x) if len(idx) == 2 else m.get_ATrhoXiXjXk_table_batch(idx[0], idx[1], idx[2], T, rho, x)

    def isoprops(self, spec, T, rv):
        return self.model(spec).isochoric_properties_batch(T, rv)

    def isochoric(self, spec, T, rv):
        m = self.model(spec)
        f, g, H = m.build_Psir_fgradHessian_batch(T, rv)
        return f, g, H, m.get_ln_fugacity_coefficients_batch(T, rv), m.get_fugacity_coefficients_batch(T, rv)


# SAFT-type models: the reference evaluates the hard-sphere term as a difference of O(1) summands
# (zeta_2^3/(zeta_3 (1-zeta_3)^2) against (zeta_2^3/zeta_3^2) ln(1-zeta_3), saftvrmie.hpp:697, pcsaft.hpp:143) and Eq. A15's J
# integral as a difference of O(lambda) powers, so its OWN values carry an absolute rounding noise of a few 1e-16 * S_HS that no
# second implementation can reproduce bit for bit.  S_HS is the absolute floor scale for these families (parity.py rule).
# Measured on the reference itself (tools/ref_self_consistency.py): its Ar00 from the 1/T sweep and from the rho sweep -- the
# same number mathematically -- differ by up to 3e-12 relative (1.5e-15 absolute) for SAFT-VR-Mie at low density, and by 2e-11
# relative (7e-14 absolute) for pure 4C water, where the closed-form site fraction (association.hpp:403-418) divides an
# O(kappa) difference of O(1) terms by kappa ~ rho.  The floor scales below are that noise with a margin.
SAFT_SUMMAND_SCALE = 0.02
SUMMAND_SCALE = {"GSAFT_WATER": 1.0}


def extra_scale(name):
    return SUMMAND_SCALE.get(name, SAFT_SUMMAND_SCALE if name.startswith(("MIE", "GSAFT")) else 0.0)


def ar_tables(be, oracle, name, n=3000):
    """Lambda_nm for n<=2, m<=4 (every compile-time get_Ar{i}{j} of the reference), Ar{n}0n up to 4, Ar0{n}n up to 6."""
    spec, T, rho, x = cases.points(name, n)
    m = oracle.RefModel(spec)
    ref = cases.full_table_oracle(m, T, rho, x)
    r40 = m.Arn0(4, T, rho, x); r06 = m.Ar0n(6, T, rho, x)
    # family scale in Taylor-coefficient space over everything the reference offers at this point
    fam = np.maximum(parity.coeff_scale(ref), np.maximum(parity.coeff_scale(r40[:, None, :]), parity.coeff_scale(r06[None, :, :])))
    fam = np.maximum(fam, extra_scale(name))
    got = be.ar_table(spec, 2, 4, T, rho, x)
    for i in range(3):
        for j in range(5):
            parity.check("%s Ar%d%d" % (name, i, j), got[i, j], ref[i, j], fam * parity.fact(i, j))
    g40 = be.ar_table(spec, 4, 0, T, rho, x)[:, 0]
    for k in range(5):
        parity.check("%s Ar%d0n" % (name, k), g40[k], r40[k], fam * parity.fact(k))
    g06 = be.ar_table(spec, 0, 6, T, rho, x)[0]
    for k in range(7):
        parity.check("%s Ar0%dn" % (name, k), g06[k], r06[k], fam * parity.fact(k))
    # the headline call itself: get_Ar04n, and how often the floor is needed for the low orders
    g04 = be.ar_table(spec, 0, 4, T, rho, x)[0]
    for k in range(5):
        parity.check("%s Ar04n[%d]" % (name, k), g04[k], r06[k], fam * parity.fact(k), max_floor_frac=(0.02 if (k < 3 and name in ("GERG10", "GERG21")) else 1.0))


COMBOS_ATRHOXI = [(0, 0, 1), (0, 0, 2), (0, 0, 3), (1, 0, 0), (1, 0, 1), (1, 0, 2), (1, 0, 3), (0, 1, 0), (0, 1, 1), (0, 1, 2), (0, 1, 3),
                  (2, 0, 0), (2, 0, 1), (2, 0, 2), (2, 0, 3), (1, 1, 0), (1, 1, 1), (1, 1, 2), (1, 1, 3), (0, 2, 0), (0, 2, 1), (0, 2, 2), (0, 2, 3)]


def atrhoxi(be, oracle, name, n=400):
    """all 23 (NT, ND, NXi) combinations of get_ATrhoXi_runtime (derivs.hpp:261-284) for every component"""
    spec, T, rho, x = cases.points(name, n)
    m = oracle.RefModel(spec)
    got = be.ATrhoXi_table(spec, T, rho, x)
    nc = x.shape[0]
    fam = np.maximum(parity.family_scale(got, (0, 1, 2, 3)), extra_scale(name))
    for i in range(nc):
        for (NT, ND, NX) in COMBOS_ATRHOXI:
            ref = m.ATrhoXi(NT, ND, i, NX, T, rho, x)
            # third composition derivatives are differences of O(family) terms in both implementations
            parity.check("%s ATrhoXi NT%d ND%d X%d^%d" % (name, NT, ND, i, NX), got[i, NT, ND, NX], ref, fam * (30.0 if NX == 3 else 1.0))


COMBOS_XIXJ = [(0, 0, 1, 0), (0, 0, 2, 0), (0, 0, 0, 1), (0, 0, 0, 2), (0, 0, 1, 1), (1, 0, 1, 0), (1, 0, 2, 0), (1, 0, 0, 1), (1, 0, 0, 2), (1, 0, 1, 1),
               (0, 1, 1, 0), (0, 1, 2, 0), (0, 1, 0, 1), (0, 1, 0, 2), (0, 1, 1, 1)]
COMBOS_XIXJXK = [(0, 0, 0, 1, 1), (0, 0, 1, 0, 1), (0, 0, 1, 1, 0), (0, 0, 1, 1, 1), (1, 0, 0, 1, 1), (1, 0, 1, 0, 1), (1, 0, 1, 1, 0), (1, 0, 1, 1, 1),
                 (0, 1, 0, 1, 1), (0, 1, 1, 0, 1), (0, 1, 1, 1, 0), (0, 1, 1, 1, 1)]


def cross_composition(be, oracle, name, n=200):
    """get_ATrhoXiXj (15 combinations, derivs.hpp:294-309) and get_ATrhoXiXjXk (12 combinations, :316-328) for a few index tuples"""
    spec, T, rho, x = cases.points(name, n)
    m = oracle.RefModel(spec)
    nc = x.shape[0]
    pairs = [(0, 1), (nc - 1, 0)] if nc > 2 else [(0, 1)]
    for (i, j) in pairs:
        got = be.cross_table(spec, (i, j), T, rho, x)
        fam = np.maximum(parity.family_scale(got, (0, 1, 2, 3)), extra_scale(name))
        for (NT, ND, NXi, NXj) in COMBOS_XIXJ:
            ref = m.ATrhoXiXj(NT, ND, i, NXi, j, NXj, T, rho, x)
            parity.check("%s ATrhoXiXj NT%d ND%d x%d^%d x%d^%d" % (name, NT, ND, i, NXi, j, NXj), got[NT, ND, NXi, NXj], ref, fam)
    if nc >= 3:
        for (i, j, k) in [(0, 1, 2), (nc - 1, 0, 1)]:
            got = be.cross_table(spec, (i, j, k), T, rho, x)
            fam = np.maximum(parity.family_scale(got, (0, 1, 2, 3, 4)), extra_scale(name))
            for (NT, ND, a, b, c) in COMBOS_XIXJXK:
                ref = m.ATrhoXiXjXk(NT, ND, i, a, j, b, k, c, T, rho, x)
                parity.check("%s ATrhoXiXjXk NT%d ND%d %d%d%d" % (name, NT, ND, a, b, c), got[NT, ND, a, b, c], ref, fam)


def isochoric(be, oracle, name, n=600, lnphi=True):
    spec, T, rho, x = cases.points(name, n)
    rv = np.ascontiguousarray(x * rho)
    m = oracle.RefModel(spec)
    rf, rg, rH = m.fgradHessian(T, rv)
    nc = rv.shape[0]
    rH = rH.reshape(nc, nc, -1)
    f, g, H, lp, ph = be.isochoric(spec, T, rv)
    sg = np.maximum(parity.family_scale(rg, 0), extra_scale(name) * m.get_R(x[:, 0]) * T)
    parity.check(name + " Psir", f, rf, np.maximum(np.max(np.abs(rg * rv), axis=0), sg * rho))
    parity.check(name + " grad Psir", g, rg, sg)
    # rho_j * H_ij has the units of the gradient: scale the Hessian rows by that
    sH = np.maximum(parity.family_scale(rH, (0, 1)), sg / np.min(rv, axis=0))
    parity.check(name + " Hessian Psir", H, rH, sH)
    parity.check(name + " phi", ph, m.isochoric_array(3, T, rv), 0.0, rtol=5e-12)
    if lnphi and not name.startswith(("MIE", "GSAFT")):   # (the oracle's ln(phi) family is wired for GERG / cubic / PC-SAFT; phi covers the rest)
        # ln phi = grad/(RT) - ln Z: both parts are O(grad/RT), the difference can be arbitrarily small
        RT = m.get_R(x[:, 0]) * T
        parity.check(name + " ln phi", lp, m.lnphi_family(0, T, rv), sg / RT)


def isoprops(be, oracle, name, n=300):
    """the derived isochoric properties against the reference's AbstractModel methods (teqpcpp.hpp:52-75) and its ln(phi) derivative family"""
    spec, T, rho, x = cases.points(name, n)
    rv = np.ascontiguousarray(x * rho)
    m = oracle.RefModel(spec)
    got = be.isoprops(spec, T, rv)
    nc = rv.shape[0]
    R = m.get_R(x[:, 0]); RT = R * T
    rg = m.isochoric_array(0, T, rv)
    sg = np.maximum(parity.family_scale(rg, 0), extra_scale(name) * RT)              # scale of dPsir/drho_i
    sp = np.maximum(sg * rho, np.abs(m.isochoric_scalar(0, T, rv)))                  # scale of pressures
    parity.check(name + " pr", got["pr"], m.isochoric_scalar(0, T, rv), sp)
    parity.check(name + " splus", got["splus"], m.isochoric_scalar(1, T, rv), sg / RT)
    tg = m.isochoric_array(5, T, rv)
    st = np.maximum(parity.family_scale(tg, 0), sg / T)
    parity.check(name + " dpdT", got["dpdT"], m.isochoric_scalar(2, T, rv), st * rho + R * rho)
    parity.check(name + " d2PsirdTdrhoi", got["d2PsirdTdrhoi"], tg, st)
    parity.check(name + " dpdrhovec", got["dpdrhovec"], m.isochoric_array(6, T, rv), RT + sg)
    pmv = m.isochoric_array(4, T, rv)
    parity.check(name + " pmv", got["pmv"], pmv, parity.family_scale(pmv, 0), rtol=1e-11)
    parity.check(name + " chempotVLE", got["chempotVLE"], m.isochoric_array(1, T, rv), sg + RT * 10)
    parity.check(name + " dchempotdT", got["dchempotdT"], m.isochoric_array(2, T, rv), st + R * 10)
    if not name.startswith(("MIE", "GSAFT")):
        # d/dv: the reference returns an Eigen expression that refers to a destroyed temporary (derivs.hpp:1362-1366, auto return of
        # `get_..._drho(...)*drhodv`), so its own output is garbage (measured: 1e-304 or -0); the documented definition is checked instead
        for what, key, sc in [(1, "dlnphi_dT", st / RT + sg / RT / T), (2, "dlnphi_drho", sg / RT / rho), (4, "dlnphi_dv", sg / RT * rho)]:
            ref = m.lnphi_family(what, T, rv) if what != 4 else m.lnphi_family(2, T, rv) * (-rho * rho)
            parity.check(name + " " + key, got[key], ref, np.maximum(sc, parity.family_scale(ref, 0)), rtol=1e-11)
        ref = m.lnphi_family(3, T, rv)
        parity.check(name + " dlnphi_dx", got["dlnphi_dx"], ref, parity.family_scale(ref, 0), rtol=1e-11)


def gerg_vs_aga8_golden(be):
    """GERG-2008, 21 components, 187 validation state points: tau^i delta^j derivatives from the NIST AGA8 analytic code."""
    g = json.load(open(os.path.join(HERE, "golden", "aga8_gerg2008.json")))
    spec = {"kind": "GERG2008resid", "model": {"names": g["components"]}}
    pts = g["points"]
    T = np.array([p["T"] for p in pts]); rho = np.array([p["rho"] for p in pts]); x = np.array([p["x"] for p in pts]).T.copy()
    got = be.ar_table(spec, 2, 2, T, rho, x)
    for key, (i, j) in {"ar00": (0, 0), "ar01": (0, 1), "ar02": (0, 2), "ar10": (1, 0), "ar11": (1, 1), "ar20": (2, 0)}.items():
        np.testing.assert_allclose(got[i, j], [p[key] for p in pts], rtol=2e-11, atol=1e-13, err_msg=key)


def edge_cases(be, oracle):
    """rho = 0 (teqp's powi branch, GERG.hpp:110-114), x_i = 0 (skipped reducing pairs, GERG.hpp:189), empty batch"""
    spec = cases.SPECS["GERG10"][0]
    m = oracle.RefModel(spec)
    z = np.array(cases.SPECS["GERG10"][3])
    T = np.array([300.0, 300.0, 250.0]); rho = np.array([0.0, 5000.0, 12000.0])
    x = np.tile(z[:, None], (1, 3))
    x[3:, 1] = 0.0; x[:, 1] /= x[:, 1].sum()          # seven absent components
    x[:, 2] = 0.0; x[0, 2] = 1.0                       # pure methane through the mixture model
    ref = m.Ar0n(4, T, rho, x)
    got = be.ar_table(spec, 0, 4, T, rho, x)[0]
    parity.check("GERG edge cases", got, ref, parity.family_scale(ref, 0))
    assert np.all(got[:, 0] == 0.0)
    empty = be.ar_table(spec, 0, 4, np.zeros(0), np.zeros(0), np.zeros((10, 0)))
    assert empty.shape[-1] == 0
    # cubic at rho = 0
    spec2, T2, rho2, x2 = cases.points("PR2k", 4)
    rho2[:] = 0.0
    ref2 = oracle.RefModel(spec2).Ar0n(4, T2, rho2, x2)
    got2 = be.ar_table(spec2, 0, 4, T2, rho2, x2)[0]
    parity.check("PR rho=0", got2, ref2, 1e-300)
