"""Synthetic state-point workloads of BASELINE.json's configs and the deterministic counter-hash generator,
in numpy.  The device-side generator (teqpb200_synth_points in csrc/tq_capi.cu) produces the same uniforms from the same
(seed, point index, stream index), so CPU-side checks and GPU-side benches see the same points.

    u_j(k) = (splitmix64(seed ^ (k * 0x9E3779B97F4A7C15 + j)) >> 11) * 2^-53
    T = Tlo + (Thi - Tlo) u_0 ;  rho = exp(ln rho_lo + (ln rho_hi - ln rho_lo) u_1)
    x_i = z0_i exp(0.2 (2 u_{2+i} - 1)) / sum_i(...)
"""
import numpy as np

SEED = 20261104
_M = np.uint64(0xFFFFFFFFFFFFFFFF)
_G = np.uint64(0x9E3779B97F4A7C15)


def _u01(seed, k, j):
    with np.errstate(over="ignore"):
        z = np.uint64(seed) ^ (k.astype(np.uint64) * _G + np.uint64(j))
        z = z + _G
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return (z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def synth_points(npts, k0, Tlo, Thi, rholo, rhohi, z0, seed=SEED):
    """Returns T[npts], rho[npts], x[ncomp][npts] (float64, C-contiguous)."""
    k = np.arange(k0, k0 + npts, dtype=np.uint64)
    T = Tlo + (Thi - Tlo) * _u01(seed, k, 0)
    rho = np.exp(np.log(rholo) + (np.log(rhohi) - np.log(rholo)) * _u01(seed, k, 1))
    z0 = np.asarray(z0, dtype=np.float64)
    x = np.empty((len(z0), npts))
    s = np.zeros(npts)
    for i in range(len(z0)):
        x[i] = z0[i] * np.exp(0.2 * (2.0 * _u01(seed, k, 2 + i) - 1.0))
        s += x[i]          # summed in index order, then divided (same op order as the device generator)
    for i in range(len(z0)):
        x[i] /= s
    return T, rho, x


# AGA8 natural-gas compositions used by the reference's GERG tests (src/tests/catch_test_GERG.cxx:136 and :130), mole percent
_C3_NAMES = ["methane", "nitrogen", "carbondioxide", "ethane", "propane", "isobutane", "n-butane", "isopentane", "n-pentane", "n-hexane"]
_C3_Z = [96.370276, 0.291926, 0.298014, 1.400282, 0.780001, 0.14425, 0.460146, 0.0759, 0.0549, 0.124306]
# C5: all 21 GERG-2008 components in the component order of the AGA8 test code (src/tests/catch_test_GERG.cxx:512) with the AGA8 gas row of
# catch_test_GERG.cxx:130 (mole percent; 15 of 21 non-zero).  SURVEY.md section 8d: the zeros become 1e-4 (mole fraction) before
# renormalising so that the Hessian / d ln phi / dx are defined; the trace components are where H_ij/x_i loses digits.
_GERG21 = ["methane", "nitrogen", "carbondioxide", "ethane", "propane", "isobutane", "n-butane", "isopentane", "n-pentane", "n-hexane",
           "n-heptane", "n-octane", "n-nonane", "n-decane", "hydrogen", "oxygen", "carbonmonoxide", "water", "hydrogensulfide", "helium", "argon"]
_C5_ROW = [98.110602, 0.813399, 0.1209, 0.611199, 0.2153, 0.0339, 0.0453, 0.0115, 0.00925, 0.006, 0.0061, 0.0028, 0.0003, 0.00005, 0, 0, 0, 0, 0,
           0.0134, 0]
_C5_Z = [v / 100.0 if v != 0 else 1e-4 for v in _C5_ROW]


# C4: methane + ethane (library values) + water (Dufal-2015 values, src/tests/catch_test_SAFTgeneric.cxx:104-128) + methane + ethane (library values,
# include/teqp/models/saftvrmie.hpp:51-52).  Only water carries association sites, and it is the LAST component: the reference reads
# the Dufal pair tables of (i, j) as those of (i, N-1), so with water anywhere else its own association
# would be evaluated with epsilon_HB(water, last) = 0 and vanish
_C4_SPEC = {"kind": "genericSAFT", "model": {
    "nonpolar": {"kind": "SAFT-VR-Mie", "model": {"coeffs": [
        {"name": "Methane", "BibTeXKey": "Lafitte-JCP-2001", "m": 1.0, "sigma_Angstrom": 3.7412, "epsilon_over_k": 153.36, "lambda_r": 12.650, "lambda_a": 6.0},
        {"name": "Ethane", "BibTeXKey": "Lafitte-JCP-2001", "m": 1.4373, "sigma_Angstrom": 3.7257, "epsilon_over_k": 206.12, "lambda_r": 12.400, "lambda_a": 6.0},
        {"name": "Water", "BibTeXKey": "Dufal-2015", "m": 1.0, "sigma_Angstrom": 3.0555, "epsilon_over_k": 418.00, "lambda_r": 35.823, "lambda_a": 6.0}]}},
    "association": {"kind": "Dufal", "model": {
        "sigma / m": [3.7412e-10, 3.7257e-10, 3.0555e-10],
        "epsilon / J/mol": [1275.1063705805507, 1713.7770350136118, 3475.445374388054],
        "lambda_r": [12.650, 12.400, 35.823],
        "epsilon_HB / J/mol": [0.0, 0.0, 13303.140189045183],
        "K_HB / m^3": [0.0, 0.0, 496.66e-30],
        "kmat": [[0.0, 0.0, 0.0], [0.0, 0.0, 0.0], [0.0, 0.0, 0.0]],
        "Delta_rule": "Dufal",
        "molecule_sites": [[], [], ["e", "e", "H", "H"]]}}}}


def _norm(z):
    z = np.asarray(z, dtype=np.float64)
    return (z / z.sum()).tolist()


WORKLOADS = {
    # C1: Peng-Robinson pure methane (values of src/tests/catch_test_cubics.cxx:28-30), Ar00..Ar04
    "C1": dict(spec={"kind": "PR", "model": {"Tcrit / K": [190.564], "pcrit / Pa": [4599200], "acentric": [0.011]}},
               T=(100.0, 600.0), rho=(1.0, 25000.0), z0=[1.0], npts=10**6, what="Ar0n", nmax=4, bytes_per_point=56, bound="hbm", x_broadcast=True),
    # C2: PC-SAFT methane + ethane (library values, include/teqp/models/pcsaft.hpp:53-54)
    "C2": dict(spec={"kind": "PCSAFT", "model": {"names": ["Methane", "Ethane"]}},
               T=(150.0, 450.0), rho=(1.0, 20000.0), z0=[0.5, 0.5], npts=10**7, what="Arxy+lnphi", bytes_per_point=96),
    # C3: GERG-2008 10-component natural gas, Ar00..Ar04 -- the headline
    "C3": dict(spec={"kind": "GERG2008resid", "model": {"names": _C3_NAMES}},
               T=(250.0, 450.0), rho=(10.0, 25000.0), z0=_norm(_C3_Z), npts=10**8, what="Ar0n", nmax=4, bytes_per_point=136),
    # C4: SAFT-VR-Mie ternary with Dufal association on water: the get_ATrhoXi table (23 combinations x 3 components)
    "C4": dict(spec=_C4_SPEC, T=(280.0, 500.0), rho=(10.0, 30000.0), z0=[0.3, 0.3, 0.4], npts=10**7, what="ATrhoXi", bytes_per_point=40 + 8 * 108),
    # C5: GERG-2008 all 21 components, isochoric Hessian / fugacity composition derivatives; 10^6 points per GPU (weak), 8*10^6 total (strong)
    "C5": dict(spec={"kind": "GERG2008resid", "model": {"names": _GERG21}},
               T=(250.0, 450.0), rho=(10.0, 25000.0), z0=_norm(_C5_Z), npts=10**6, npts_strong=8 * 10**6, what="fgradHessian", nmax=4,
               bytes_per_point=8 * 22 + 8 * (1 + 21 + 441)),
}
# C5 with the fugacity composition derivatives in the same step (BASELINE.json configs[4] names both)
WORKLOADS["C5x"] = dict(WORKLOADS["C5"], what="fgradHessian+dlnphi", bytes_per_point=8 * 22 + 8 * (1 + 3 * 21 + 2 * 441))


def workload_points(name, npts, k0=0):
    w = WORKLOADS[name]
    return synth_points(npts, k0, w["T"][0], w["T"][1], w["rho"][0], w["rho"][1], w["z0"])


def shard_range(npts, rank, world):
    """Contiguous even split of [0, npts) over `world` ranks / devices, remainder to the last one -- the same rule the
    multi-device host driver in csrc/tq_capi.cu (run_host) applies."""
    per = npts // world
    a = per * rank
    b = npts if rank == world - 1 else a + per
    return a, b


def pr_saturation_guess(T, Tc, pc, acentric):
    """rough saturation densities for starting values: Wilson's vapour pressure, then the liquid and vapour roots of the Peng-Robinson cubic"""
    R = 8.31446261815324
    T = np.asarray(T, dtype=np.float64)
    p = pc * np.exp(5.373 * (1 + acentric) * (1 - Tc / T))
    mm = 0.37464 + 1.54226 * acentric - 0.26992 * acentric**2
    a = 0.45723552892138218938 * R**2 * Tc**2 / pc * (1 + mm * (1 - np.sqrt(T / Tc)))**2
    b = 0.077796073903888455972 * R * Tc / pc
    rL = np.empty_like(T); rV = np.empty_like(T)
    for k in range(T.size):
        A = a[k] * p[k] / (R * T[k])**2; B = b * p[k] / (R * T[k])
        Z = np.roots([1, -(1 - B), A - 3 * B * B - 2 * B, -(A * B - B * B - B**3)])
        Z = np.sort(Z[np.abs(Z.imag) < 1e-9].real)
        Z = Z[Z > B]
        rL[k] = p[k] / (Z[0] * R * T[k]); rV[k] = p[k] / (Z[-1] * R * T[k])
    return rL, rV
