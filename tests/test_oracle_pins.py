"""Pins the oracle (reference templates + clean-room AD shim) against values that do NOT depend on the shim:
closed forms, the reference's stored golden scalars, and the AGA8 analytic GERG-2008 implementation (tests/golden)."""
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def test_vdW1_closed_forms(oracle):
    # /root/reference/src/tests/catch_tests.cxx:206-259
    a, b = 0.2283, 4.278e-5
    m = oracle.RefModel({"kind": "vdW1", "model": {"a": a, "b": b}})
    T = np.array([300.0, 450.0]); rho = np.array([3.0, 9000.0]); x = np.ones((1, 2))
    r = m.Ar0n(4, T, rho, x)
    y = b * rho
    np.testing.assert_allclose(r[2], (y / (y - 1)) ** 2, rtol=1e-14)
    np.testing.assert_allclose(r[3], -2 * (y / (y - 1)) ** 3, rtol=1e-14)
    np.testing.assert_allclose(r[4], 6 * (y / (y - 1)) ** 4, rtol=1e-14)


def test_pcsaft_golden_scalar(oracle):
    # /root/reference/src/tests/catch_test_PCSAFT.cxx:27
    m = oracle.RefModel({"kind": "PCSAFT", "model": {"names": ["Methane"]}})
    v = m.Ar0n(1, np.array([200.0]), np.array([300.0]), np.ones((1, 1)))[0, 0]
    assert v == pytest.approx(-0.032400020930842724, rel=1e-14)


def test_gerg2008_vs_aga8_golden(oracle):
    g = json.load(open(os.path.join(HERE, "golden", "aga8_gerg2008.json")))
    m = oracle.RefModel({"kind": "GERG2008resid", "model": {"names": g["components"]}})
    pts = g["points"]
    T = np.array([p["T"] for p in pts]); rho = np.array([p["rho"] for p in pts]); x = np.array([p["x"] for p in pts]).T.copy()
    for key, (i, j) in {"ar00": (0, 0), "ar01": (0, 1), "ar02": (0, 2), "ar10": (1, 0), "ar11": (1, 1), "ar20": (2, 0)}.items():
        ref = np.array([p[key] for p in pts])
        got = m.ArNM(i, j, T, rho, x)
        np.testing.assert_allclose(got, ref, rtol=2e-11, atol=1e-13, err_msg=key)
    # (the tabulated pressures of that table are NOT asserted: the reference itself has that comparison commented out,
    #  catch_test_GERG.cxx:696-698, and ~10 % of the rows deviate at the 1e-4 level)


def test_R_values(oracle):
    assert oracle.RefModel({"kind": "GERG2008resid", "model": {"names": ["methane", "ethane"]}}).get_R([0.5, 0.5]) == 8.314472
    assert oracle.RefModel({"kind": "PR", "model": {"Tcrit / K": [190.564], "pcrit / Pa": [4599200], "acentric": [0.011]}}).get_R([1.0]) == 8.31446261815324


def _pin(oracle, name, n):
    import ctypes as C
    out = (C.c_double * n)()
    assert getattr(oracle.lib(), name)(out) == 0, name
    return list(out)


def test_saftvrmie_clapeyron_goldens(oracle):
    """the reference's stored Clapeyron.jl values for SAFT-VR-Mie ethane at 300 K, 12000 mol/m3 (src/tests/catch_test_SAFTVRMie.cxx:57-80),
    at the reference's own tolerance (Catch2 Approx: 1.2e-5 relative; SURVEY App. E #5: they agree to ~2e-7, so they pin the terms, not 1e-12)"""
    a1kB, a2kB2, a3kB3, mono, chain = _pin(oracle, "ref_pin_saftvrmie_core", 5)
    for got, ref in [(a1kB, -694.2608061145818), (a2kB2, -6741.101051705587), (a3kB3, -81372.77460911816), (mono, -1.322739797471788),
                     (chain, -0.0950261207746853)]:
        assert got == pytest.approx(ref, rel=1.2e-5)


def test_association_goldens(oracle):
    """Delta matrix entries and site fractions of ethanol + water (CR1 / CS), src/tests/catch_test_association.cxx:28-52, WithinRel 1e-8"""
    d00, d03, d33, xa0, xa3 = _pin(oracle, "ref_pin_association", 5)
    assert d00 == pytest.approx(5.85623687e-27, rel=1e-8)
    assert d03 == pytest.approx(4.26510827e-27, rel=1e-8)
    assert d33 == pytest.approx(2.18581242e-27, rel=1e-8)
    assert xa0 == pytest.approx(0.06258400385436955, rel=1e-8)
    assert xa3 == pytest.approx(0.10938445109190545, rel=1e-8)


def test_dufal_water_site_fraction_golden(oracle):
    """non-bonded site fraction of Dufal-2015 water at 290 K (Clapeyron.jl), src/tests/catch_test_SAFTgeneric.cxx:160-180; agrees to 7e-15"""
    (xa,) = _pin(oracle, "ref_pin_dufal_water", 1)
    assert xa == pytest.approx(0.08922902098124778, rel=1e-13)


@pytest.mark.parametrize("name", ["PCSAFT2", "GERG2004_3", "MIE3"])
def test_shim_nested_duals_to_depth_5_against_binary128(oracle, name):
    """the AD shim's HigherOrderDual arithmetic at depth 4 and 5 (get_ATrhoXi with NT + ND + NXi = 4, 5: derivs.hpp:247-259 builds
    HigherOrderDual<iT + iD + iXi>) against an independent route: truncated-Taylor sweeps of the same templates in binary128
    (oracle/ref_truth.cpp uses Real<N, T> only, never Dual).  Agreement at 1e-9 of the block scale pins the Dual product / quotient / exp / log /
    pow rules to fifth order."""
    import cases
    spec, T, rho, x = cases.points(name, 40)
    m = oracle.RefModel(spec); tm = oracle.TruthModel(spec)
    tt = tm.ATrhoXi_table(T, rho, x)
    for i in range(x.shape[0]):
        for (NT, ND, NX) in [(2, 0, 3), (0, 2, 3), (1, 1, 3), (2, 0, 2), (1, 1, 2), (0, 2, 2), (1, 0, 3), (0, 1, 3)]:
            got = m.ATrhoXi(NT, ND, i, NX, T, rho, x)
            S = np.max(np.abs(tt[:, :NT + 1, :ND + 1, :NX + 1]), axis=(0, 1, 2, 3))
            assert np.all(np.abs(got - tt[i, NT, ND, NX]) <= 1e-9 * S), (name, i, NT, ND, NX)
