"""CPU tier: the kernels' per-point functions replayed on the host (tests/emul) against the oracle and the golden files.
This checks the MATHEMATICS the GPU will execute (same source, same parameter blocks) where no GPU is available."""
import pytest

import parity_suite as ps

AR_CASES = ["PR1", "PR2k", "SRK3", "cubicTwuMC", "vdW1", "vdW2", "PCSAFT1", "PCSAFT2", "PCSAFT3k", "GERG10", "GERG21", "GERG2004_3", "GERG_H2O_CO2",
            "MIE1", "MIE3", "GSAFT_WATER", "GSAFT_C4", "GSAFT_CR1"]


@pytest.mark.parametrize("name", AR_CASES)
def test_ar_tables(emul, oracle, name):
    ps.ar_tables(ps.EmulBackend(emul), oracle, name, n=1500)


@pytest.mark.parametrize("name", ["PR2k", "SRK3", "vdW2", "PCSAFT2", "PCSAFT3k", "MIE3", "GSAFT_C4", "GSAFT_CR1", "GERG2004_3", "GERG_H2O_CO2", "GERG10"])
def test_atrhoxi(emul, oracle, name):
    ps.atrhoxi(ps.EmulBackend(emul), oracle, name, n=200)


@pytest.mark.parametrize("name", ["PR2k", "SRK3", "PCSAFT2", "PCSAFT3k", "GERG2004_3", "GERG_H2O_CO2", "GERG10", "GERG21", "MIE3", "GSAFT_C4", "GSAFT_CR1"])
def test_isochoric(emul, oracle, name):
    ps.isochoric(ps.EmulBackend(emul), oracle, name, n=300)


@pytest.mark.parametrize("name", ["PR2k", "SRK3", "GERG2004_3", "GERG_H2O_CO2"])
def test_cross_composition(emul, oracle, name):
    ps.cross_composition(ps.EmulBackend(emul), oracle, name, n=100)


@pytest.mark.parametrize("name", ["PR2k", "SRK3", "PCSAFT2", "PCSAFT3k", "GERG2004_3", "GERG_H2O_CO2", "GERG10"])
def test_isoprops(emul, oracle, name):
    ps.isoprops(ps.EmulBackend(emul), oracle, name, n=100)


def test_gerg_vs_aga8_golden(emul):
    ps.gerg_vs_aga8_golden(ps.EmulBackend(emul))


def test_edge_cases(emul, oracle):
    ps.edge_cases(ps.EmulBackend(emul), oracle)
