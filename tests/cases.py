"""Model specs and input ranges shared by the CPU (host replay) and GPU (C ABI) parity tests."""
import numpy as np

from teqp_b200.workloads import WORKLOADS, synth_points, workload_points

SPECS = {
    "PR1": ({"kind": "PR", "model": {"Tcrit / K": [190.564], "pcrit / Pa": [4599200], "acentric": [0.011]}}, (100, 600), (1, 25000), [1.0]),
    "PR2k": ({"kind": "PR", "model": {"Tcrit / K": [190.564, 305.322], "pcrit / Pa": [4599200, 4872200], "acentric": [0.011, 0.099],
                                      "kmat": [[0, 0.05], [0.05, 0]]}}, (150, 450), (1, 15000), [0.5, 0.5]),
    "SRK3": ({"kind": "SRK", "model": {"Tcrit / K": [190.564, 305.322, 369.8], "pcrit / Pa": [4599200, 4872200, 4248000],
                                       "acentric": [0.011, 0.099, 0.152]}}, (200, 450), (1, 12000), [0.4, 0.3, 0.3]),
    "cubicTwuMC": ({"kind": "cubic", "model": {"type": "PR", "Tcrit / K": [190.564, 305.322], "pcrit / Pa": [4599200, 4872200],
                                               "acentric": [0.011, 0.099], "R / J/mol/K": 8.3144598,
                                               "alpha": [{"type": "Twu", "c": [0.1, 0.9, 2.0]}, {"type": "Mathias-Copeman", "c": [0.4, -0.1, 0.3]}]}},
                   (150, 300), (1, 15000), [0.6, 0.4]),
    "vdW1": ({"kind": "vdW1", "model": {"a": 0.2283, "b": 4.278e-5}}, (150, 600), (1, 15000), [1.0]),
    "vdW2": ({"kind": "vdW", "model": {"Tcrit / K": [190.564, 305.322], "pcrit / Pa": [4599200, 4872200]}}, (200, 450), (1, 9000), [0.5, 0.5]),
    "PCSAFT1": ({"kind": "PCSAFT", "model": {"names": ["Methane"]}}, (150, 450), (1, 20000), [1.0]),
    "PCSAFT2": ({"kind": "PCSAFT", "model": {"names": ["Methane", "Ethane"]}}, (150, 450), (1, 20000), [0.5, 0.5]),
    "PCSAFT3k": ({"kind": "PCSAFT", "model": {"coeffs": [
        {"name": "Methane", "m": 1.0, "sigma_Angstrom": 3.7039, "epsilon_over_k": 150.03, "BibTeXKey": "Gross-IECR-2001"},
        {"name": "Ethane", "m": 1.6069, "sigma_Angstrom": 3.5206, "epsilon_over_k": 191.42, "BibTeXKey": "Gross-IECR-2001"},
        {"name": "Propane", "m": 2.0020, "sigma_Angstrom": 3.6184, "epsilon_over_k": 208.11, "BibTeXKey": "Gross-IECR-2001"}],
        "kmat": [[0, 0.01, 0.02], [0.01, 0, 0.03], [0.02, 0.03, 0]], "ab": "Liang-IECR-2014"}}, (200, 450), (1, 14000), [0.4, 0.3, 0.3]),
    "MIE1": ({"kind": "SAFT-VR-Mie", "model": {"names": ["Methane"]}}, (150, 450), (1, 20000), [1.0]),
    "MIE3": ({"kind": "SAFT-VR-Mie", "model": {"names": ["Methane", "Ethane", "Propane"], "kmat": [[0, 0.01, 0.02], [0.01, 0, 0.03], [0.02, 0.03, 0]]}},
             (200, 500), (1, 14000), [0.4, 0.3, 0.3]),
    "GSAFT_WATER": ({"kind": "genericSAFT", "model": {
        "nonpolar": {"kind": "SAFT-VR-Mie", "model": {"coeffs": [WORKLOADS["C4"]["spec"]["model"]["nonpolar"]["model"]["coeffs"][2]]}},
        "association": {"kind": "Dufal", "model": {"sigma / m": [3.0555e-10], "epsilon / J/mol": [3475.445374388054], "lambda_r": [35.823],
                                                   "epsilon_HB / J/mol": [13303.140189045183], "K_HB / m^3": [496.66e-30], "kmat": [[0.0]],
                                                   "Delta_rule": "Dufal", "molecule_sites": [["e", "e", "H", "H"]]}}}},
                    (290, 600), (10, 50000), [1.0]),
    "GSAFT_C4": (WORKLOADS["C4"]["spec"], WORKLOADS["C4"]["T"], WORKLOADS["C4"]["rho"], WORKLOADS["C4"]["z0"]),
    "GSAFT_CR1": ({"kind": "genericSAFT", "model": {
        "nonpolar": {"kind": "SAFT-VR-Mie", "model": {"names": ["Methane", "Ethane"]}},
        "association": {"kind": "canonical", "model": {"b / m^3/mol": [3.1e-5, 4.9e-5], "beta": [0.03, 0.008], "epsilon / J/mol": [12000.0, 18000.0],
                                                       "Delta_rule": "CR1", "molecule_sites": [["B", "P"], ["N", "N", "P"]],
                                                       "options": {"alpha": 0.4}}}}},
                  (250, 500), (10, 15000), [0.55, 0.45]),
    "GERG10": (WORKLOADS["C3"]["spec"], WORKLOADS["C3"]["T"], WORKLOADS["C3"]["rho"], WORKLOADS["C3"]["z0"]),
    "GERG21": (WORKLOADS["C5"]["spec"], WORKLOADS["C5"]["T"], WORKLOADS["C5"]["rho"], WORKLOADS["C5"]["z0"]),
    "GERG2004_3": ({"kind": "GERG2004resid", "model": {"names": ["nitrogen", "methane", "ethane"]}}, (150, 450), (1, 20000), [0.2, 0.7, 0.1]),
    "GERG_H2O_CO2": ({"kind": "GERG2008resid", "model": {"names": ["water", "carbondioxide", "hydrogen", "helium"]}}, (300, 700), (1, 30000), [0.3, 0.4, 0.2, 0.1]),
}


def points(name, n, k0=0):
    spec, Tr, rr, z0 = SPECS[name]
    T, rho, x = synth_points(n, k0, Tr[0], Tr[1], rr[0], rr[1], z0)
    return spec, T, rho, x


def full_table_oracle(m, T, rho, x):
    """[3][5][n] table of Lambda_nm from the reference's compile-time get_Ar{i}{j} (teqpcpp.hpp:22-37)."""
    tab = np.empty((3, 5, T.size))
    for i in range(3):
        for j in range(5):
            tab[i, j] = m.ArNM(i, j, T, rho, x)
    return tab
