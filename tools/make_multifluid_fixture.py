#!/usr/bin/env python
"""Build-container tool: writes tests/golden/multifluid_specs.json -- self-contained `{"kind": "multifluid", ...}` model specifications
(components, BIP and departure collections INLINE) for the parity tests of the JSON-driven multifluid kind, so that the same spec can be
given to libteqpb200 and to the oracle on the GPU box, where /root/reference does not exist.

Only what the factory reads is kept from each fluid file of the reference (/root/reference/teqp/fluiddata/dev/fluids/<name>.json):
EOS[0].alphar, EOS[0].gas_constant, EOS[0].STATES.reducing, INFO identifiers; from dev/mixtures only the entries of the pairs used.
These are published equation-of-state coefficients (test INPUT fixtures, a few kB); the fluid library itself is not copied."""
import json, os
ROOT = "/root/reference/teqp/fluiddata"
HERE = os.path.dirname(os.path.abspath(__file__))


def fluid(name):
    j = json.load(open(os.path.join(ROOT, "dev", "fluids", name + ".json")))
    e = j["EOS"][0]
    return {"EOS": [{"alphar": e["alphar"], "gas_constant": e["gas_constant"], "STATES": {"reducing": {"T": e["STATES"]["reducing"]["T"], "rhomolar": e["STATES"]["reducing"]["rhomolar"]}}}],
            "INFO": {k: j["INFO"][k] for k in ("NAME", "CAS", "REFPROP_NAME") if k in j["INFO"]}}


BIP = json.load(open(os.path.join(ROOT, "dev", "mixtures", "mixture_binary_pairs.json")))
DEP = json.load(open(os.path.join(ROOT, "dev", "mixtures", "mixture_departure_functions.json")))


def spec(names, flags=None):
    comps = [fluid(n) for n in names]
    cas = {c["INFO"]["CAS"] for c in comps}
    bips = [b for b in BIP if b["CAS1"] in cas and b["CAS2"] in cas]
    fns = {b.get("function", "") for b in bips} - {""}
    deps = [d for d in DEP if d["Name"] in fns]
    m = {"components": comps}
    if len(comps) > 1:
        m["BIP"] = bips
    if deps:                      # (the reference rejects a "departure" key that holds an empty array, multifluid.hpp:1112-1150)
        m["departure"] = deps
    if flags:
        m["flags"] = flags
    return {"kind": "multifluid", "model": m}


out = {
    # Power + Gaussian pure terms, GERG-2008 departure functions, mole-fraction-weighted R
    "MF_CH4_N2_O2": spec(["Methane", "Nitrogen", "Oxygen"]),
    # a binary with a departure function, constant (CODATA) gas constant
    "MF_CH4_C2H6_CODATA": spec(["Methane", "Ethane"], {"Rmodel": "CODATA"}),
    # single component: the short circuit that ignores the mole fraction (multifluid.hpp:172-175)
    "MF_N2": spec(["Nitrogen"]),
    # Lemmon2005 (R125) + Exponential-type departure (R32-R125)
    "MF_R32_R125": spec(["R32", "R125"]),
    # non-analytic critical-region terms (CO2 of Span-Wagner) with nitrogen and a GERG-2008 departure function: jet sweeps only
    "MF_CO2_N2": spec(["CarbonDioxide", "Nitrogen"]),
    # GaoB terms (ammonia 2020) with argon; Lorentz-Berthelot estimate for the missing pair
    "MF_NH3_AR": spec(["Ammonia", "Argon"], {"estimate": "Lorentz-Berthelot"}),
}
path = os.path.join(HERE, "..", "tests", "golden", "multifluid_specs.json")
json.dump(out, open(path, "w"), separators=(",", ":"))
print("wrote", path, os.path.getsize(path), "bytes")
for k, v in out.items():
    print(k, [c["INFO"]["NAME"] for c in v["model"]["components"]], "BIP", len(v["model"].get("BIP", [])), "dep", [d["Name"] for d in v["model"].get("departure", [])],
          [t["type"] for c in v["model"]["components"] for t in c["EOS"][0]["alphar"]])
