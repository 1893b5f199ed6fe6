#!/usr/bin/env python
"""Parity adjudication table: libteqpb200 (GPU through the C ABI, or the CPU replay of the same kernel functions) and
the double-precision oracle, both against the binary128 evaluation of the reference's model templates.  Errors in units of the family
scale S (Taylor-coefficient space).  Writes profiles/<tag>_truth_adjudication_<backend>.json and prints a table.

    python tools/adjudicate.py [--backend emul|gpu] [--n 2000] [--tag r02]
"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_ref, parity_suite as ps  # noqa: E402

CONFIGS = [("C1", "PR1", ("ar",)), ("C2", "PCSAFT2", ("ar", "iso")), ("C3", "GERG10", ("ar",)), ("C4", "GSAFT_C4", ("ar", "atrhoxi")), ("C5", "GERG21", ("ar", "iso"))]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="emul"); ap.add_argument("--n", type=int, default=2000); ap.add_argument("--tag", default="r02")
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    if a.backend == "gpu":
        be = ps.GpuBackend()
    else:
        import emul_lib
        be = ps.EmulBackend(emul_lib)
    out = {}
    for cfg, name, what in CONFIGS:
        if a.only and cfg not in a.only.split(","):
            continue
        rows, bad = ps.truth_adjudication(be, oracle_ref, name, n=a.n, what=what)
        out[cfg] = {"case": name, "points": a.n, "rows": rows, "violations": [r["output"] for r in bad]}
        print("== %s (%s), backend %s: errors / family scale (max | p99.9)" % (cfg, name, a.backend))
        print("%-34s %21s %21s %21s %9s" % ("output", "|gpu-truth|", "|oracle-truth|", "|gpu-oracle|", "rel>1e-12"))
        for r in rows:
            print("%-34s %10.2e %10.2e %10.2e %10.2e %10.2e %10.2e %9.4f%s" % (r["output"], r["gpu_vs_truth_max"], r["gpu_vs_truth_p999"], r["oracle_vs_truth_max"],
                  r["oracle_vs_truth_p999"], r["gpu_vs_oracle_max"], r["gpu_vs_oracle_p999"], r["frac_points_rel_gt_1e-12"], "  <-- VIOLATION" if r in bad else ""))
        sys.stdout.flush()
    path = os.path.join(ROOT, "profiles", "%s_truth_adjudication_%s.json" % (a.tag, a.backend))
    json.dump(out, open(path, "w"), indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
