#!/usr/bin/env python
"""Secondary measurements: the BASELINE.json configs other than the headline (bench.py measures C3).  One JSON line per config:
device-resident throughput (CUDA events on torch's current stream, which the library launches on), algorithmic HBM GB/s, and the
reference's CPU path (oracle/_ref, all host threads) on a bounded sample of the same synthetic stream.

    python tools/bench_configs.py [--only C5] [--steps 5] [--warmup 3] [--no-cpu]
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--scale", type=float, default=1.0, help="multiply every config's point count (ncu runs use a small one)")
    args = ap.parse_args()
    import torch
    import teqp_b200
    from teqp_b200.workloads import SEED, WORKLOADS
    assert torch.cuda.is_available(), "no CUDA device; teqp_b200 has no CPU path"
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    lib = teqp_b200.lib()
    dp = C.POINTER(C.c_double)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = peaks.get("hbm_gbs", 6650.0)

    def synth(w, n):
        nc = len(w["z0"])
        T = torch.empty(n, dtype=torch.float64, device=dev); rho = torch.empty_like(T); x = torch.empty((nc, n), dtype=torch.float64, device=dev)
        err = C.create_string_buffer(500)
        z0 = (C.c_double * nc)(*w["z0"])
        lib.teqpb200_set_stream(C.c_void_p(torch.cuda.current_stream().cuda_stream))
        rc = lib.teqpb200_synth_points(C.c_longlong(n), C.c_longlong(0), C.c_ulonglong(SEED), C.c_double(w["T"][0]), C.c_double(w["T"][1]),
                                       C.c_double(w["rho"][0]), C.c_double(w["rho"][1]), z0, C.c_int(nc),
                                       C.cast(C.c_void_p(T.data_ptr()), dp), C.cast(C.c_void_p(rho.data_ptr()), dp), C.cast(C.c_void_p(x.data_ptr()), dp),
                                       err, C.c_int(500))
        assert rc == 0, err.value.decode()
        return T, rho, x

    def timed(fn):
        for _ in range(args.warmup):
            fn()
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
        l0 = teqp_b200.kernel_launches()
        ev[0].record()
        for s in range(args.steps):
            fn()
            ev[s + 1].record()
        torch.cuda.synchronize()
        return ev[0].elapsed_time(ev[-1]) / args.steps, (teqp_b200.kernel_launches() - l0) // args.steps

    def cpu(fn_of_n, n0, seconds=8.0):
        if args.no_cpu:
            return None
        import oracle_ref
        if not oracle_ref.available():
            return None
        t0 = time.perf_counter(); fn_of_n(n0); dt = time.perf_counter() - t0
        n = int(max(n0, min(n0 * seconds / max(dt, 1e-6), 2e7)))
        t0 = time.perf_counter(); fn_of_n(n); dt = time.perf_counter() - t0
        return {"value": n / dt, "unit": "points/s", "kind": "reference", "sample": "%d points, %.1f s wall" % (n, dt)}

    def emit(name, what, n, ms, launches, bpp, cb, extra=None):
        line = {"config": name, "what": what, "points": n, "ms_per_step": ms, "value": n / (ms * 1e-3), "unit": "points/s", "dtype": "f64",
                "gpu_launches_per_step": int(launches), "algorithmic_bytes_per_point": bpp,
                "hbm": {"achieved_GBs": bpp * n / (ms * 1e-3) / 1e9, "peak_GBs": hbm, "frac": bpp * n / (ms * 1e-3) / 1e9 / hbm},
                "cpu_baseline": cb}
        if extra:
            line.update(extra)
        print(json.dumps(line), flush=True)

    only = set(args.only.split(",")) if args.only else None

    def want(c):
        return only is None or c in only

    def refmodel(spec):
        import oracle_ref
        m = oracle_ref.RefModel(spec)
        return m, m.nthreads

    from teqp_b200.workloads import workload_points

    # ---- C1: Peng-Robinson pure methane, Ar00..Ar04: the reference-sized 10^6 grid and a 10^8 batch for the bandwidth figure
    if want("C1"):
        w = WORKLOADS["C1"]
        m = teqp_b200.make_model(w["spec"])
        for n in (int(1e6 * args.scale), int(1e8 * args.scale)):
            T, rho, x = synth(w, n)
            out = torch.empty((5, n), dtype=torch.float64, device=dev)
            ms, nl = timed(lambda: m.get_Ar0n_batch(4, T, rho, x, out=out))
            cb = None
            if n <= 2e6 and not args.no_cpu:
                rm, nt = refmodel(w["spec"])
                cb = cpu(lambda k: rm.Ar0n(4, *workload_points("C1", k)), 200000)
                if cb:
                    cb["cores"] = nt
            emit("C1", "PR pure methane get_Ar04n", n, ms, nl, 8 * 3 + 40, cb, {"note": "x[1][npts] is read too: 64 B/pt moved"})
            del T, rho, x, out
    # ---- C2: PC-SAFT methane+ethane: Ar_nm (n+m<=2) in one sweep + ln(phi)
    if want("C2"):
        w = WORKLOADS["C2"]
        m = teqp_b200.make_model(w["spec"])
        n = int(1e7 * args.scale)
        T, rho, x = synth(w, n)
        rv = (x * rho).contiguous()
        pairs = [(0, 0), (0, 1), (0, 2), (1, 0), (1, 1), (2, 0)]
        out = torch.empty((6, n), dtype=torch.float64, device=dev)

        def step():
            m.get_Arxy_batch(pairs, T, rho, x, out=out)
            m.get_ln_fugacity_coefficients_batch(T, rv)
        ms, nl = timed(step)
        cb = None
        if not args.no_cpu:
            rm, nt = refmodel(w["spec"])

            def ref(k):
                Tn, rn, xn = workload_points("C2", k)
                for (a, b) in pairs:
                    rm.ArNM(a, b, Tn, rn, xn)
                rm.lnphi_family(0, Tn, np.ascontiguousarray(xn * rn))
            cb = cpu(ref, 50000)
            if cb:
                cb["cores"] = nt
        emit("C2", "PC-SAFT CH4+C2H6: Ar00,01,02,10,11,20 + ln phi", n, ms, nl, 32 + 64, cb)
        del T, rho, x, rv, out
    # ---- C4: SAFT-VR-Mie ternary + Dufal association on water: the whole get_ATrhoXi table (36 entries x 3 components, of which
    # the reference exposes 23 x 3) in one launch
    if want("C4"):
        w = WORKLOADS["C4"]
        m = teqp_b200.make_model(w["spec"])
        n = int(1e7 * args.scale)
        T, rho, x = synth(w, n)
        out = torch.empty((3, 3, 3, 4, n), dtype=torch.float64, device=dev)
        ms, nl = timed(lambda: m.get_ATrhoXi_table_batch(T, rho, x, out=out))
        cb = None
        if not args.no_cpu:
            rm, nt = refmodel(w["spec"])
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            from parity_suite import COMBOS_ATRHOXI

            def ref(k):
                Tn, rn, xn = workload_points("C4", k)
                for i in range(3):
                    for (a, b, c) in COMBOS_ATRHOXI:
                        rm.ATrhoXi(a, b, i, c, Tn, rn, xn)
            cb = cpu(ref, 32 * nt, seconds=10.0)
            if cb:
                cb["cores"] = nt
        emit("C4", "SAFT-VR-Mie CH4+C2H6+H2O with Dufal association: get_ATrhoXi table (23 combos x 3 comps; 36 x 3 written)", n, ms, nl, 40 + 8 * 108, cb)
        del T, rho, x, out
        T, rho, x = synth(w, n)
        out = torch.empty((5, n), dtype=torch.float64, device=dev)
        ms, nl = timed(lambda: m.get_Ar0n_batch(4, T, rho, x, out=out))
        emit("C4-Ar04n", "same model: get_Ar04n", n, ms, nl, 40 + 40, None)
        del T, rho, x, out
    # ---- C5: GERG-2008 21 components: Psir, gradient, Hessian (441) and ln phi
    if want("C5"):
        w = WORKLOADS["C5"]
        m = teqp_b200.make_model(w["spec"])
        n = int(1e6 * args.scale)
        T, rho, x = synth(w, n)
        rv = (x * rho).contiguous()
        ms, nl = timed(lambda: m.build_Psir_fgradHessian_batch(T, rv))
        cb = None
        if not args.no_cpu:
            rm, nt = refmodel(w["spec"])

            def ref(k):
                Tn, rn, xn = workload_points("C5", k)
                rm.fgradHessian(Tn, np.ascontiguousarray(xn * rn))
            cb = cpu(ref, 64 * nt, seconds=10.0)
            if cb:
                cb["cores"] = nt
        emit("C5", "GERG-2008 21 comps: build_Psir_fgradHessian (f, grad[21], H[441])", n, ms, nl, 8 * 22 + 8 * (1 + 21 + 441), cb)
        ms2, nl2 = timed(lambda: m.get_ln_fugacity_coefficients_batch(T, rv))
        emit("C5-lnphi", "GERG-2008 21 comps: ln phi[21]", n, ms2, nl2, 8 * 22 + 8 * 21, None)
        del T, rho, x, rv


if __name__ == "__main__":
    main()
