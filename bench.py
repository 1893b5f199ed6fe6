#!/usr/bin/env python
"""bench.py -- headline benchmark: GERG-2008 10-component natural gas, Ar00..Ar04 (get_Ar04n), FP64, state points / s.

    python bench.py --gpus N --steps K --warmup W            our arm (one process per GPU; torchrun for N > 1)
    python bench.py --impl reference --gpus N ...            the reference's CPU path (oracle/_ref) on the host cores
    python bench.py --workload C1|C2|C4|C5 ...               the other BASELINE.json configs, same JSON line

A "step" is one pass of the hot path over the workload's synthetic state points (BASELINE.json configs[2], SURVEY.md section 8d:
10^8 points IN TOTAL, split evenly over the N GPUs -- strong scaling; independent points => shards, no collective).
  value    device-resident leg: every rank evaluates its contiguous slice (inputs already in HBM), CUDA events, max over ranks.
  e2e      the library's own multi-GPU host driver: ONE process (rank 0) calls teqpb200_set_devices([0..N-1]) and the public
           host-pointer entry point on the WHOLE workload in pinned host memory; H2D and D2H copies are inside the timed region.
  roofline the kernel's executed FP64 flops per point (counted from the SASS of the cubin that is actually loaded) / launch time.
"""
import argparse
import collections
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "state points/sec (Ar0n n<=4, FP64)"
UNIT = "points/s"
REF_LABEL = "reference templates + clean-room AD shim (oracle/_ref), g++ -O3 -fno-tree-vectorize -march=x86-64-v3"


# Only the ONE JSON line may reach stdout: libraries that print to fd 1 (NCCL's version banner, for instance) are diverted to stderr.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C3")
    ap.add_argument("--points", type=int, default=0, help="state points in total per step (default: the workload's full size)")
    ap.add_argument("--e2e-points", type=int, default=0, help="points in total per step for the host-buffer (e2e) leg (default: the same)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target duration of the CPU baseline sample")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the short C1/C2/C4/C5 lines the default N=1 run appends")
    return ap.parse_args()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons for one GPU while the timed region runs (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [s.strip() for s in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------------------------------------------------------
# workloads: what one step calls, for device-resident tensors and for host arrays (same public methods of teqp_b200.Model)
# ----------------------------------------------------------------------------------------------------------------------------
ARXY_PAIRS = [(0, 0), (0, 1), (0, 2), (1, 0), (1, 1), (2, 0)]


def describe(wname, w, nc, total):
    what = w["what"]
    if what == "Ar0n":
        call = "get_Ar0%dn" % w.get("nmax", 4)
    elif what == "Arxy+lnphi":
        call = "get_Arxy_batch[Ar00,01,02,10,11,20] + get_ln_fugacity_coefficients_batch"
    elif what == "ATrhoXi":
        call = "get_ATrhoXi_table_batch (23 combinations x %d components; 36 x %d rows written)" % (nc, nc)
    elif what == "fgradHessian+dlnphi":
        call = "build_Psir_fgradHessian_batch + isochoric_properties_batch[dlnphi_dT, dlnphi_drho, dlnphi_dx[%d]]" % (nc * nc)
    else:
        call = "build_Psir_fgradHessian_batch (Psir, grad[%d], Hessian[%d])" % (nc, nc * nc)
    fam = w["spec"]["kind"]
    return "%s: %s, %d components, %s over %d points in total (SURVEY 8d synthetic grid)" % (wname, fam, nc, call, total)


def make_steps(w, model, nc, n, xp):
    """returns (alloc_outputs(n) -> outs, step(inputs, outs), rows_out) for array module xp in {'torch', 'numpy'}"""
    what = w["what"]
    nmax = w.get("nmax", 4)
    if what == "Ar0n":
        return (lambda mk: mk((nmax + 1, n))), (lambda I, O: model.get_Ar0n_batch(nmax, I["T"], I["rho"], I["x"], out=O)), nmax + 1
    if what == "Arxy+lnphi":
        def alloc(mk):
            return (mk((len(ARXY_PAIRS), n)), mk((nc, n)))

        def step(I, O):
            model.get_Arxy_batch(ARXY_PAIRS, I["T"], I["rho"], I["x"], out=O[0])
            model.get_ln_fugacity_coefficients_batch(I["T"], I["rv"], out=O[1])
        return alloc, step, len(ARXY_PAIRS) + nc
    if what == "ATrhoXi":
        return (lambda mk: mk((nc, 3, 3, 4, n))), (lambda I, O: model.get_ATrhoXi_table_batch(I["T"], I["rho"], I["x"], out=O)), 36 * nc
    if what == "fgradHessian":
        return (lambda mk: (mk((1, n)), mk((nc, n)), mk((nc * nc, n)))), (lambda I, O: model.build_Psir_fgradHessian_batch(I["T"], I["rv"], out=O)), 1 + nc + nc * nc
    if what == "fgradHessian+dlnphi":
        # C5 as BASELINE.json words it: "isochoric Hessian/fugacity composition derivatives": Psir, gradient, Hessian AND d ln phi / dx (+ d/dT, d/drho)
        which = ["dlnphi_dT", "dlnphi_drho", "dlnphi_dx"]

        def step(I, O):
            if os.environ.get("TEQPB200_C5X_TWO_CALLS"):   # A/B: the two separate entry points
                model.build_Psir_fgradHessian_batch(I["T"], I["rv"], out=O[:3])
                model.isochoric_properties_batch(I["T"], I["rv"], which=which, out=dict(zip(which, O[3:])))
            else:                                          # one call, one launch for the GERG models
                model.fgradHessian_and_properties_batch(I["T"], I["rv"], which=which, out=O[:3], out_props=dict(zip(which, O[3:])))
        return (lambda mk: (mk((1, n)), mk((nc, n)), mk((nc * nc, n)), mk((nc, n)), mk((nc, n)), mk((nc * nc, n)))), step, 1 + 3 * nc + 2 * nc * nc
    raise SystemExit("unknown workload kind " + what)


def needs_rv(w):
    return w["what"] in ("Arxy+lnphi", "fgradHessian", "fgradHessian+dlnphi")


def rows_in(w, nc):
    if w.get("x_broadcast"):
        return 2
    return {"Ar0n": 2 + nc, "ATrhoXi": 2 + nc, "Arxy+lnphi": 2 + 2 * nc, "fgradHessian": 1 + nc, "fgradHessian+dlnphi": 1 + nc}[w["what"]]


def algorithmic_bytes(w, nc):
    """inputs + requested outputs, FP64 SoA.  C1 is quoted with x implicit (56 B/pt); the kernel reads its x row (64)."""
    return w["bytes_per_point"]


def cpu_reference_rate(wname, seconds):
    """Times the reference's own templates (oracle/_ref) on all host cores over a bounded sample of the workload."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_ref
    from teqp_b200.workloads import WORKLOADS, workload_points
    if not oracle_ref.available():
        return None
    w = WORKLOADS[wname]
    m = oracle_ref.RefModel(w["spec"])
    cores = m.nthreads
    what = w["what"]
    nmax = w.get("nmax", 4)

    def call(n):
        T, rho, x = workload_points(wname, n)
        if what == "Ar0n":
            return m.Ar0n(nmax, T, rho, x)
        if what == "Arxy+lnphi":
            for (a, b) in ARXY_PAIRS:
                m.ArNM(a, b, T, rho, x)
            return m.lnphi_family(0, T, np.ascontiguousarray(x * rho))
        if what == "ATrhoXi":
            from parity_suite import COMBOS_ATRHOXI
            for i in range(x.shape[0]):
                for (a, b, c) in COMBOS_ATRHOXI:
                    out = m.ATrhoXi(a, b, i, c, T, rho, x)
            return out
        if what == "fgradHessian+dlnphi":
            rv = np.ascontiguousarray(x * rho)
            m.fgradHessian(T, rv)
            for q in (1, 2, 3):
                out = m.lnphi_family(q, T, rv)
            return out
        return m.fgradHessian(T, np.ascontiguousarray(x * rho))[0]

    n0 = {"Ar0n": 4096, "Arxy+lnphi": 2048, "ATrhoXi": 16, "fgradHessian": 32, "fgradHessian+dlnphi": 16}[what] * cores
    call(n0)  # warm-up, pages in the library
    t0 = time.perf_counter(); call(n0); dt = time.perf_counter() - t0
    rate = n0 / dt
    n = int(max(n0, min(rate * seconds, 5e7)))
    t0 = time.perf_counter(); call(n); dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": "%d points of %s (first points of the same synthetic stream), %s, %.1f s wall, %d host threads" % (n, wname, REF_LABEL, dt, cores)}


def config_block(wname, w, nc, total):
    return {"workload": describe(wname, w, nc, total), "points_total": total}


def run_reference(args):
    from teqp_b200.workloads import WORKLOADS
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w = WORKLOADS[args.workload]
    nc = len(w["z0"])
    total = args.points or w.get("npts_strong" if args.gpus > 1 else "npts", w["npts"])
    per_step = max(1.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    vals, last = [], None
    for s in range(args.warmup + args.steps):
        r = cpu_reference_rate(args.workload, per_step)
        if r is None:
            emit({"impl": "reference", "unavailable": "oracle/_ref/libteqp_ref_oracle.so is not built"})
            return
        if s >= args.warmup:
            vals.append(r["value"])
        last = r
    v = float(np.mean(vals))
    last["value"] = v
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": None, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_block(args.workload, w, nc, total),
            "cpu_baseline": last, "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    emit(line)


# ----------------------------------------------------------------------------------------------------------------------------
# executed FP64 flops per point of the kernel that is actually loaded: static SASS count of the straight-line generated kernel
# ----------------------------------------------------------------------------------------------------------------------------
def generated_kernel_flops(model, NT, ND):
    """GERG models run a per-model generated kernel without loops: every FP64 instruction of the cubin executes once per state point
    (the only branches are the npts guard and the rho = 0 / x_i = 0 special cases).  Counts DFMA/DMUL/DADD from `cuobjdump -sass`."""
    import ctypes as C
    import teqp_b200
    lib = teqp_b200.lib()
    n = C.c_ulonglong(0)
    err = C.create_string_buffer(2000)
    if lib.teqpb200_generated_kernel(model._uuid, NT, ND, 2, None, C.byref(n), err, 2000) != 0 or n.value == 0:
        return None
    buf = C.create_string_buffer(n.value)
    if lib.teqpb200_generated_kernel(model._uuid, NT, ND, 2, buf, C.byref(n), err, 2000) != 0:
        return None
    with tempfile.NamedTemporaryFile(suffix=".cubin") as f:
        f.write(buf.raw[:n.value]); f.flush()
        try:
            sass = subprocess.run(["cuobjdump", "-sass", f.name], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=120).stdout
        except Exception:
            return None
    ops = collections.Counter()
    for ln in sass.splitlines():
        mm = re.match(r"\s+/\*[0-9a-f]+\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if mm:
            ops[mm.group(2).split(".")[0]] += 1
    if not ops:
        return None
    tot = sum(ops.values())
    fp64 = ops["DFMA"] + ops["DMUL"] + ops["DADD"] + ops["DSETP"] + ops["DMNMX"]
    return {"flops_per_point": 2 * ops["DFMA"] + ops["DMUL"] + ops["DADD"], "dfma": ops["DFMA"], "dmul": ops["DMUL"], "dadd": ops["DADD"],
            "sass_instructions": tot, "fp64_instructions": fp64, "fp64_share_of_instructions": fp64 / tot,
            "source": "static SASS count of the loaded per-model cubin (cuobjdump -sass; straight-line code, every instruction runs once per point)"}


def profile_flops(wname):
    """ncu-measured executed flops per point committed under profiles/ (cross-check for the static count; the only source for the
    template kernels of the other families)."""
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "flops_per_point.json")))
        return prof.get(wname)
    except Exception:
        return None


def run_ours(args):
    import ctypes as C
    import torch
    import teqp_b200
    from teqp_b200.workloads import SEED, WORKLOADS, shard_range
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; teqp_b200 has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    lib = teqp_b200.lib()
    lib.teqpb200_last_call_device_ms.restype = C.c_double
    dp = C.POINTER(C.c_double)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def synth(w, n, k0):
        nc = len(w["z0"])
        T = torch.empty(n, dtype=torch.float64, device=dev); rho = torch.empty_like(T); x = torch.empty((nc, n), dtype=torch.float64, device=dev)
        err = C.create_string_buffer(500)
        z0 = (C.c_double * nc)(*w["z0"])
        lib.teqpb200_set_stream(C.c_void_p(torch.cuda.current_stream().cuda_stream))
        rc = lib.teqpb200_synth_points(C.c_longlong(n), C.c_longlong(k0), C.c_ulonglong(SEED), C.c_double(w["T"][0]), C.c_double(w["T"][1]),
                                       C.c_double(w["rho"][0]), C.c_double(w["rho"][1]), z0, C.c_int(nc),
                                       C.cast(C.c_void_p(T.data_ptr()), dp), C.cast(C.c_void_p(rho.data_ptr()), dp), C.cast(C.c_void_p(x.data_ptr()), dp),
                                       err, C.c_int(500))
        if rc:
            raise RuntimeError(err.value.decode())
        I = {"T": T, "rho": rho, "x": x}
        if w.get("x_broadcast"):
            # one composition for every point (pure fluid: x = [1]): the library's TEQPB200_BROADCAST_X path, no x column in HBM (56 B/pt)
            I["x"] = torch.tensor(w["z0"], dtype=torch.float64, device=dev)
        if needs_rv(w):
            I["rv"] = (x * rho).contiguous()
        return I

    def device_leg(wname, total, steps, warmup, sample_clocks):
        """every rank evaluates its contiguous slice of [0, total); returns dict with value, times, launches, clocks, checksum"""
        w = WORKLOADS[wname]
        nc = len(w["z0"])
        a, b = shard_range(total, rank, world)
        P = b - a
        model = teqp_b200.make_model(w["spec"])
        I = synth(w, P, a)
        alloc, step, rows_out = make_steps(w, model, nc, P, "torch")
        O = alloc(lambda shape: torch.empty(shape, dtype=torch.float64, device=dev))
        # clocks / throttle reasons are sampled from the warm-up steps on (same kernel, same load): the timed region of the headline is ~0.1 s,
        # one nvidia-smi sample at its 100 ms period
        sampler = ClockSampler(local) if sample_clocks else None
        if sampler:
            sampler.start()
        for _ in range(warmup):
            step(I, O)
        barrier()
        l0 = teqp_b200.kernel_launches()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        ev[0].record()
        for s in range(steps):
            step(I, O)    # inputs + outputs of one step are far larger than the 126 MB L2 (stated in config.l2): no flush needed
            ev[s + 1].record()
        barrier()
        launches = teqp_b200.kernel_launches() - l0
        clocks = sampler.stop() if sampler else None
        total_ms = ev[0].elapsed_time(ev[-1])
        kern_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
        tmax = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        total_ms = float(tmax.item())
        first = O[0] if isinstance(O, tuple) else O
        first = first.reshape(-1, P)
        checksum = float(first[:, :: max(1, P // 1024)].sum().item())
        res = {"model": model, "w": w, "nc": nc, "P": P, "a": a, "total": total, "value": total * steps / (total_ms * 1e-3), "total_ms": total_ms,
               "kern_ms": float(np.mean(kern_ms)), "launches": int(launches), "clocks": clocks, "checksum": checksum, "rows_out": rows_out, "I": I}
        del O
        return res

    # ---------------- device-resident leg (strong scaling over the ranks) ----------------
    wname = args.workload
    w = WORKLOADS[wname]
    total = args.points or w.get("npts_strong" if world > 1 else "npts", w["npts"])
    R = device_leg(wname, total, args.steps, args.warmup, True)
    model, nc, P = R["model"], R["nc"], R["P"]
    I = R.pop("I")

    # ---------------- end-to-end leg: ONE process drives all N devices through the library's host-pointer call ----------------
    e2e = None
    Pe = args.e2e_points or total
    if not args.no_e2e:
        if rank == 0:
            hin = {"T": teqp_b200.pinned_empty((Pe,))}
            if not w["what"].startswith("fgradHessian"):
                hin["rho"] = teqp_b200.pinned_empty((Pe,))
                if not w.get("x_broadcast"):
                    hin["x"] = teqp_b200.pinned_empty((nc, Pe))
            if needs_rv(w):
                hin["rv"] = teqp_b200.pinned_empty((nc, Pe))
            # fill the pinned buffers with the SAME synthetic stream, generated on the device in slices and copied down
            sl = 1 << 24
            for c0 in range(0, Pe, sl):
                n = min(sl, Pe - c0)
                J = synth(w, n, c0)
                for k in hin:
                    if k == "x" and w.get("x_broadcast"):
                        continue
                    torch.from_numpy(hin[k])[..., c0:c0 + n].copy_(J[k])
                del J
            if w.get("x_broadcast"):
                hin["x"] = np.array(w["z0"], dtype=np.float64)
            alloc_h, step_h, _ = make_steps(w, model, nc, Pe, "numpy")
            HO = alloc_h(lambda shape: teqp_b200.pinned_empty(shape))
        del I
        torch.cuda.empty_cache()
        barrier()
        if rank == 0:
            teqp_b200.set_devices(list(range(world)))       # the north-star "batched host driver that shards the batch across the GPUs"
            for _ in range(max(1, min(args.warmup, 2))):
                step_h(hin, HO)
            l0 = teqp_b200.kernel_launches()
            t0 = time.perf_counter()
            for s in range(args.steps):
                step_h(hin, HO)                              # returns when the results are in the host buffers
            e2e_s = time.perf_counter() - t0
            e2e_launches = teqp_b200.kernel_launches() - l0
            dev_ms = [lib.teqpb200_last_call_device_ms(C.c_int(d)) for d in range(world)]
            first = HO[0] if isinstance(HO, tuple) else HO
            first = first.reshape(-1, Pe)
            e2e = {"value": Pe * args.steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": 8 * rows_in(w, nc) * Pe, "d2h_bytes_per_step": 8 * R["rows_out"] * Pe,
                   "points_per_step": Pe, "driver": "one process, teqpb200_set_devices(%s), host-pointer call, one host thread + 3 streams per device" % list(range(world)),
                   "device_ms_last_call_per_gpu": [round(v, 3) for v in dev_ms], "gpu_launches": int(e2e_launches),
                   "checksum": float(first[:, :: max(1, Pe // 1024)].sum())}
            teqp_b200.set_devices([])
            lib.teqpb200_release_scratch()
        barrier()
    else:
        del I

    if rank == 0:
        bpp = algorithmic_bytes(w, nc)
        kms = R["kern_ms"]
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        ach = bpp * P / (kms * 1e-3) / 1e9
        line = {"metric": METRIC, "value": R["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": R["total_ms"] / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": dict(config_block(wname, w, nc, total), points_per_gpu=P,
                               l2="inputs+outputs (%.2f GB per GPU per step) exceed the 126 MB L2; no flush" % ((8 * (rows_in(w, nc) + R["rows_out"])) * P / 1e9),
                               checksum=R["checksum"]),
                "clocks": R["clocks"], "gpu_launches": R["launches"] + (e2e["gpu_launches"] if e2e else 0),
                "e2e": e2e, "roofline": None}
        hbm_view = {"achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                    "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback 6650 (B200_PROFILING.md)", "algorithmic_bytes_per_point": bpp}
        fl = None
        if w["what"] == "Ar0n" and w["spec"]["kind"].startswith("GERG"):
            fl = generated_kernel_flops(model, 0, w.get("nmax", 4))
        pf = profile_flops(wname)
        if fl is None and pf:
            fl = {"flops_per_point": pf["flops_per_point"], "source": pf.get("source")}
        tf = C.c_double(0)
        err = C.create_string_buffer(300)
        peak_meas = tf.value if lib.teqpb200_fp64_peak(C.byref(tf), err, C.c_int(300)) == 0 else 0.0
        sm_mhz = (R["clocks"] or {}).get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
        peak_theory = 148 * 64 * 2 * sm_mhz * 1e6 / 1e12
        if fl and peak_meas > 0 and w.get("bound", "fp64") == "fp64":
            # bound by the FP64 FMA pipe (arithmetic intensity far above the ~5 flop/B ridge): the roofline is the measured DFMA peak
            achieved = fl["flops_per_point"] * P / (kms * 1e-3) / 1e12
            line["roofline"] = {"bound": "fp64", "achieved": achieved, "peak": peak_meas, "unit": "TFLOP/s", "frac": achieved / peak_meas,
                                "traffic": (pf or {}).get("dram_bytes_per_point") and (pf["dram_bytes_per_point"] * P),
                                "kernel_ms": kms, "points_per_launch": P, "flops": fl,
                                "peak_source": "teqpb200_fp64_peak(): 8 independent DFMA chains/thread, measured in this run on this GPU",
                                "peak_theoretical": peak_theory, "frac_of_theoretical": achieved / peak_theory,
                                "peak_theoretical_source": "148 SMs x 64 FP64 lanes x 2 x %.0f MHz" % sm_mhz,
                                "ncu_profile": pf, "hbm": hbm_view}
        else:
            line["roofline"] = dict(hbm_view, bound="hbm", traffic=(pf or {}).get("dram_bytes_per_point") and (pf["dram_bytes_per_point"] * P),
                                    kernel_ms=kms, points_per_launch=P, fp64_peak_measured=peak_meas)
        if not args.no_cpu and world == 1:
            cb = cpu_reference_rate(wname, args.cpu_seconds)
            if cb:
                line["cpu_baseline"] = cb
        if world == 1 and wname == "C3" and not args.no_secondary:
            # the other BASELINE.json configs, device-resident, short (3 warm-up + 3 timed steps each): recorded by the driver with the headline
            sec = []
            for name, npts2 in (("C1", 0), ("C1", 10**8), ("C2", 0), ("C4", 0), ("C5", 0), ("C5x", 0)):
                try:
                    ww = WORKLOADS[name]
                    S = device_leg(name, npts2 or ww["npts"], 3, 3, False)
                    S.pop("I"); S.pop("model")
                    b2 = algorithmic_bytes(ww, S["nc"])
                    sec.append({"config": describe(name, ww, S["nc"], S["total"]), "value": S["value"], "unit": UNIT, "ms_per_step": S["total_ms"] / 3,
                                "gpu_launches": S["launches"], "algorithmic_bytes_per_point": b2,
                                "hbm_frac": b2 * S["P"] / (S["kern_ms"] * 1e-3) / 1e9 / hbm_peak, "checksum": S["checksum"]})
                    torch.cuda.empty_cache()
                except Exception as e:  # a secondary line must never take the headline down
                    sec.append({"config": name, "error": str(e)[:300]})
            # the batched pure-fluid saturation solver (teqp::pure_VLE_T, 10 Newton iterations per isotherm) on the C1 model
            try:
                from teqp_b200.workloads import pr_saturation_guess
                nv = 10**6
                Tg = np.linspace(105.0, 185.0, 2048)
                rLg, rVg = pr_saturation_guess(Tg, 190.564, 4599200.0, 0.011)
                Tv = np.linspace(105.0, 185.0, nv)
                dTv = torch.from_numpy(Tv).cuda(); dLv = torch.from_numpy(np.interp(Tv, Tg, rLg)).cuda(); dVv = torch.from_numpy(np.exp(np.interp(Tv, Tg, np.log(rVg)))).cuda()
                dzv = torch.ones(1, dtype=torch.float64, device="cuda")
                mv = teqp_b200.make_model(WORKLOADS["C1"]["spec"])
                l0 = teqp_b200.lib().teqpb200_kernel_launches()
                for _ in range(3):
                    ov = mv.pure_VLE_T_batch(dTv, dLv, dVv, 10, dzv)
                torch.cuda.synchronize()
                lv = (teqp_b200.lib().teqpb200_kernel_launches() - l0) // 3
                ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                ev0.record()
                for _ in range(3):
                    ov = mv.pure_VLE_T_batch(dTv, dLv, dVv, 10, dzv)
                ev1.record(); torch.cuda.synchronize()
                msv = ev0.elapsed_time(ev1) / 3
                okv = bool(torch.isfinite(ov[0]).all() and (ov[0] > 1.2 * ov[1]).all())
                sec.append({"config": "V1: PR methane, pure_VLE_T_batch (teqp::pure_VLE_T + dpsatdT_pure, 10 Newton iterations) over %d isotherms 105..185 K" % nv,
                            "value": nv / msv * 1e3, "unit": "isotherms/s", "ms_per_step": msv, "gpu_launches": int(lv), "all_converged_two_phase": okv,
                            "checksum": float(ov[2].sum().item())})
            except Exception as e:
                sec.append({"config": "V1", "error": str(e)[:300]})
            line["secondary"] = sec
        emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
