#!/usr/bin/env python
"""bench.py -- headline benchmark: GERG-2008 10-component natural gas, Ar00..Ar04 (get_Ar04n), FP64, state points / s.

    python bench.py --gpus N --steps K --warmup W            our arm (one process per GPU; torchrun for N > 1)
    python bench.py --impl reference --gpus N ...            the reference's CPU path (oracle/_ref) on the host cores

A "step" is one pass of the hot path over one batch of synthetic state points (BASELINE.json configs[2], SURVEY.md section 8d:
10^8 points per GPU, independent points => shards, no collective, weak scaling).  `value` times the kernel with inputs resident
in HBM; `e2e` times the public host-buffer call (pinned host arrays in, results out) with the copies inside the timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "state points/sec (Ar0n n<=4, FP64)"
UNIT = "points/s"


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
    ap.add_argument("--points", type=int, default=0, help="state points per GPU per step (default: the workload's full size)")
    ap.add_argument("--e2e-points", type=int, default=1 << 25, help="points per GPU per step for the host-buffer (e2e) leg")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target duration of the CPU baseline sample")
    ap.add_argument("--no-cpu", action="store_true")
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


def cpu_reference_rate(wname, nmax, seconds, label):
    """Times the reference's own templates (oracle/_ref) on all host cores over a bounded sample of the workload."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_ref
    from teqp_b200.workloads import WORKLOADS, workload_points
    if not oracle_ref.available():
        return None
    w = WORKLOADS[wname]
    m = oracle_ref.RefModel(w["spec"])
    cores = m.nthreads
    T, rho, x = workload_points(wname, 4096 * cores)
    m.Ar0n(nmax, T, rho, x)  # warm-up, pages in the library
    t0 = time.perf_counter(); m.Ar0n(nmax, T, rho, x); dt = time.perf_counter() - t0
    rate = T.size / dt
    n = int(max(4096 * cores, min(rate * seconds, 5e7)))
    T, rho, x = workload_points(wname, n)
    t0 = time.perf_counter(); out = m.Ar0n(nmax, T, rho, x); dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": "%d points of %s (first points of the same synthetic stream), %s, %.1f s wall, %d host threads" % (n, wname, label, dt, cores),
            "_check": float(out[0, 0])}


def run_reference(args):
    from teqp_b200.workloads import WORKLOADS
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w = WORKLOADS[args.workload]
    nmax = w.get("nmax", 4)
    per_step = max(1.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    vals, last = [], None
    for s in range(args.warmup + args.steps):
        r = cpu_reference_rate(args.workload, nmax, per_step, "reference templates + AD shim, g++ -O2")
        if r is None:
            emit({"impl": "reference", "unavailable": "oracle/_ref/libteqp_ref_oracle.so is not built"})
            return
        if s >= args.warmup:
            vals.append(r["value"])
        last = r
    v = float(np.mean(vals))
    last.pop("_check", None)
    last["value"] = v
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": None, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "%s: GERG-2008 residual, %d components, get_Ar0%dn (SURVEY 8d synthetic grid; bounded sample per step, see cpu_baseline.sample)"
                                   % (args.workload, len(w["spec"]["model"].get("names", [])), nmax),
                       "components": w["spec"]["model"].get("names", w["spec"]["kind"])},
            "cpu_baseline": last, "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    emit(line)


def run_ours(args):
    import ctypes as C
    import torch
    import teqp_b200
    from teqp_b200.workloads import SEED, WORKLOADS
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; teqp_b200 has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    w = WORKLOADS[args.workload]
    nmax = w.get("nmax", 4)
    P = args.points or w["npts"]
    nc = len(w["z0"])
    model = teqp_b200.make_model(w["spec"])
    lib = teqp_b200.lib()
    dp = C.POINTER(C.c_double)

    def synth(n, k0):
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
        return T, rho, x

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident leg ----------------
    T, rho, x = synth(P, rank * P)
    out = torch.empty((nmax + 1, P), dtype=torch.float64, device=dev)
    for _ in range(args.warmup):
        model.get_Ar0n_batch(nmax, T, rho, x, out=out)
    barrier()
    sampler = ClockSampler(local); sampler.start()
    l0 = teqp_b200.kernel_launches()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record()
    for s in range(args.steps):
        model.get_Ar0n_batch(nmax, T, rho, x, out=out)   # inputs (96 B/pt x P) are far larger than the 126 MB L2: no flush needed
        ev[s + 1].record()
    barrier()
    launches = teqp_b200.kernel_launches() - l0
    clocks = sampler.stop()
    total_ms = ev[0].elapsed_time(ev[-1])
    kern_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    tmax = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    total_ms = float(tmax.item())
    value = world * P * args.steps / (total_ms * 1e-3)
    checksum = float(out[:, :: max(1, P // 1024)].sum().item())
    del out

    # ---------------- end-to-end leg: pinned host buffers through the public host-pointer call ----------------
    Pe = min(P, args.e2e_points)
    hT = teqp_b200.pinned_empty((Pe,)); hr = teqp_b200.pinned_empty((Pe,)); hx = teqp_b200.pinned_empty((nc, Pe)); ho = teqp_b200.pinned_empty((nmax + 1, Pe))
    hT[:] = T[:Pe].cpu().numpy(); hr[:] = rho[:Pe].cpu().numpy(); hx[:] = x[:, :Pe].cpu().numpy()
    del T, rho, x
    torch.cuda.empty_cache()
    for _ in range(max(1, min(args.warmup, 2))):
        model.get_Ar0n_batch(nmax, hT, hr, hx, out=ho)
    barrier()
    t0 = time.perf_counter()
    for s in range(args.steps):
        model.get_Ar0n_batch(nmax, hT, hr, hx, out=ho)   # returns when the results are in ho
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * Pe * args.steps / float(te.item())
    e2e_check = float(ho[:, :: max(1, Pe // 1024)].sum())

    if rank == 0:
        bpp = w["bytes_per_point"]
        kms = float(np.mean(kern_ms))
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        ach = bpp * P / (kms * 1e-3) / 1e9
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": {"workload": "%s: GERG-2008 residual, %d components, get_Ar0%dn over %d points per GPU (SURVEY 8d synthetic grid)" % (args.workload, nc, nmax, P),
                           "points_per_gpu": P, "l2": "inputs+outputs (%.1f GB per step) exceed L2; no flush" % (bpp * P / 1e9),
                           "e2e_points_per_gpu": Pe, "checksum": checksum, "e2e_checksum": e2e_check},
                "clocks": clocks, "gpu_launches": int(launches),
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 8 * (2 + nc) * Pe, "d2h_bytes_per_step": 8 * (nmax + 1) * Pe},
                "roofline": None}
        hbm_view = {"achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                    "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback 6650 (B200_PROFILING.md)",
                    "algorithmic_bytes_per_point": bpp}
        fp = fp64_info(lib, kms, P)
        if fp and "frac" in fp:
            # the kernel is bound by the FP64 FMA pipe (arithmetic intensity ~ 51 flop/B against a ridge of ~ 5 flop/B): the roofline
            # that bounds it is the measured DFMA peak; the HBM view is kept beside it
            line["roofline"] = {"bound": "fp64", "achieved": fp["achieved_tflops"], "peak": fp["peak_tflops_measured"], "unit": "TFLOP/s",
                                "frac": fp["frac"], "traffic": (fp.get("dram_bytes_per_point_ncu") or 0) * P or None,
                                "flops_per_point": fp["flops_per_point"], "fp64_pipe_active_pct_ncu": fp.get("fp64_pipe_active_pct_ncu"),
                                "peak_source": "teqpb200_fp64_peak(): 8 independent DFMA chains/thread, measured in this run",
                                "flops_source": fp.get("source"), "hbm": hbm_view}
        else:
            line["roofline"] = dict(hbm_view, bound="hbm", traffic=None)
        if not args.no_cpu and world == 1:
            cb = cpu_reference_rate(args.workload, nmax, args.cpu_seconds, "reference templates + AD shim, g++ -O2")
            if cb:
                cb.pop("_check", None)
                line["cpu_baseline"] = cb
        emit(line)
    if dist is not None:
        dist.destroy_process_group()


def fp64_info(lib, kernel_ms, P):
    """FP64-pipe view of the roofline: measured DFMA peak of this GPU and the kernel's executed flops per point (from the ncu
    instruction counts committed under profiles/, if present)."""
    import ctypes as C
    info = {}
    try:
        tf = C.c_double(0)
        err = C.create_string_buffer(300)
        if lib.teqpb200_fp64_peak(C.byref(tf), err, C.c_int(300)) == 0:
            info["peak_tflops_measured"] = tf.value
    except Exception:
        return None
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "flops_per_point.json")))
        fpp = prof["C3_Ar04n"]["flops_per_point"]
        info["flops_per_point"] = fpp
        info["achieved_tflops"] = fpp * P / (kernel_ms * 1e-3) / 1e12
        if "peak_tflops_measured" in info:
            info["frac"] = info["achieved_tflops"] / info["peak_tflops_measured"]
        info["source"] = prof["C3_Ar04n"].get("source", "profiles/")
        info["fp64_pipe_active_pct_ncu"] = prof["C3_Ar04n"].get("fp64_pipe_active_pct_ncu")
        info["dram_bytes_per_point_ncu"] = prof["C3_Ar04n"].get("dram_bytes_per_point")
    except Exception:
        pass
    return info


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
