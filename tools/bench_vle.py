#!/usr/bin/env python
"""GPU timing of the batched pure-fluid saturation solver (pure_VLE_T_batch) beside the reference's own teqp::pure_VLE_T on the host
cores: N isotherms of methane, 10 Newton iterations each, device-resident inputs, CUDA events on the launching stream.  One JSON line per model."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import teqp_b200
import cases
import oracle_ref
import parity_suite as ps

N = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
MAXITER = 10
for name in ("PR1", "GERG10", "PCSAFT1"):
    z, Tc, pc, w, (Tlo, Thi) = ps.VLE_CASES[name]
    spec = cases.SPECS[name][0]
    T = np.linspace(Tlo, Thi, N)
    nb = 4096   # the starting values come from a numpy cubic solve per point: interpolate them from a coarse grid
    Tg = np.linspace(Tlo, Thi, nb)
    rLg, rVg = ps.pr_saturation_guess(Tg, Tc, pc, w)
    rL0 = np.interp(T, Tg, rLg); rV0 = np.exp(np.interp(T, Tg, np.log(rVg)))
    m = teqp_b200.make_model(spec)
    dT, dL, dV = (torch.from_numpy(a).cuda() for a in (T, rL0, rV0))
    dz = torch.tensor(z, dtype=torch.float64, device="cuda")
    l0 = teqp_b200.lib().teqpb200_kernel_launches()
    for _ in range(2):
        out = m.pure_VLE_T_batch(dT, dL, dV, MAXITER, dz)
    torch.cuda.synchronize()
    l1 = teqp_b200.lib().teqpb200_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 5
    e0.record()
    for _ in range(steps):
        out = m.pure_VLE_T_batch(dT, dL, dV, MAXITER, dz)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    gL, gV = out[0].cpu().numpy(), out[1].cpu().numpy()
    # CPU reference on a bounded sample of the same isotherms
    ns = 20000
    idx = np.linspace(0, N - 1, ns).astype(np.int64)
    ref = oracle_ref.RefModel(spec)
    t0 = time.perf_counter()
    rL, rV, _, _ = ref.pure_VLE_T(T[idx], rL0[idx], rV0[idx], MAXITER, z)
    dt = time.perf_counter() - t0
    ok = np.isfinite(rL) & (rL > 1.2 * rV)
    err = float(np.max(np.abs(gL[idx][ok] / rL[ok] - 1)))
    print(json.dumps({"metric": "isotherms/s (pure_VLE_T, %d Newton iterations)" % MAXITER, "model": name, "value": N / ms * 1e3, "ms_per_step": ms, "isotherms": N,
                      "gpu_launches_per_step": (l1 - l0) // 2, "cpu_reference": {"value": ns / dt, "cores": ref.nthreads, "sample": "%d of the same isotherms" % ns},
                      "max_rel_diff_rhoL_vs_reference_on_sample": err, "converged_fraction_reference": float(ok.mean())}))

# binary T-x flash (mix_VLE_Tx_batch): methane + ethane, Peng-Robinson and PC-SAFT
for name in ("PR2k", "PCSAFT2"):
    spec = cases.SPECS[name][0]
    NB = 2048
    rng = np.random.RandomState(7)
    Tb = 150.0 + 30.0 * rng.rand(NB); xb = 0.05 + 0.4 * rng.rand(NB)
    rvLb, rvVb = ps.pr_mixture_guess(Tb, xb, [190.564, 305.322], [4599200.0, 4872200.0], [0.011, 0.099], ps.MIXVLE_CASES[name])
    reps = max(1, N // NB)
    T = np.tile(Tb, reps); rvL0 = np.ascontiguousarray(np.tile(rvLb, (1, reps))); rvV0 = np.ascontiguousarray(np.tile(rvVb, (1, reps)))
    xs = np.ascontiguousarray(np.tile(np.stack([xb, 1 - xb]), (1, reps)))
    n = T.size
    m = teqp_b200.make_model(spec)
    dT, dL, dV, dx = (torch.from_numpy(a).cuda() for a in (T, rvL0, rvV0, xs))
    for _ in range(2):
        out = m.mix_VLE_Tx_batch(dT, dL, dV, dx, maxiter=MAXITER)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 3
    e0.record()
    for _ in range(steps):
        out = m.mix_VLE_Tx_batch(dT, dL, dV, dx, maxiter=MAXITER)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    gc = out[2].cpu().numpy()
    ref = oracle_ref.RefModel(spec)
    ns = 4096
    t0 = time.perf_counter()
    rL, rV, rc = ref.mix_VLE_Tx(T[:ns], rvL0[:, :ns].copy(), rvV0[:, :ns].copy(), xs[:, :ns].copy(), maxiter=MAXITER)
    dt = time.perf_counter() - t0
    gL = out[0][:, :ns].cpu().numpy()
    ok = (rc == 1) | (rc == 2)
    print(json.dumps({"metric": "binary T-x flashes/s (mix_VLE_Tx, <= %d Newton iterations)" % MAXITER, "model": name, "value": n / ms * 1e3, "ms_per_step": ms, "flashes": n,
                      "converged_fraction_gpu": float(((gc == 1) | (gc == 2)).mean()),
                      "cpu_reference": {"value": ns / dt, "cores": ref.nthreads, "sample": "%d of the same problems (oracle port of VLE.hpp:76-192 on the reference's AbstractModel)" % ns},
                      "max_rel_diff_rhovecL_vs_reference_on_sample": float(np.max(np.abs(gL[:, ok] / rL[:, ok] - 1)))}))
