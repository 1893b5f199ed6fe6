#!/bin/bash
# r02zc: full GPU tier after adding the advancedPRaEres family, the no-allocation test alone, smoke, and a throughput probe of the new kernels
set -u
O=gpurun_out/r02zc
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "^E  .*(Assert|Error)|passed|failed|^FAILED" $O/pytest_gpu.log | cut -c1-330 | head -20
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke OK')" > $O/smoke.log 2>&1; echo "smoke rc=$?"
timeout 200 python - > $O/apr_probe.json 2> $O/apr_probe.err <<'P'
import json, sys, torch
sys.path.insert(0, "tests")
import cases, teqp_b200
spec = cases.SPECS["APR_CO2_N2"][0]
m = teqp_b200.make_model(spec)
n = 20_000_000
g = torch.Generator(device="cuda").manual_seed(1)
T = (200 + 300 * torch.rand(n, device="cuda", dtype=torch.float64, generator=g))
rho = torch.exp(torch.rand(n, device="cuda", dtype=torch.float64, generator=g) * 9.0)
x0 = 0.2 + 0.6 * torch.rand(n, device="cuda", dtype=torch.float64, generator=g)
x = torch.stack([x0, 1 - x0]).contiguous()
out = {}
for name, fn in [("get_Ar04n", lambda: m.get_Ar0n_batch(4, T, rho, x)), ("get_Ar20", lambda: m.get_Arxy_batch([(2, 0)], T, rho, x))]:
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): fn()
    e1.record(); torch.cuda.synchronize()
    out[name] = {"points": n, "ms": e0.elapsed_time(e1) / 5, "Gpt_per_s": n / (e0.elapsed_time(e1) / 5 * 1e-3) / 1e9}
print(json.dumps(out))
P
echo "probe rc=$?"; cat $O/apr_probe.json; tail -3 $O/apr_probe.err
