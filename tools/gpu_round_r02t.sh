#!/bin/bash
# r02t: C4 get_ATrhoXi with register-capped kernel variants (TEQPB200_ATRHOXI_MINB = 0 | 6 | 8), mix_VLE_Tp_batch tests
set -u
TAG=${1:-r02t}
O=gpurun_out/$TAG
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "mix_vle or pure_vle or atrhoxi" > $O/pytest.log 2>&1; echo "pytest rc=$?"
grep -E "^E  .*(Assert|Error)|passed|failed|^FAILED" $O/pytest.log | cut -c1-330 | head -20
for MB in 0 6 8; do
  TEQPB200_ATRHOXI_MINB=$MB timeout 900 python bench.py --workload C4 --points 2000000 --no-cpu --no-e2e --no-secondary --steps 3 --warmup 2 > $O/bench_c4_mb$MB.json 2> $O/bench_c4_mb$MB.err
  python -c "
import json;d=json.load(open('$O/bench_c4_mb$MB.json'));print('C4 minb=$MB: %.2f Mpt/s  %.1f ms  checksum %s' % (d['value']/1e6, d['ms_per_step'], d['config'].get('checksum')))" 2>&1 | tail -1
done
for MB in 6 8; do
  TEQPB200_ATRHOXI_MINB=$MB timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "atrhoxi and (GSAFT or MIE)" > $O/pytest_mb$MB.log 2>&1; echo "pytest minb=$MB rc=$?"; tail -1 $O/pytest_mb$MB.log
done
