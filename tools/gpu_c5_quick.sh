#!/bin/bash
# quick C5 / C5x iteration: the isochoric GPU tests, both bench lines, and one ncu capture with the shared-memory pipe metrics
set -u
TAG=${1:-c5q}
O=gpurun_out/$TAG
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "fused or isoprops or isochoric or truth" > $O/pytest.log 2>&1; echo "pytest rc=$?"
grep -E "^E  .*(Assert|Error)|passed|failed|^FAILED" $O/pytest.log | cut -c1-330 | head -20
for WL in C5 C5x; do
  timeout 600 python bench.py --workload $WL --no-cpu --no-e2e --no-secondary --steps 5 --warmup 3 > $O/bench_$WL.json 2> $O/bench_$WL.err
  python -c "
import json;d=json.load(open('$O/bench_$WL.json'));print('$WL: %.1f Mpt/s  %.3f ms' % (d['value']/1e6, d['ms_per_step']))" 2>&1 | tail -1
done
M="smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__inst_executed_pipe_fp64.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"
C5S="python bench.py --workload C5 --no-cpu --no-e2e --no-secondary --points 400000 --steps 2 --warmup 1"
timeout 300 $C5S > $O/plain_c5.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gerg_iso -s 1 -c 1 --metrics $M -f -o $O/prof_c5 $C5S > $O/ncu_c5.log 2>&1
echo "ncu rc=$?"
