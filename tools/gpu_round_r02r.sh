#!/bin/bash
# r02r: fused C5x entry point (build_Psir_fgradHessian_and_properties_batch): GPU test, A/B against the two separate calls, ncu of the fused launch
set -u
TAG=${1:-r02r}
O=gpurun_out/$TAG
mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "fused or isoprops or isochoric" > $O/pytest_fused.log 2>&1; echo "pytest rc=$?"
grep -E "^E  .*(Assert|Error)|passed|failed|^FAILED" $O/pytest_fused.log | cut -c1-330 | head -20
for V in fused two; do
  if [ $V = two ]; then export TEQPB200_C5X_TWO_CALLS=1; else unset TEQPB200_C5X_TWO_CALLS; fi
  timeout 600 python bench.py --workload C5x --no-cpu --no-e2e --no-secondary --steps 5 --warmup 3 > $O/bench_c5x_$V.json 2> $O/bench_c5x_$V.err
  python -c "
import json;d=json.load(open('$O/bench_c5x_$V.json'));print('C5x $V: %.1f Mpt/s  %.3f ms  launches %s' % (d['value']/1e6, d['ms_per_step'], d.get('gpu_launches')))" 2>&1 | tail -1
done
unset TEQPB200_C5X_TWO_CALLS
M="smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__inst_executed_pipe_fp64.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"
C5S="python bench.py --workload C5x --no-cpu --no-e2e --no-secondary --points 400000 --steps 2 --warmup 1"
timeout 300 $C5S > $O/plain_c5x.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gerg_iso -s 1 -c 1 --metrics $M -f -o $O/prof_c5x $C5S > $O/ncu_c5x.log 2>&1
echo "ncu c5x rc=$?"
