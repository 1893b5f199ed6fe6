#!/bin/bash
# C5 (GERG-2008, 21 components, build_Psir_fgradHessian) with the lane-cooperative kernels (tq_gerg_iso_coop.cuh): parity tests first, then an
# A/B over the lanes per point (0 = thread-per-point kernels) and the warps per CTA, then an ncu capture of the default
set -u
TAG=${1:-r02p}
O=gpurun_out/$TAG
mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -q > $O/pytest_iso.log 2>&1; echo "pytest rc=$?"
grep -E "^E  .*(Assert|Error)|passed|failed|^FAILED" $O/pytest_iso.log | cut -c1-300 | head -20
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -2 $O/pytest_noalloc_alone.log
for L in 0 4 8 16; do
  for W in 16 8; do
    [ $L = 0 ] && [ $W = 8 ] && continue
    for WL in C5 C5x; do
      TEQPB200_ISO_COOP_L=$L TEQPB200_ISO_COOP_WARPS=$W timeout 600 python bench.py --workload $WL --no-cpu --no-e2e --steps 5 --warmup 3 > $O/bench_${WL}_L${L}_W$W.json 2> $O/bench_${WL}_L${L}_W$W.err
      python -c "
import json;d=json.load(open('$O/bench_${WL}_L${L}_W$W.json'));print('$WL L=$L W=$W: %.1f Mpt/s  %.3f ms' % (d['value']/1e6, d['ms_per_step']))" 2>&1 | tail -1
    done
  done
done
SHORT="python bench.py --workload C5 --no-cpu --no-e2e --points 400000 --steps 2 --warmup 1"
timeout 300 $SHORT > $O/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gerg_iso -s 1 -c 1 \
    --metrics smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__inst_executed_pipe_fp64.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active \
    -f -o $O/prof_c5 $SHORT > $O/ncu_full.log 2>&1
echo "ncu rc=$?"
