#!/bin/bash
# ncu --set full captures of the C5 Hessian kernel and the three C4 get_ATrhoXi part kernels, plus the C4 A/B (small jets vs rectangular jet)
set -u
O=gpurun_out/${1:-r02i}
mkdir -p $O
M="smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__inst_executed_pipe_fp64.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"
# C4 A/B, device-resident, 10^6 points
for v in split rect; do
  if [ $v = rect ]; then export TEQPB200_ATRHOXI_RECT=1; else unset TEQPB200_ATRHOXI_RECT; fi
  timeout 600 python bench.py --workload C4 --points 1000000 --no-cpu --no-e2e --steps 3 --warmup 2 > $O/bench_C4_$v.json 2> $O/bench_C4_$v.err; echo "C4 $v rc=$?"
  python -c "
import json;d=json.load(open('$O/bench_C4_$v.json'));print('C4 $v: %.2f Mpt/s, launches %d' % (d['value']/1e6, d['gpu_launches']))"
done
unset TEQPB200_ATRHOXI_RECT
S5="python bench.py --workload C5 --points 200000 --no-cpu --no-e2e --steps 1 --warmup 1"
timeout 300 $S5 > $O/plain_c5.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:gerg_isochoric -s 1 -c 1 --metrics $M -f -o $O/prof_c5 $S5 > $O/ncu_c5.log 2>&1; echo "ncu c5 rc=$?"
S4="python bench.py --workload C4 --points 100000 --no-cpu --no-e2e --steps 1 --warmup 1"
timeout 300 $S4 > $O/plain_c4.log 2>&1 && timeout 1200 ncu --set full --clock-control none --import-source on -k regex:split_ATrhoXi -s 3 -c 3 --metrics $M -f -o $O/prof_c4 $S4 > $O/ncu_c4.log 2>&1; echo "ncu c4 rc=$?"
ls -la $O
