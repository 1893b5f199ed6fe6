#!/bin/bash
# usage: tools/gpu_ncu_kernel.sh <workload> <points> <kernel-regex> <out-name>   -- one ncu --set full capture of one kernel of a bench workload
set -u
mkdir -p gpurun_out
SHORT="python bench.py --workload $1 --points $2 --no-cpu --no-e2e --no-secondary --steps 1 --warmup 1"
timeout 300 $SHORT > gpurun_out/plain_$4.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$3 -s 1 -c 1 \
    --metrics smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__inst_executed_pipe_fp64.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active \
    -f -o gpurun_out/prof_$4 $SHORT > gpurun_out/ncu_$4.log 2>&1
echo "ncu $4 rc=$?"
