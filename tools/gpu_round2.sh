#!/bin/bash
# GPU round for the secondary configs: parity tests, tools/bench_configs.py, ncu capture of the GERG isochoric kernel.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python tools/bench_configs.py --only C5 > gpurun_out/bench_configs.jsonl 2> gpurun_out/bench_configs.err; echo "bench_configs rc=$?"
cat gpurun_out/bench_configs.jsonl; tail -3 gpurun_out/bench_configs.err
SHORT="python tools/bench_configs.py --only C5 --no-cpu --scale 0.2 --steps 1 --warmup 1"
timeout 300 $SHORT > gpurun_out/plain_c5.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gerg_isochoric -s 1 -c 1 \
    --metrics smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__inst_executed_pipe_fp64.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active \
    -f -o gpurun_out/prof_gerg_iso $SHORT > gpurun_out/ncu_iso.log 2>&1
echo "ncu iso rc=$?"
