#!/bin/bash
# One gpurun call (round 2): GPU parity tests, smoke, bench (our arm + reference arm), ncu launch list and one full capture of the
# generated GERG kernel.   Usage: gpurun --timeout 2400 -- 'bash tools/gpu_round_r02.sh [tag]'
set -u
TAG=${1:-r02}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi > $O/nvidia_smi.txt 2>&1
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
cut -c1-3000 $O/bench.json; tail -3 $O/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?"
cat $O/bench_ref.json
SHORT="python bench.py --no-cpu --no-e2e --no-secondary --points 4000000 --steps 2 --warmup 1"
timeout 300 $SHORT > $O/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches.csv $SHORT > $O/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 300 $SHORT > $O/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tqj -s 1 -c 1 \
    --metrics smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__inst_executed_pipe_fp64.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active \
    -f -o $O/prof_gerg $SHORT > $O/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $O
