#!/bin/bash
# Full refresh (round 2): all GPU tests, smoke, bench (+ reference arm), launch list, full ncu capture of the generated kernel, adjudication table on the GPU
set -u
TAG=${1:-r02h}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi > $O/nvidia_smi.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
grep -E "^E  .*(Assert|Error)|passed|failed|^FAILED" $O/pytest_gpu.log | cut -c1-330 | head -40
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
timeout 1200 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
cut -c1-1500 $O/bench.json; tail -n 3 $O/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?"
SHORT="python bench.py --no-cpu --no-e2e --no-secondary --points 4000000 --steps 2 --warmup 1"
timeout 300 $SHORT > $O/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches.csv $SHORT > $O/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 300 $SHORT > $O/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tqj -s 1 -c 1 \
    --metrics smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__inst_executed_pipe_fp64.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active \
    -f -o $O/prof_gerg $SHORT > $O/ncu_full.log 2>&1
echo "ncu full rc=$?"
timeout 900 python tools/adjudicate.py --backend gpu --n 2000 --tag r02 > $O/adjudicate_gpu.log 2>&1; echo "adjudicate rc=$?"; cp profiles/r02_truth_adjudication_gpu.json $O/ 2>/dev/null
grep -c VIOLATION $O/adjudicate_gpu.log || true
