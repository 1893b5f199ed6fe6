#!/bin/bash
# Final refresh of round 2: all GPU tests, the no-alloc test alone, smoke, bench (+ reference arm), launch list, ncu of the headline kernel and of the C5 / C5x kernels,
# adjudication on the GPU, VLE solver timings
set -u
TAG=${1:-r02u}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi > $O/nvidia_smi.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
grep -E "^E  .*(Assert|Error)|passed|failed|^FAILED" $O/pytest_gpu.log | cut -c1-330 | head -40
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -2 $O/pytest_noalloc_alone.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
timeout 1200 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
cut -c1-600 $O/bench.json; tail -n 3 $O/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?"
SHORT="python bench.py --no-cpu --no-e2e --no-secondary --points 4000000 --steps 2 --warmup 1"
timeout 300 $SHORT > $O/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches.csv $SHORT > $O/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
M="smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__inst_executed_pipe_fp64.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"
timeout 300 $SHORT > $O/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tqj -s 1 -c 1 --metrics $M -f -o $O/prof_gerg $SHORT > $O/ncu_full.log 2>&1
echo "ncu full rc=$?"
for WL in C5 C5x; do
  S="python bench.py --workload $WL --no-cpu --no-e2e --no-secondary --points 400000 --steps 2 --warmup 1"
  timeout 300 $S > $O/plain_$WL.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:gerg_iso -s 1 -c 1 --metrics $M -f -o $O/prof_$WL $S > $O/ncu_$WL.log 2>&1
  echo "ncu $WL rc=$?"
done
timeout 900 python tools/adjudicate.py --backend gpu --n 2000 --tag r02 > $O/adjudicate_gpu.log 2>&1; echo "adjudicate rc=$?"; cp profiles/r02_truth_adjudication_gpu.json $O/ 2>/dev/null
grep -c VIOLATION $O/adjudicate_gpu.log || true
timeout 900 python tools/bench_vle.py 2000000 > $O/bench_vle.jsonl 2> $O/bench_vle.err; echo "bench_vle rc=$?"
