#!/bin/bash
# Full refresh of the current tree (round 2, after the lane-cooperative C5 kernels, the no-allocation entry point and the ideal-gas kinds):
# all GPU tests, the no-alloc test alone, smoke, bench (+ reference arm), launch list, ncu of the headline kernel and of the C5 kernel,
# C5 A/B over the lanes per point, adjudication table on the GPU
set -u
TAG=${1:-r02p}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi > $O/nvidia_smi.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
grep -E "^E  .*(Assert|Error)|passed|failed|^FAILED" $O/pytest_gpu.log | cut -c1-330 | head -40
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -2 $O/pytest_noalloc_alone.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
timeout 1200 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
cut -c1-1500 $O/bench.json; tail -n 3 $O/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?"
SHORT="python bench.py --no-cpu --no-e2e --no-secondary --points 4000000 --steps 2 --warmup 1"
timeout 300 $SHORT > $O/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches.csv $SHORT > $O/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
M="smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__inst_executed_pipe_fp64.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"
timeout 300 $SHORT > $O/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tqj -s 1 -c 1 --metrics $M -f -o $O/prof_gerg $SHORT > $O/ncu_full.log 2>&1
echo "ncu full rc=$?"
# C5: lanes per point (0 = thread-per-point kernels)
for L in 0 4 8 16; do
  for WL in C5 C5x; do
    TEQPB200_ISO_COOP_L=$L timeout 600 python bench.py --workload $WL --no-cpu --no-e2e --no-secondary --steps 5 --warmup 3 > $O/bench_${WL}_L${L}.json 2> $O/bench_${WL}_L${L}.err
    python -c "
import json;d=json.load(open('$O/bench_${WL}_L${L}.json'));print('$WL L=$L: %.1f Mpt/s  %.3f ms' % (d['value']/1e6, d['ms_per_step']))" 2>&1 | tail -1
  done
done
C5S="python bench.py --workload C5 --no-cpu --no-e2e --no-secondary --points 400000 --steps 2 --warmup 1"
timeout 300 $C5S > $O/plain_c5.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gerg_iso -s 1 -c 1 --metrics $M -f -o $O/prof_c5 $C5S > $O/ncu_c5.log 2>&1
echo "ncu c5 rc=$?"
timeout 900 python tools/adjudicate.py --backend gpu --n 2000 --tag r02 > $O/adjudicate_gpu.log 2>&1; echo "adjudicate rc=$?"; cp profiles/r02_truth_adjudication_gpu.json $O/ 2>/dev/null
grep -c VIOLATION $O/adjudicate_gpu.log || true
# ideal-gas kind with 21 components through the generic isoprops / cross kernels, in its own process under the sanitizer (a 21-component jet-sweep
# isoprops launch of the GERG family faulted in r02f)
timeout 600 compute-sanitizer --tool memcheck python tools/probe_ideal21.py > $O/probe_ideal21.log 2>&1; echo "ideal21 probe rc=$?"; tail -n 6 $O/probe_ideal21.log
