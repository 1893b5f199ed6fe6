#!/bin/bash
# headline iteration: the GERG / multifluid GPU tests, the bench line (no CPU leg), one ncu capture of the generated kernel
set -u
TAG=${1:-hq}
O=gpurun_out/$TAG
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "GERG or multifluid or golden or edge or headline or truth or shards" > $O/pytest.log 2>&1; echo "pytest rc=$?"
grep -E "^E  .*(Assert|Error)|passed|failed|^FAILED" $O/pytest.log | cut -c1-330 | head -20
timeout 900 python bench.py --no-cpu --no-secondary --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('$O/bench.json'));r=d['roofline'];print('C3: %.3f Gpt/s  %.3f ms  frac %.3f  flops/pt %s  e2e %.3f' % (d['value']/1e9, d['ms_per_step'], r['frac'], r['flops']['flops_per_point'], d['e2e']['value']/1e9))"
M="smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__inst_executed_pipe_fp64.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"
SHORT="python bench.py --no-cpu --no-e2e --no-secondary --points 4000000 --steps 2 --warmup 1"
timeout 300 $SHORT > $O/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tqj -s 1 -c 1 --metrics $M -f -o $O/prof_gerg $SHORT > $O/ncu_full.log 2>&1
echo "ncu rc=$?"
