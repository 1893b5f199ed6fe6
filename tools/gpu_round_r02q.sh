#!/bin/bash
# r02q: the GPU tests added after r02p (genericSAFT with PC-SAFT, mix_VLE_Tx_batch, ...), the 21-component ideal-gas probe in its own process, the VLE solver timings
set -u
TAG=${1:-r02q}
O=gpurun_out/$TAG
mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
grep -E "^E  .*(Assert|Error)|passed|failed|^FAILED" $O/pytest_gpu.log | cut -c1-330 | head -40
timeout 600 python tools/probe_ideal21.py > $O/probe_ideal21.log 2>&1; echo "ideal21 probe rc=$?"; tail -n 6 $O/probe_ideal21.log
timeout 900 python tools/bench_vle.py 2000000 > $O/bench_vle.jsonl 2> $O/bench_vle.err; echo "bench_vle rc=$?"; cat $O/bench_vle.jsonl | cut -c1-700; tail -n 3 $O/bench_vle.err
