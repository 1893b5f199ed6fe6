#!/bin/bash
# Multi-GPU round: sharder tests, strong-scaling bench under torchrun (as the driver launches it), the library's one-process e2e leg,
# the C5 sweep, and the host<->device bandwidth table.   usage: gpurun --gpus N -- 'bash tools/gpu_multi.sh N tag'
set -u
N=${1:-2}
O=gpurun_out/${2:-r02m}_n$N
mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 600 python -m pytest tests -m gpu -q -k "sharder or error_path" > $O/pytest_sharder.log 2>&1; echo "pytest sharder rc=$?"; tail -3 $O/pytest_sharder.log
timeout 300 python tools/pcie_probe.py --gpus $N > $O/pcie_probe.jsonl 2> $O/pcie_probe.err; cat $O/pcie_probe.jsonl
NS="1 $N"; if [ $N -eq 8 ]; then NS="4 8"; fi; if [ $N -eq 4 ]; then NS="1 2 4"; fi
for n in $NS; do
  if [ $n -eq 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517"; fi
  timeout 900 $L bench.py --gpus $n --steps 5 --warmup 3 --no-secondary > $O/bench_n$n.json 2> $O/bench_n$n.err; echo "bench n=$n rc=$?"
  python -c "
import json;d=json.load(open('$O/bench_n$n.json'));print('n=$n value %.3f Gpt/s  e2e %.3f Gpt/s  dev_ms %s' % (d['value']/1e9, d['e2e']['value']/1e9, d['e2e']['device_ms_last_call_per_gpu']))"
  timeout 600 $L bench.py --gpus $n --workload C5 --steps 5 --warmup 3 --no-cpu > $O/bench_c5_n$n.json 2> $O/bench_c5_n$n.err; echo "bench C5 n=$n rc=$?"
  python -c "
import json;d=json.load(open('$O/bench_c5_n$n.json'));print('C5 n=$n value %.1f Mpt/s  e2e %.1f Mpt/s' % (d['value']/1e6, d['e2e']['value']/1e6))"
done
for f in $O/*.err; do echo "== $f"; tail -n 2 $f; done | head -60
