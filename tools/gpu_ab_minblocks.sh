#!/bin/bash
# A/B of the generated kernel's occupancy request (TEQPB200_JIT_MINBLOCKS): headline bench, device-resident leg only
set -u
O=gpurun_out/${1:-r02b}
mkdir -p $O
for n in ${NS:-0 4 5}; do
  TEQPB200_JIT_MINBLOCKS=$n timeout 300 python bench.py --no-cpu --no-e2e --no-secondary --steps 5 --warmup 3 > $O/bench_minblocks$n.json 2> $O/bench_minblocks$n.err
  echo "minblocks=$n rc=$?"; python -c "
import json;d=json.load(open('$O/bench_minblocks$n.json'));print(d['value']/1e9,'Gpt/s frac',d['roofline']['frac'],d['roofline']['flops']['flops_per_point'])"
done
