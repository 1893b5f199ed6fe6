#!/bin/bash
# Full refresh: parity tests, smoke, bench (ours + reference arm), secondary configs, ncu launch list + full capture of the headline kernel
set -u
mkdir -p gpurun_out
bash tools/gpu_round.sh
timeout 900 python tools/bench_configs.py > gpurun_out/bench_configs.jsonl 2> gpurun_out/bench_configs.err; echo "bench_configs rc=$?"
cut -c1-400 gpurun_out/bench_configs.jsonl
