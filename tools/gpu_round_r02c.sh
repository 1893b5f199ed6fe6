#!/bin/bash
set -u
O=gpurun_out/${1:-r02c}
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
tail -30 $O/pytest_gpu.log
NS="6 7" bash tools/gpu_ab_minblocks.sh $(basename $O)
