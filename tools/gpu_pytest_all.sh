#!/bin/bash
set -u
O=gpurun_out/${1:-r02d}
mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
grep -E "^E  .*(Assert|Error)|passed|failed|^FAILED" $O/pytest_gpu.log | cut -c1-330 | head -80
