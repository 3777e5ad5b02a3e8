#!/bin/bash
# Runs each kernel-family test group in its own process (a trap in one kernel kills only its
# group) with a hard timeout; logs go to gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
for grp in "$@"; do
  echo "=== $grp" | tee -a gpurun_out/check.log
  timeout 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "$grp" -x --timeout 120 2>&1 | tail -25 | tee -a gpurun_out/check.log
done
