#!/bin/bash
# Engine tests + smoke + a short bench; each in its own process with a timeout.
mkdir -p gpurun_out
export SDAAS_ROOT=/tmp/sdaas
echo "=== smoke" | tee gpurun_out/engine.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -15 | tee -a gpurun_out/engine.log
echo "=== engine tests" | tee -a gpurun_out/engine.log
timeout 900 python -m pytest tests/test_engine_gpu.py -q -m gpu -x --timeout 600 2>&1 | tail -40 | tee -a gpurun_out/engine.log
for impl in "$@"; do
  echo "=== bench $impl" | tee -a gpurun_out/engine.log
  timeout 900 python bench.py --impl $impl --steps 3 --warmup 3 2>&1 | tail -12 | tee -a gpurun_out/engine.log
done
