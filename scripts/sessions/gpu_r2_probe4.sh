#!/bin/bash
set -x
mkdir -p gpurun_out
./scripts/probes/pipe_rates > gpurun_out/p4_pipe_rates.txt 2>&1
cat gpurun_out/p4_pipe_rates.txt
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -x -q -m gpu > gpurun_out/p4_pytest.txt 2>&1
echo "pytest rc=$?"
tail -5 gpurun_out/p4_pytest.txt
SW_NATIVE_LIB=$PWD/chiaswarm_b200/_lib/libswarm_sm100_trace.so timeout 600 python scripts/trace_igemm.py > gpurun_out/p4_trace.txt 2>&1
echo "trace rc=$?"
timeout 600 python bench.py --steps 6 --warmup 3 > gpurun_out/p4_bench.json 2> gpurun_out/p4_bench.err
cut -c1-200 gpurun_out/p4_bench.json
