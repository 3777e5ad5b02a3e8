#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/p7_pytest.txt 2>&1
echo "pytest rc=$?"
tail -5 gpurun_out/p7_pytest.txt
SW_NATIVE_LIB=$PWD/chiaswarm_b200/_lib/libswarm_sm100_trace.so timeout 600 python scripts/trace_gn.py > gpurun_out/p7_gn.txt 2>&1
cat gpurun_out/p7_gn.txt
SW_NATIVE_LIB=$PWD/chiaswarm_b200/_lib/libswarm_sm100_trace.so timeout 600 python scripts/trace_igemm.py conv > gpurun_out/p7_trace_conv.txt 2>&1
grep -A1 auto gpurun_out/p7_trace_conv.txt
timeout 600 python bench.py --steps 6 --warmup 3 > gpurun_out/p7_bench.json 2> gpurun_out/p7_bench.err
cut -c1-200 gpurun_out/p7_bench.json
