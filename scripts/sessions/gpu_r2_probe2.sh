#!/bin/bash
# round-2 probe 2: after the warp-uniform role loops: kernel tests, phase trace, attention bench, step bench
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -x -q -m gpu > gpurun_out/p2_pytest.txt 2>&1
echo "pytest rc=$?"
tail -5 gpurun_out/p2_pytest.txt
SW_NATIVE_LIB=$PWD/chiaswarm_b200/_lib/libswarm_sm100_trace.so timeout 600 python scripts/trace_igemm.py > gpurun_out/p2_trace.txt 2>&1
echo "trace rc=$?"
timeout 300 python scripts/bench_attn.py > gpurun_out/p2_attn.txt 2>&1
cat gpurun_out/p2_attn.txt
timeout 600 python bench.py --steps 6 --warmup 3 > gpurun_out/p2_bench.json 2> gpurun_out/p2_bench.err
cat gpurun_out/p2_bench.json
