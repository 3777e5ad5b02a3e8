#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/p9_attn.txt
for st in 0 700 1100 1500; do
  echo "=== SW_ATTN_STAGGER=$st" >> gpurun_out/p9_attn.txt
  SW_ATTN_STAGGER=$st timeout 300 python scripts/bench_attn.py 2>&1 | grep -E "4096, 4096|1024, 1024" >> gpurun_out/p9_attn.txt
done
cat gpurun_out/p9_attn.txt
SW_NATIVE_LIB=$PWD/chiaswarm_b200/_lib/libswarm_sm100_trace.so timeout 300 python scripts/trace_attn.py > gpurun_out/p9_attn_trace.txt 2>&1
cat gpurun_out/p9_attn_trace.txt
timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "attn or groupnorm or gn" > gpurun_out/p9_pytest.txt 2>&1
tail -3 gpurun_out/p9_pytest.txt
timeout 300 python scripts/trace_gn.py 2>&1 | grep -v "^    " | cut -c1-110 > gpurun_out/p9_gn.txt
cat gpurun_out/p9_gn.txt
timeout 600 python bench.py --steps 6 --warmup 3 > gpurun_out/p9_bench.json 2> gpurun_out/p9_bench.err
cut -c1-200 gpurun_out/p9_bench.json
