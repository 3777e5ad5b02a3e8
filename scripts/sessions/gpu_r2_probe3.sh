#!/bin/bash
set -x
mkdir -p gpurun_out
SW_NATIVE_LIB=$PWD/chiaswarm_b200/_lib/libswarm_sm100_trace.so timeout 600 python scripts/trace_gn.py > gpurun_out/p3_gn.txt 2>&1
cat gpurun_out/p3_gn.txt
SW_IGEMM_TUNING=0 timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/p3_bench_notune.json 2> gpurun_out/p3_bench_notune.err
cat gpurun_out/p3_bench_notune.json | cut -c1-200
timeout 600 python scripts/profile_kernels_plan.py > gpurun_out/p3_step_profile.txt 2>&1
head -40 gpurun_out/p3_step_profile.txt
