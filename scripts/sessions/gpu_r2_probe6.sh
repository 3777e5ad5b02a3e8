#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 120 ./scripts/probes/tmem_rate > gpurun_out/p6_tmem_rate.txt 2>&1
cat gpurun_out/p6_tmem_rate.txt
SW_NATIVE_LIB=$PWD/chiaswarm_b200/_lib/libswarm_sm100_trace.so timeout 600 python scripts/trace_igemm.py conv > gpurun_out/p6_trace_conv.txt 2>&1
cat gpurun_out/p6_trace_conv.txt
