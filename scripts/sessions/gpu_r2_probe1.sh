#!/bin/bash
# round-2 probe 1: phase trace of the linear shapes + library kernel listing
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/p1_smi.txt
SW_NATIVE_LIB=$PWD/chiaswarm_b200/_lib/libswarm_sm100_trace.so timeout 600 python scripts/trace_igemm.py > gpurun_out/p1_trace.txt 2>&1
echo "trace rc=$?"
timeout 300 python scripts/lib_kernels.py > gpurun_out/p1_lib_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size,launch__cluster_dim_x,launch__cluster_dim_y,launch__shared_mem_per_block_dynamic,launch__registers_per_thread --clock-control none -s 21 -c 30 --csv --log-file gpurun_out/p1_lib_kernels.csv python scripts/lib_kernels.py > gpurun_out/p1_lib_ncu.log 2>&1
echo "ncu rc=$?"
tail -5 gpurun_out/p1_trace.txt
