#!/bin/bash
mkdir -p gpurun_out
for cl in 2 4 8 16; do
  echo "=== SW_GN_MAXCL=$cl" >> gpurun_out/p8_gn.txt
  SW_GN_MAXCL=$cl SW_NATIVE_LIB=$PWD/chiaswarm_b200/_lib/libswarm_sm100_trace.so timeout 300 python scripts/trace_gn.py 2>&1 | grep -v "^    " >> gpurun_out/p8_gn.txt
done
cat gpurun_out/p8_gn.txt | cut -c1-120
SW_NATIVE_LIB=$PWD/chiaswarm_b200/_lib/libswarm_sm100_trace.so timeout 300 python scripts/trace_attn.py > gpurun_out/p8_attn_trace.txt 2>&1
cat gpurun_out/p8_attn_trace.txt
