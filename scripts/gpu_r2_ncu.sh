#!/bin/bash
mkdir -p gpurun_out
timeout 120 ncu --set full --clock-control none --import-source on -k regex:k_igemm_lean_rs -s 6 -c 1 -f -o gpurun_out/ncu_fused_lean_rs \
  python scripts/rs_probe.py > gpurun_out/m20_ncu.log 2>&1
tail -2 gpurun_out/m20_ncu.log
timeout 170 compute-sanitizer --tool racecheck --print-limit 20 python -m pytest tests/test_kernels_gpu.py -x -q \
  -k "test_linear and 128-64-64 or test_layernorm and 77 or test_groupnorm and 4096-320-0 or test_flash_attn_causal and 45" > gpurun_out/m20_racecheck.txt 2>&1
tail -12 gpurun_out/m20_racecheck.txt
timeout 120 compute-sanitizer --tool memcheck --print-limit 20 python -m pytest tests/test_kernels_gpu.py -x -q \
  -k "test_conv3x3_stride2_implicit and 10-18 or test_upsample2x_folded_into_conv and 5-9" > gpurun_out/m20_memcheck.txt 2>&1
tail -8 gpurun_out/m20_memcheck.txt
