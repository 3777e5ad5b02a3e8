#!/bin/bash
# 2-GPU: new kernel test (packed attention with seq_q != seq_kv), frame-split t2v (small + zeroscope)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "frames" 2>&1 | tail -5 > gpurun_out/m3_attn_frames.txt
cat gpurun_out/m3_attn_frames.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 \
  scripts/bench_t2v.py --arch tiny-t2v --width 64 --height 64 --frames 8 --steps 3 --runs 1 --tol 5e-2 > gpurun_out/m3_t2v_tiny.txt 2>&1
tail -3 gpurun_out/m3_t2v_tiny.txt
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29552 \
  scripts/bench_t2v.py --steps 25 --runs 3 > gpurun_out/m3_t2v_w2.txt 2>&1
tail -3 gpurun_out/m3_t2v_w2.txt
