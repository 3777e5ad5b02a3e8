#!/bin/bash
# 8-GPU validation: fused collectives at world 8, patch engine 2x4, frame-split t2v 2x4, bench N=8 (with sdxl_patch)
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 \
  scripts/dist_check.py > gpurun_out/m14_dist_w8.txt 2>&1
grep "err fused\|DIST" gpurun_out/m14_dist_w8.txt | cut -c1-400
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29562 \
  scripts/tp_block_check.py > gpurun_out/m14_tp_w8.txt 2>&1
grep "TP_BLOCK" gpurun_out/m14_tp_w8.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29552 \
  scripts/bench_t2v.py --steps 25 --runs 3 > gpurun_out/m14_t2v_w8.txt 2>&1
tail -2 gpurun_out/m14_t2v_w8.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 \
  bench.py --gpus 8 --steps 4 --warmup 3 > gpurun_out/m14_bench8.json 2> gpurun_out/m14_bench8.err
cat gpurun_out/m14_bench8.json | cut -c1-3000
tail -3 gpurun_out/m14_bench8.err
