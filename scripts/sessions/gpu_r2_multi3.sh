#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 \
  scripts/dist_check.py > gpurun_out/m4_dist_w2.txt 2>&1
tail -8 gpurun_out/m4_dist_w2.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29562 \
  scripts/dist_check.py --skew-ms 50 --no-timing > gpurun_out/m4_dist_w2_skew.txt 2>&1
tail -6 gpurun_out/m4_dist_w2_skew.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 \
  scripts/bench_t2v.py --arch tiny-t2v --width 64 --height 64 --frames 8 --steps 3 --runs 1 --tol 5e-2 --guidance 1.0 > gpurun_out/m4_t2v_tiny.txt 2>&1
tail -3 gpurun_out/m4_t2v_tiny.txt
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29552 \
  scripts/bench_t2v.py --steps 25 --runs 3 --guidance 1.0 > gpurun_out/m4_t2v_w2_g1.txt 2>&1
tail -3 gpurun_out/m4_t2v_w2_g1.txt
