#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 \
  scripts/dist_check.py > gpurun_out/m5_dist_w2.txt 2>&1
grep -v "^$" gpurun_out/m5_dist_w2.txt | tail -8 | cut -c1-700
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29562 \
  scripts/dist_check.py --skew-ms 50 --no-timing > gpurun_out/m5_dist_w2_skew.txt 2>&1
tail -6 gpurun_out/m5_dist_w2_skew.txt | cut -c1-300
