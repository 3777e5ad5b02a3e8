#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 \
  scripts/dist_check.py > gpurun_out/m7_dist_w2.txt 2>&1
grep "err fused\|DIST" gpurun_out/m7_dist_w2.txt | cut -c1-600
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29562 \
  scripts/dist_check.py --skew-ms 50 --no-timing > gpurun_out/m7_dist_w2_skew.txt 2>&1
grep "DIST" gpurun_out/m7_dist_w2_skew.txt
for g in 1.0 7.5; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29563 \
  scripts/patch_check.py --arch sdxl --size 1024 --steps 30 --guidance $g --runs 2 > gpurun_out/m7_patch_sdxl_g$g.txt 2>&1
tail -2 gpurun_out/m7_patch_sdxl_g$g.txt
done
