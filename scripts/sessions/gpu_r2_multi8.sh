#!/bin/bash
mkdir -p gpurun_out
for mode in tp kv; do
SW_PATCH_ATTN=$mode timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29563 \
  scripts/patch_check.py --arch sd15 --size 512 --steps 3 --guidance 1.0 --runs 2 > gpurun_out/m12_patch_sd15_$mode.txt 2>&1
echo $mode; tail -2 gpurun_out/m12_patch_sd15_$mode.txt | cut -c1-600
SW_PATCH_ATTN=$mode timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29564 \
  scripts/patch_check.py --arch sdxl --size 1024 --steps 30 --guidance 1.0 --runs 2 --skip-single > gpurun_out/m12_patch_sdxl_$mode.txt 2>&1
tail -2 gpurun_out/m12_patch_sdxl_$mode.txt | cut -c1-600
done
