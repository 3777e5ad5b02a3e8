#!/bin/bash
mkdir -p gpurun_out
for d in 16 32 64 0; do
SW_FUSED_DBG=$d timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2958$((d/16)) \
  scripts/dist_check.py > gpurun_out/m6_dist_dbg$d.txt 2>&1
echo "DBG $d"; grep "reducescatter" gpurun_out/m6_dist_dbg$d.txt | cut -c1-90;  grep "reducescatter" gpurun_out/m6_dist_dbg$d.txt | grep -o "us_fused_local_only.*" | cut -c1-120
done
