#!/bin/bash
echo "sdxl kv (signal fence restored):"; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29572 \
  scripts/patch_check.py --arch sdxl --size 1024 --steps 2 --guidance 1.0 2>&1 | grep rel_err | grep -o "rel_err[^,]*"
