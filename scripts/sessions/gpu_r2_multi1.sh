#!/bin/bash
# 2-GPU validation: dist tests (fused collectives, skew, patch engine, gang serving) + bench at N=2
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_dist_gpu.py -x -q 2>&1 | tail -40 > gpurun_out/m1_dist_pytest.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 \
  bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/m1_bench2.json 2> gpurun_out/m1_bench2.err
tail -5 gpurun_out/m1_dist_pytest.txt
cat gpurun_out/m1_bench2.json
tail -5 gpurun_out/m1_bench2.err
