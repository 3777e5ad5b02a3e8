#!/bin/bash
# 4-GPU validation: dist tests at world 4 (+ skew), patch engine, gang serving with two gangs, t2v, bench N=4
mkdir -p gpurun_out
timeout 2400 python -m pytest tests/test_dist_gpu.py -x -q -k "4" 2>&1 | tail -15 > gpurun_out/m8_dist_pytest_w4.txt
tail -4 gpurun_out/m8_dist_pytest_w4.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29552 \
  scripts/bench_t2v.py --steps 25 --runs 3 > gpurun_out/m8_t2v_w4.txt 2>&1
tail -2 gpurun_out/m8_t2v_w4.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29541 \
  bench.py --gpus 4 --steps 4 --warmup 3 > gpurun_out/m8_bench4.json 2> gpurun_out/m8_bench4.err
cat gpurun_out/m8_bench4.json | cut -c1-3000
tail -3 gpurun_out/m8_bench4.err
