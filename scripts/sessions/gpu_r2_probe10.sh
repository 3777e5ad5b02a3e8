#!/bin/bash
mkdir -p gpurun_out
./scripts/probes/pipe_rates > gpurun_out/p10_pipe_rates.txt 2>&1
cat gpurun_out/p10_pipe_rates.txt
timeout 600 python bench.py --steps 6 --warmup 3 > gpurun_out/p10_bench.json 2> gpurun_out/p10_bench.err
cut -c1-200 gpurun_out/p10_bench.json
