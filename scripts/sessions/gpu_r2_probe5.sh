#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python scripts/bench_kmajor.py > gpurun_out/p5_kmajor.txt 2>&1
cat gpurun_out/p5_kmajor.txt
