#!/bin/bash
mkdir -p gpurun_out
bash scripts/gpu_ncu_full.sh
timeout 200 python scripts/profile_step.py 3 64 > gpurun_out/plain.log 2>&1
timeout 120 python bench.py --steps 60 --warmup 8 > gpurun_out/bench_n1.log 2>gpurun_out/bench_n1.err; tail -1 gpurun_out/bench_n1.log | cut -c1-200
