#!/bin/bash
mkdir -p gpurun_out
timeout 280 python bench.py --impl reference --gpus 1 --steps 4 --warmup 0 > gpurun_out/bench_ref_n1.log 2>gpurun_out/bench_ref.err; echo "ref n1 rc=$?"; tail -1 gpurun_out/bench_ref_n1.log | cut -c1-300
timeout 280 python bench.py --impl reference --gpus 2 --steps 2 --warmup 0 > gpurun_out/bench_ref_n2.log 2>>gpurun_out/bench_ref.err; echo "ref n2 rc=$?"; tail -1 gpurun_out/bench_ref_n2.log | cut -c1-300
bash scripts/gpu_ncu_full.sh
timeout 200 python scripts/profile_step.py 3 64 > gpurun_out/plain.log 2>&1
