#!/bin/bash
# Runs on the GPU box via gpurun: kernel tests in isolated processes (a hung kernel only loses its group), then bench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
PT="python -m pytest -q --timeout 180 --timeout-method=thread -p no:cacheprovider"
timeout 400 $PT tests/test_gpu_kernels.py -k "not gemm" > gpurun_out/t_kernels.log 2>&1; echo "kernels(non-gemm) rc=$?"
timeout 400 $PT tests/test_gpu_kernels.py -k "gemm" > gpurun_out/t_gemm.log 2>&1; echo "gemm rc=$?"
timeout 500 $PT tests/test_gpu_engine.py > gpurun_out/t_engine.log 2>&1; echo "engine rc=$?"
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/t_kernels.log gpurun_out/t_gemm.log gpurun_out/t_engine.log | head -40
grep -E "^E  " gpurun_out/t_engine.log | cut -c1-300 | head -20
tail -2 gpurun_out/bench.log | cut -c1-600; tail -5 gpurun_out/bench.err
