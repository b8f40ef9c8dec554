#!/bin/bash
# quick GPU iteration: kernel tests (gemm group isolated), per-layer timings, bench
mkdir -p gpurun_out
PT="python -m pytest -q --timeout 180 --timeout-method=thread -p no:cacheprovider"
timeout 400 $PT tests/test_gpu_kernels.py -k "gemm" > gpurun_out/t_gemm.log 2>&1; echo "gemm rc=$?"
timeout 400 $PT tests/test_gpu_kernels.py -k "not gemm" > gpurun_out/t_kernels.log 2>&1; echo "kernels rc=$?"
timeout 500 $PT tests/test_gpu_engine.py > gpurun_out/t_engine.log 2>&1; echo "engine rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/t_gemm.log gpurun_out/t_kernels.log gpurun_out/t_engine.log | head -20
timeout 300 python scripts/bench_layers.py > gpurun_out/layers.log 2>&1; echo "layers rc=$?"; cat gpurun_out/layers.log | tail -40
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"
tail -1 gpurun_out/bench.log | cut -c1-330; tail -3 gpurun_out/bench.err
