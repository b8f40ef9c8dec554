#!/bin/bash
# gpurun --gpus 2 --timeout 240 -- 'bash scripts/gpu_final2.sh': the multi-GPU tests on the final tree (2 GPUs)
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout 180 -p no:cacheprovider 2>&1 | tail -8 | tee gpurun_out/t_multi_final.log
