#!/bin/bash
# gpurun --gpus N --timeout 1500 -- 'bash scripts/gpu_cfg.sh N': BASELINE.json configs 3-5 on N GPUs -> gpurun_out/configs_nN.jsonl
N=${1:-2}
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
OUT=gpurun_out/configs_n$N.jsonl; : > $OUT
timeout 300 python benchmarks/configs.py --config elastic --gpus $N 2> gpurun_out/cfg_elastic_n$N.err | tee -a $OUT | cut -c1-900; tail -2 gpurun_out/cfg_elastic_n$N.err
for fps in 30 60; do timeout 300 python benchmarks/configs.py --config live --gpus $N --fps $fps 2> gpurun_out/cfg_live_n$N.err | tee -a $OUT | cut -c1-400; done; tail -2 gpurun_out/cfg_live_n$N.err
timeout 400 python benchmarks/configs.py --config multi_client --gpus $N 2> gpurun_out/cfg_multi_n$N.err | tee -a $OUT | cut -c1-600; tail -2 gpurun_out/cfg_multi_n$N.err
