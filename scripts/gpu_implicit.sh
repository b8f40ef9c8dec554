#!/bin/bash
# gpurun --timeout 300 -- 'bash scripts/gpu_implicit.sh': im2col-TMA tile check, implicit conv numerics, then suite + A/B bench
mkdir -p gpurun_out
set -o pipefail
timeout 120 python -m pytest tests/test_gpu_kernels.py -q -x -p no:cacheprovider -k "im2col_tma_tile" 2>&1 | tail -25 | tee gpurun_out/implicit_tile.log || { echo "TILE TEST FAILED"; exit 1; }
timeout 120 python -m pytest tests/test_gpu_kernels.py -q -x -p no:cacheprovider -k "conv3x3s2_implicit" 2>&1 | tail -25 | tee gpurun_out/implicit_conv.log || { echo "CONV TEST FAILED"; exit 1; }
timeout 200 python -m pytest tests -m gpu -x -q --timeout 120 -p no:cacheprovider 2>&1 | tail -6 | tee gpurun_out/t_all_gpu.log
for v in 1 0; do
DVC_IMPLICIT_CONV=$v timeout 120 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_implicit_$v.json 2> gpurun_out/bench_implicit_$v.err; echo "bench implicit=$v rc=$?"
tail -1 gpurun_out/bench_implicit_$v.json | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('value',d['value'],'p50',d['config']['p50_batch_latency_ms'],'e2e',d['e2e']['value'],'sustained',(d['sustained'] or {}).get('value'),'launches',d['gpu_launches'])"
done
