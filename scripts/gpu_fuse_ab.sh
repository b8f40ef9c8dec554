#!/bin/bash
# gpurun --timeout 300 -- 'bash scripts/gpu_fuse_ab.sh': same-box A/B of the fused dw->pw launch for single blocks inside the full step
mkdir -p gpurun_out
for v in 0 conv3 0 conv3; do
DVC_FUSE_DWPW=$v timeout 120 python bench.py --steps 40 --warmup 5 --sustain-s 0 > gpurun_out/bench_fuse_$v.json 2> gpurun_out/bench_fuse_$v.err; echo "bench fuse=$v rc=$?"
tail -1 gpurun_out/bench_fuse_$v.json | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('value',d['value'],'p50',d['config']['p50_batch_latency_ms'],'e2e',d['e2e']['value'],'launches',d['gpu_launches'])"
done
