#!/bin/bash
# First thing to run next: the fused depthwise->pointwise kernel at the flagship shapes, then the whole engine with it.
#   gpurun --timeout 600 -- 'bash scripts/gpu_fused_eval.sh'
mkdir -p gpurun_out
timeout 120 python scripts/check_dwpw_fused.py > gpurun_out/fused_check_full.log 2>&1; echo "fused check rc=$?"; tail -5 gpurun_out/fused_check_full.log
PT="python -m pytest -q --timeout 120 --timeout-method=thread -p no:cacheprovider"
DVC_FUSE_DWPW=1 timeout 300 $PT tests/test_gpu_engine.py -m gpu -x 2>&1 | tail -3        # engine vs CPU oracle with the fused ops in the plan
for f in 0 1 0 1; do DVC_FUSE_DWPW=$f timeout 120 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('fuse=$f', d['value'], d['e2e']['value'], d['config']['p50_batch_latency_ms'], d['gpu_launches'])"; done
