#!/bin/bash
mkdir -p gpurun_out
timeout 100 python scripts/sweep_dw_split.py 2>&1 | tail -15
PT="python -m pytest -q --timeout 120 --timeout-method=thread -p no:cacheprovider"
timeout 200 $PT tests/test_gpu_kernels.py -m gpu -x -k "dwconv" 2>&1 | tail -2
for sp in 1 0; do DVC_DW_SPLIT=$sp timeout 120 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('split=$sp', d['value'], d['e2e']['value'], d['config']['p50_batch_latency_ms'])"; done
