#!/bin/bash
mkdir -p gpurun_out
PT="python -m pytest -q --timeout 120 --timeout-method=thread -p no:cacheprovider"
timeout 300 $PT tests/test_gpu_kernels.py tests/test_gpu_engine.py tests/test_zz_no_alloc.py -m gpu -x 2>&1 | tail -5
timeout 120 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print(d['value'], d['e2e']['value'], d['config']['p50_batch_latency_ms'], d['clocks']['samples'])"
