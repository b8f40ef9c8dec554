#!/bin/bash
mkdir -p gpurun_out
PT="python -m pytest -q --timeout 120 --timeout-method=thread -p no:cacheprovider"
timeout 300 $PT tests/test_gpu_kernels.py tests/test_gpu_engine.py -m gpu -x 2>&1 | tail -5
timeout 120 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print(d['value'], d['e2e']['value'], d['config']['p50_batch_latency_ms'], d['clocks']['samples'])"
timeout 60 python scripts/profile_ops.py >/dev/null 2>&1
python - <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.ops import kernels as K
raw = torch.from_numpy(np.random.default_rng(0).integers(0, 256, (64, 300, 300, 3), dtype=np.uint8)).cuda()
t = K.make_area_tables(300, 300, 400, 400, "cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
ts = []
for _ in range(7):
    flush.zero_(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); K.resize_u8(raw, t); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b) * 1e3)
print("resize 64x300x300->400x400 us:", min(ts))
PY
