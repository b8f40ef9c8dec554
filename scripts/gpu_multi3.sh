#!/bin/bash
mkdir -p gpurun_out
i=0
for cfg in "smi 200" "nvml 100" "nvml 20"; do
set -- $cfg; i=$((i+1))
DVC_CLOCK_BACKEND=$1 DVC_CLOCK_PERIOD_MS=$2 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2952$i bench.py --gpus 2 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('$1 $2', d['value'], d['e2e']['value'], d['config']['p50_batch_latency_ms'], d['clocks'])"
done
