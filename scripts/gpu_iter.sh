#!/bin/bash
mkdir -p gpurun_out
for c in 4 6 8 3; do timeout 120 python bench.py --concurrency $c 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('conc=$c', d['value'], d['e2e']['value'], d['config']['p50_batch_latency_ms'], d['clocks']['samples'])"; done
