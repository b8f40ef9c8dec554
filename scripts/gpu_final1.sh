#!/bin/bash
# gpurun --timeout 300 -- 'bash scripts/gpu_final1.sh': full single-GPU suite, smoke and the bench line on the final tree
mkdir -p gpurun_out
timeout 200 python -m pytest tests -m gpu -x -q --timeout 120 -p no:cacheprovider 2>&1 | tail -6 | tee gpurun_out/t_all_gpu_final.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 120 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_final_n1.json 2> gpurun_out/bench_final_n1.err; echo "bench rc=$?"
tail -1 gpurun_out/bench_final_n1.json | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('value',d['value'],'p50',d['config']['p50_batch_latency_ms'],'e2e',d['e2e']['value'],'sustained',(d['sustained'] or {}).get('value'),'launches',d['gpu_launches'],'clocks',d['clocks'])"
