#!/bin/bash
# gpurun --timeout 1500 -- 'bash scripts/gpu_one.sh': one-GPU round: full GPU test suite, smoke, bench at several lane counts, torch baseline
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q --timeout 900 -p no:cacheprovider 2>&1 | tail -6 | tee gpurun_out/t_all_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
for c in 4 6 8; do
timeout 300 python bench.py --steps 20 --warmup 5 --concurrency $c > gpurun_out/bench_n1_c$c.log 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -1 gpurun_out/bench_n1_c$c.log | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('conc',d['config']['batches_in_flight'],'value',d['value'],'ms/step',d['ms_per_step'],'p50',d['config']['p50_batch_latency_ms'],'e2e',d['e2e']['value'],'sustained',(d['sustained'] or {}).get('value'),'clocks',d['clocks'])"; tail -2 gpurun_out/bench_n1.err | cut -c1-300
done
timeout 300 python bench.py --impl torch --steps 20 --warmup 5 > gpurun_out/bench_torch_n1.log 2> gpurun_out/bench_torch.err; echo "torch rc=$?"; tail -1 gpurun_out/bench_torch_n1.log | cut -c1-400; tail -3 gpurun_out/bench_torch.err | cut -c1-300
