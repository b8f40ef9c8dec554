#!/bin/bash
# gpurun --gpus 8 --timeout 900 -- 'bash scripts/gpu_n8_bench.sh [Ns]': bench.py at the given GPU counts on one box + dispatch timeline at 8
NS=${1:-"8 4 2 1"}
mkdir -p gpurun_out
show() { tail -1 $1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.readline()); print('N',d['n_gpus'],'value',d['value'],'ms/step',d['ms_per_step'],'p50',d['config'].get('p50_batch_latency_ms'),'e2e',d['e2e']['value'],'sharded',(d.get('e2e_sharded') or {}).get('value'),'sustained',(d.get('sustained') or {}).get('value'),'coll',d.get('collectives_in_timed_regions'),'clocks',(d.get('clocks') or {}).get('sm_mhz'),(d.get('clocks') or {}).get('reasons'))
except Exception as e: print('unparsable', e)"; }
for N in $NS; do
if [ "$N" = "1" ]; then timeout 200 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.log 2> gpurun_out/bench_n1.err
else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.log 2> gpurun_out/bench_n$N.err; fi
echo "bench N=$N rc=$?"; show gpurun_out/bench_n$N.log
done
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1 --master-port 29577 scripts/diag_timeline.py 2>&1 | grep "window\|gpu" | cut -c1-330 | tee gpurun_out/timeline_n8.log
