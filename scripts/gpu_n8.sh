#!/bin/bash
# gpurun --gpus 8 --timeout 1200 -- 'bash scripts/gpu_n8.sh': the whole box - bench, elastic 8->4->8, mailbox checks
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.log 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; tail -1 gpurun_out/bench_n$N.log | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('N',d['n_gpus'],'value',d['value'],'ms/step',d['ms_per_step'],'p50',d['config']['p50_batch_latency_ms'],'e2e',d['e2e']['value'],'sharded',(d['e2e_sharded'] or {}).get('value'),'sustained',(d['sustained'] or {}).get('value'),'clocks',d['clocks'])"; tail -3 gpurun_out/bench_n$N.err | cut -c1-300
timeout 300 python benchmarks/configs.py --config elastic --gpus $N 2> gpurun_out/cfg_elastic_n$N.err | tee gpurun_out/elastic_n$N.json | cut -c1-1200; tail -2 gpurun_out/cfg_elastic_n$N.err | cut -c1-300
timeout 600 python scripts/check_mailbox.py --gpus $N > gpurun_out/check_mailbox_n$N.log 2>&1; echo "check_mailbox rc=$?"; grep -v "^  File\|^    " gpurun_out/check_mailbox_n$N.log | tail -8 | cut -c1-400
