#!/bin/bash
# gpurun --gpus 2 --timeout 1200 -- 'bash scripts/gpu_mbx2.sh [N]': elastic mailbox path + torchrun front-end + bench at N GPUs
N=${1:-2}
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
timeout 600 python scripts/check_mailbox.py --gpus $N > gpurun_out/check_mailbox_n$N.log 2>&1; echo "check_mailbox rc=$?"; grep -v "^  File\|^    " gpurun_out/check_mailbox_n$N.log | tail -12
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29533 scripts/check_cluster.py > gpurun_out/check_cluster_n$N.log 2>&1; echo "check_cluster rc=$?"; grep "cluster check" gpurun_out/check_cluster_n$N.log | sort -u
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.log 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; tail -1 gpurun_out/bench_n$N.log | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('N',d['n_gpus'],'value',d['value'],'ms/step',d['ms_per_step'],'p50',d['config']['p50_batch_latency_ms'],'e2e',d['e2e']['value'],'sharded',(d['e2e_sharded'] or {}).get('value'),'sustained',(d['sustained'] or {}).get('value'),'clocks',d['clocks'])"; tail -3 gpurun_out/bench_n$N.err
