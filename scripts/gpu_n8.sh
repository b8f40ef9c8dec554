#!/bin/bash
# gpurun --gpus 8 --timeout 1500 -- 'bash scripts/gpu_n8.sh': the whole box -> gpurun_out/{bench_n*.log, bench_torch_n*.log, configs_n8.jsonl, check_*_n8.log}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n8.txt 2>&1
show() { tail -1 $1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.readline()); print(d.get('impl'),'N',d['n_gpus'],'value',d['value'],'ms/step',d['ms_per_step'],'p50',d['config'].get('p50_batch_latency_ms'),'e2e',d['e2e']['value'],'sharded',(d.get('e2e_sharded') or {}).get('value'),'sustained',(d.get('sustained') or {}).get('value'),'clocks',(d.get('clocks') or {}).get('sm_mhz'),(d.get('clocks') or {}).get('reasons'))
except Exception as e: print('unparsable', e)"; }
for N in 8 4 2; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.log 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"; show gpurun_out/bench_n$N.log
done
for N in 8 4 2; do
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 2954$N bench.py --impl torch --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_torch_n$N.log 2> gpurun_out/bench_torch_n$N.err; echo "torch N=$N rc=$?"; show gpurun_out/bench_torch_n$N.log
done
OUT=gpurun_out/configs_n8.jsonl; : > $OUT
timeout 300 python benchmarks/configs.py --config elastic --gpus 8 2> gpurun_out/cfg_elastic_n8.err | tee -a $OUT | cut -c1-700
for fps in 30 60; do timeout 300 python benchmarks/configs.py --config live --gpus 8 --fps $fps 2> gpurun_out/cfg_live_n8.err | tee -a $OUT | cut -c1-300; done
timeout 400 python benchmarks/configs.py --config multi_client --gpus 8 2> gpurun_out/cfg_multi_n8.err | tee -a $OUT | cut -c1-500
tail -2 gpurun_out/cfg_*_n8.err | cut -c1-300
timeout 500 python scripts/check_mailbox.py --gpus 8 > gpurun_out/check_mailbox_n8.log 2>&1; echo "check_mailbox rc=$?"; grep "mailbox check\|Error" gpurun_out/check_mailbox_n8.log | cut -c1-300
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1 --master-port 29561 scripts/check_cluster.py > gpurun_out/check_cluster_n8.log 2>&1; echo "check_cluster rc=$?"; grep "cluster check" gpurun_out/check_cluster_n8.log | sort -u | cut -c1-300
