#!/bin/bash
# quick 2-GPU iteration: mailbox tests, elastic check, bench at two lane counts
mkdir -p gpurun_out
python -m pytest -q -x --timeout 600 -p no:cacheprovider tests/test_gpu_mailbox.py -m gpu 2>&1 | tail -3
timeout 600 python scripts/check_mailbox.py --gpus 2 > gpurun_out/check_mailbox_n2.log 2>&1; echo "check_mailbox rc=$?"; grep "mailbox check\|Error" gpurun_out/check_mailbox_n2.log | cut -c1-200
for c in 4 6; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 20 --warmup 5 --concurrency $c > gpurun_out/bench_n2_c$c.log 2> gpurun_out/bench_n2.err; echo "bench rc=$?"; tail -1 gpurun_out/bench_n2_c$c.log | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('conc',d['config']['batches_in_flight'],'N',d['n_gpus'],'value',d['value'],'ms/step',d['ms_per_step'],'p50',d['config']['p50_batch_latency_ms'],'e2e',d['e2e']['value'],'sharded',(d['e2e_sharded'] or {}).get('value'),'sustained',(d['sustained'] or {}).get('value'))"; tail -3 gpurun_out/bench_n2.err | grep -v "^\*\|OMP" | cut -c1-300
done
