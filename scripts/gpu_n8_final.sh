#!/bin/bash
mkdir -p gpurun_out
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 scripts/check_cluster.py > gpurun_out/cluster_check_n8.log 2>&1; echo "cluster check rc=$?"
grep -c "cluster check ok" gpurun_out/cluster_check_n8.log
for n in 4 8; do
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 50 --warmup 5 > gpurun_out/bench_n$n.log 2>gpurun_out/bench_n$n.err
echo "bench n=$n rc=$?"; tail -1 gpurun_out/bench_n$n.log | cut -c1-1800; tail -2 gpurun_out/bench_n$n.err | cut -c1-300
done
