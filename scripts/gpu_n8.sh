#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/check_cluster.py > gpurun_out/cluster_check_n$N.log 2>&1; echo "cluster check rc=$?"
grep "cluster check ok" gpurun_out/cluster_check_n$N.log; tail -3 gpurun_out/cluster_check_n$N.log | cut -c1-300
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n$N.log 2>gpurun_out/bench_n$N.err
echo "bench n=$N rc=$?"; tail -1 gpurun_out/bench_n$N.log | cut -c1-1300; tail -3 gpurun_out/bench_n$N.err | cut -c1-300
timeout 150 python benchmarks/configs.py --config elastic --gpus $N > gpurun_out/elastic_n$N.log 2>gpurun_out/elastic_n$N.err; echo "elastic rc=$?"; cat gpurun_out/elastic_n$N.log; tail -3 gpurun_out/elastic_n$N.err | cut -c1-300
