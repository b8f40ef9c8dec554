#!/bin/bash
# 2-GPU validation after kernel changes: cluster bit-identity check, multi-GPU tests, bench N=2.
mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/check_cluster.py > gpurun_out/cluster_check.log 2>&1; echo "cluster check rc=$?"
tail -4 gpurun_out/cluster_check.log
timeout 300 python -m pytest -q --timeout 200 --timeout-method=thread -p no:cacheprovider tests/test_gpu_multi.py -m gpu -x 2>&1 | tail -3
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 > gpurun_out/bench_n2.log 2>gpurun_out/bench_n2.err
echo "bench n=2 rc=$?"; tail -1 gpurun_out/bench_n2.log | cut -c1-1400; tail -3 gpurun_out/bench_n2.err
