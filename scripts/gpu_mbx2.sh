#!/bin/bash
# gpurun --gpus 2 --timeout 1200 -- 'bash scripts/gpu_mbx2.sh': elastic mailbox path + torchrun front-end + bench at N=2
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
timeout 600 python scripts/check_mailbox.py --gpus 2 > gpurun_out/check_mailbox_n2.log 2>&1; echo "check_mailbox rc=$?"; tail -12 gpurun_out/check_mailbox_n2.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29533 scripts/check_cluster.py > gpurun_out/check_cluster_n2.log 2>&1; echo "check_cluster rc=$?"; grep "cluster check" gpurun_out/check_cluster_n2.log; tail -5 gpurun_out/check_cluster_n2.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_n2.log 2> gpurun_out/bench_n2.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_n2.log; tail -5 gpurun_out/bench_n2.err
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.log 2> gpurun_out/bench_n1.err; echo "bench1 rc=$?"; tail -2 gpurun_out/bench_n1.log; tail -3 gpurun_out/bench_n1.err
