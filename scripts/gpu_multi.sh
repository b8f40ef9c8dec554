#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/check_cluster.py > gpurun_out/cluster_check.log 2>&1; echo "cluster check rc=$?"
tail -5 gpurun_out/cluster_check.log
for n in 1 $N; do
  if [ $n -eq 1 ]; then
    timeout 300 python bench.py --gpus 1 --steps 30 --warmup 5 > gpurun_out/bench_n1.log 2>gpurun_out/bench_n1.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --steps 30 --warmup 5 > gpurun_out/bench_n$n.log 2>gpurun_out/bench_n$n.err
  fi
  echo "bench n=$n rc=$?"; tail -1 gpurun_out/bench_n$n.log | cut -c1-900; tail -3 gpurun_out/bench_n$n.err
done
