#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out; rm -f gpurun_out/configs_n*.log
for c in "live --fps 30" "live --fps 60" "elastic" "multi_client"; do
  timeout 150 python benchmarks/configs.py --config $c --gpus $N >> gpurun_out/configs_n$N.log 2>>gpurun_out/configs.err; echo "config $c rc=$?"
done
cat gpurun_out/configs_n$N.log; grep -vE "^\s*$" gpurun_out/configs.err | tail -8
timeout 120 python bench.py --impl torch --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_torch_n1.log 2>gpurun_out/bench_torch.err; echo "torch n1 rc=$?"; tail -1 gpurun_out/bench_torch_n1.log | cut -c1-330; tail -3 gpurun_out/bench_torch.err
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --impl torch --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_torch_n$N.log 2>gpurun_out/bench_torch_n$N.err; echo "torch n$N rc=$?"; tail -1 gpurun_out/bench_torch_n$N.log | cut -c1-330
