#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
PT="python -m pytest -q --timeout 300 --timeout-method=thread -p no:cacheprovider"
timeout 600 $PT tests -m gpu -x > gpurun_out/t_all_gpu.log 2>&1; echo "all gpu tests rc=$?"; tail -3 gpurun_out/t_all_gpu.log
for c in "live --fps 30" "live --fps 60" "elastic" "multi_client"; do
  timeout 400 python benchmarks/configs.py --config $c --gpus $N >> gpurun_out/configs_n$N.log 2>gpurun_out/configs.err; echo "config $c rc=$?"
done
timeout 300 python benchmarks/configs.py --config live --fps 60 --gpus 1 >> gpurun_out/configs_n1.log 2>>gpurun_out/configs.err
cat gpurun_out/configs_n$N.log gpurun_out/configs_n1.log; tail -5 gpurun_out/configs.err
timeout 300 python bench.py --impl torch --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_torch_n1.log 2>gpurun_out/bench_torch.err; echo "torch n1 rc=$?"; tail -1 gpurun_out/bench_torch_n1.log | cut -c1-600; tail -3 gpurun_out/bench_torch.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --impl torch --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_torch_n$N.log 2>gpurun_out/bench_torch_n$N.err; echo "torch n$N rc=$?"; tail -1 gpurun_out/bench_torch_n$N.log | cut -c1-600
timeout 500 python bench.py --impl reference --gpus 1 --steps 4 --warmup 0 > gpurun_out/bench_ref_n1.log 2>gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -1 gpurun_out/bench_ref_n1.log | cut -c1-700; nproc
