#!/bin/bash
mkdir -p gpurun_out
PT="python -m pytest -q --timeout 300 --timeout-method=thread -p no:cacheprovider"
timeout 500 $PT tests -m gpu -x > gpurun_out/t_all_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/t_all_gpu.log
timeout 100 python -c "import __graft_entry__ as g; g.build(); g.smoke()" 2>&1 | tail -1
timeout 200 python bench.py > gpurun_out/bench_n1.log 2>gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -1 gpurun_out/bench_n1.log | cut -c1-300
