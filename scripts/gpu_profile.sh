#!/bin/bash
mkdir -p gpurun_out
PT="python -m pytest -q --timeout 180 --timeout-method=thread -p no:cacheprovider"
timeout 500 $PT tests/test_gpu_engine.py > gpurun_out/t_engine.log 2>&1; echo "engine rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/t_engine.log | head
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"
tail -1 gpurun_out/bench.log | cut -c1-400
timeout 200 python scripts/profile_step.py 3 64 > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python scripts/profile_step.py 3 64 > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/plain.log; tail -3 gpurun_out/ncu.log
