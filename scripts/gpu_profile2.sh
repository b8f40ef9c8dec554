#!/bin/bash
# gpurun --timeout 400 -- 'bash scripts/gpu_profile2.sh': launch list of one step on the final tree + ncu --set full of the implicit conv
mkdir -p gpurun_out
timeout 100 python scripts/profile_step.py 3 64 0 mbx > gpurun_out/plain.log 2>&1 && \
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 400 --csv --log-file gpurun_out/launches_warm.csv python scripts/profile_step.py 3 64 0 mbx > gpurun_out/ncu.log 2>&1
echo "launch list rc=$?"; tail -2 gpurun_out/plain.log; tail -2 gpurun_out/ncu.log
timeout 100 python scripts/prof_implicit.py > gpurun_out/plain_implicit.log 2>&1 && \
timeout 200 ncu --set full --clock-control none --import-source on -k "regex:gemm_f16_kernel" -s 1 -c 1 -f -o gpurun_out/prof_implicit python scripts/prof_implicit.py > gpurun_out/ncu_implicit.log 2>&1
echo "ncu full rc=$?"; tail -1 gpurun_out/plain_implicit.log; tail -2 gpurun_out/ncu_implicit.log; ls -la gpurun_out/prof_implicit.ncu-rep
