#!/bin/bash
mkdir -p gpurun_out
timeout 200 python scripts/profile_step.py 3 64 0 mbx > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 400 --csv --log-file gpurun_out/launches_warm.csv python scripts/profile_step.py 3 64 0 mbx > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/plain.log; tail -2 gpurun_out/ncu.log
