#!/bin/bash
# gpurun --timeout 50 -- 'bash scripts/gpu_ncu_nms.sh': ncu --set full of class_nms_kernel (second launch) after its compaction change
mkdir -p gpurun_out
timeout 22 python scripts/profile_ops.py > gpurun_out/plain_ops2.log 2>&1 && \
timeout 26 ncu --set full --clock-control none --import-source on -k "regex:class_nms_kernel" -s 1 -c 1 -f -o gpurun_out/prof_nms python scripts/profile_ops.py > gpurun_out/ncu_nms.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/plain_ops2.log; tail -2 gpurun_out/ncu_nms.log
