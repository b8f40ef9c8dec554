#!/bin/bash
mkdir -p gpurun_out
timeout 200 python scripts/profile_ops.py > gpurun_out/plain_ops.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:annotate_kernel|softmax_kernel|class_nms_kernel|merge_deliver_kernel|dw3x3_tma_kernel|gemm_f16_kernel|preproc_conv0|resize_u8_kernel" -s 12 -c 12 -f -o gpurun_out/prof_ops python scripts/profile_ops.py > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/plain_ops.log; tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/*.ncu-rep
