#!/bin/bash
# gpurun --timeout 900 -- 'bash scripts/gpu_mbx1.sh': mailbox protocol on one GPU, then the whole GPU suite
mkdir -p gpurun_out
PT="python -m pytest -q --timeout 300 --timeout-method=thread -p no:cacheprovider"
timeout 600 $PT tests/test_gpu_mailbox.py -m gpu -x 2>&1 | tail -30 | tee gpurun_out/t_mailbox.log
