#!/bin/bash
mkdir -p gpurun_out
PT="python -m pytest -q --timeout 180 --timeout-method=thread -p no:cacheprovider"
timeout 300 $PT tests/test_zz_no_alloc.py -m gpu > gpurun_out/t_noalloc_alone.log 2>&1; echo "alone rc=$?"
timeout 900 $PT tests -m gpu -x > gpurun_out/t_all_gpu.log 2>&1; echo "all rc=$?"
tail -3 gpurun_out/t_noalloc_alone.log; tail -4 gpurun_out/t_all_gpu.log
