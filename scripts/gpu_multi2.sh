#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/check_cluster.py > gpurun_out/cluster_check.log 2>&1; echo "cluster check rc=$?"
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/cluster_check.log | tail -6
timeout 200 python -m pytest -q --timeout 120 --timeout-method=thread -p no:cacheprovider tests -m gpu -x 2>&1 | tail -3
