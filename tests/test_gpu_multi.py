"""Multi-GPU (runs at every N in {2, 4, 8} the box has): the mailbox data plane must equal the single-GPU engine bit
for bit, survive a leaver, a silent worker and a SIGKILLed worker process (scripts/check_mailbox.py), and its torchrun
front-end must do the same without collectives inside a phase (scripts/check_cluster.py)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _need(n):
    if not torch.cuda.is_available() or torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")


@pytest.mark.parametrize("n", [2, 4, 8])
def test_mailbox_elastic_equals_single_gpu(n):
    _need(n)
    p = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "check_mailbox.py"), "--gpus", str(n)], cwd=ROOT,
                       capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert p.stdout.count("mailbox check ok") == 5 and "SIGKILL" in p.stdout and "no lock-step" in p.stdout


@pytest.mark.parametrize("n", [2, 8])
def test_torchrun_cluster_equals_single_gpu(n):
    _need(n)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "scripts", "check_cluster.py")]
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert p.stdout.count("cluster check ok") == 4
