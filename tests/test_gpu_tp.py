"""TP on real GPUs (needs >= 2 devices; the single-GPU round-end run skips it): launches
tools/tp_check.py under torchrun - one-shot NVLink all-reduce vs NCCL, TP = 2 stack vs TP = 1."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_tp2_on_gpus():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "tools", "tp_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0 and "TP_CHECK PASS" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]
