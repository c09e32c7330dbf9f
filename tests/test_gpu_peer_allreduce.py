"""GPU (>= 2 devices): the NVLink peer-memory all-reduce fused with residual add + RMSNorm (+ fp8 quant) against NCCL
+ the unfused ops (tests/multi_gpu/check_peer_allreduce.py under torchrun).  Skipped on a 1-GPU box."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_peer_allreduce_two_ranks():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "multi_gpu", "check_peer_allreduce.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "peer all-reduce check: OK" in r.stdout
