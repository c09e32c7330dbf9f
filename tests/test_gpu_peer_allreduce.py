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


def test_peer_allreduce_stress_two_ranks():
    """>= 10 000 unsynchronised calls with random per-rank delays, alternating sizes / kernels, eager + graph
    (tests/multi_gpu/stress_peer_allreduce.py; reference: csrc/custom_all_reduce_test.cu)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29534", os.path.join(ROOT, "tests", "multi_gpu", "stress_peer_allreduce.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "peer all-reduce stress: OK" in r.stdout


def test_tp2_logits_match_tp1():
    """TP-2 decode == TP-1 decode on the same (sharded) weights, NCCL and fused peer all-reduce
    (tests/multi_gpu/check_tp_parity.py)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29535", os.path.join(ROOT, "tests", "multi_gpu", "check_tp_parity.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "tp parity check: OK" in r.stdout
