#!/usr/bin/env python3
"""Runs a few fp8 GEMM launches of one shape (for `ncu -k regex:scaled_mm_kernel`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_b200 import _custom_ops as ops  # noqa: E402

FP8 = torch.float8_e4m3fn
shapes = {"qkv": (6144, 4096), "o": (4096, 4096), "gate_up": (28672, 4096), "down": (4096, 14336)}
M = int(sys.argv[2]) if len(sys.argv) > 2 else 256
for name in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["o"]):
    N, K = shapes[name]
    ws = [torch.randn(N, K, device="cuda").clamp(-448, 448).to(FP8) for _ in range(3)]
    x = torch.randn(M, K, device="cuda").to(FP8)
    sa = torch.rand(M, 1, device="cuda") * 0.1 + 0.01
    sb = torch.rand(1, N, device="cuda") * 0.1 + 0.01
    for i in range(6):
        ops.cutlass_scaled_mm(x, ws[i % 3].t(), sa, sb, torch.bfloat16)
    torch.cuda.synchronize()
print("ok")
