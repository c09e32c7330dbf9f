#!/usr/bin/env python3
"""fp8 dense tensor peak as this box delivers it (BASELINE.md section 4 asks for it): cuBLASLt fp8 (torch._scaled_mm)
and this repo's cutlass_scaled_mm at 8192^3, best of 10 and sustained over ~2 s, in TFLOP/s."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_b200 import _custom_ops as ops  # noqa: E402

FP8 = torch.float8_e4m3fn


def bench(fn, flops):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize()
        best = min(best, a.elapsed_time(b) * 1e-3)
    n, t0 = 0, time.perf_counter()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    while time.perf_counter() - t0 < 2.0:
        for _ in range(20):
            fn()
        n += 20
        torch.cuda.synchronize()
    b.record(); b.synchronize()
    return round(flops / best / 1e12, 1), round(flops * n / (a.elapsed_time(b) * 1e-3) / 1e12, 1)


def main():
    N = 8192
    x = torch.randn(N, N, device="cuda").to(FP8)
    w = torch.randn(N, N, device="cuda").to(FP8)          # [N, K], K contiguous
    sa = torch.ones(N, 1, device="cuda")
    sb = torch.ones(1, N, device="cuda")
    one = torch.ones((), device="cuda")
    flops = 2.0 * N ** 3
    lt = bench(lambda: torch._scaled_mm(x, w.t(), scale_a=one, scale_b=one, out_dtype=torch.bfloat16), flops)
    ours = bench(lambda: ops.cutlass_scaled_mm(x, w.t(), sa, sb, torch.bfloat16), flops)
    print(json.dumps({"shape": "8192^3 fp8 e4m3 -> bf16", "cublaslt_burst_sustained_TFLOPs": lt,
                      "vllm_b200_scaled_mm_burst_sustained_TFLOPs": ours,
                      "note": "cuBLASLt with per-tensor scales; ours with per-token x per-channel scales"}))


if __name__ == "__main__":
    main()
