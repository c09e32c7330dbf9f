#!/usr/bin/env python3
"""Runs torch._scaled_mm (cuBLASLt fp8) and, when built, the reference's sm100 CUTLASS cutlass_scaled_mm once per
(shape, M) so that an ncu launch list (gpu__time_duration, grid, cluster, block, smem) shows which tile / cluster
configuration the libraries pick at decode batch sizes.  Not a benchmark: run under
  ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size,launch__cluster_size,\
launch__shared_mem_per_block_dynamic --csv python benchmarks/probe_cublaslt_kernels.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
FP8 = torch.float8_e4m3fn
SHAPES = {"qkv": (6144, 4096), "o": (4096, 4096), "gate_up": (28672, 4096), "down": (4096, 14336),
          "70b_qkv": (1280, 8192), "70b_o": (8192, 1024), "70b_gate_up": (7168, 8192), "70b_down": (8192, 3584)}


def main():
    from parity_utils import load_ref_cutlass
    ref = load_ref_cutlass()
    for name, (N, K) in SHAPES.items():
        w = torch.randn(N, K, device="cuda").clamp(-448, 448).to(FP8)
        for M in (16, 128, 256):
            x = torch.randn(M, K, device="cuda").to(FP8)
            sa = torch.rand(M, 1, device="cuda") * 0.1 + 0.01
            sb = torch.rand(1, N, device="cuda") * 0.1 + 0.01
            torch.cuda.nvtx.range_push(f"{name}_M{M}")
            torch._scaled_mm(x, w.t(), sa, sb, out_dtype=torch.bfloat16)
            if ref is not None:
                o = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
                ref.cutlass_scaled_mm(o, x, w.t(), sa, sb, None)
            torch.cuda.nvtx.range_pop()
            torch.cuda.synchronize()
            print(name, M, flush=True)


if __name__ == "__main__":
    main()
