#!/usr/bin/env python3
"""Per-CTA timeline of the fp8 GEMM kernel (debug hook vb200_debug_scaled_mm_trace: %globaltimer stamps at the
phase boundaries of every CTA).  Prints, per shape, the median over CTAs of each phase's duration in microseconds
and the kernel's span (first entry -> last exit)."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_b200 import _custom_ops as ops  # noqa: E402
from vllm_b200 import _lib  # noqa: E402

FP8 = torch.float8_e4m3fn
SHAPES = {"qkv": (6144, 4096), "o": (4096, 4096), "gate_up": (28672, 4096), "down": (4096, 14336)}
NAMES = ["entry", "setup", "tma0", "land0", "mma_end", "acc_rdy", "parked", "arrived", "fetched", "epi_end", "exit"]


def main():
    lib = _lib.load()
    fn = lib.vb200_debug_scaled_mm_trace
    fn.restype = None
    fn.argtypes = [ctypes.c_void_p]
    buf = torch.zeros(148 * 16, dtype=torch.int64, device="cuda")
    for M in (16, 256):
        for name, (N, K) in SHAPES.items():
            ws = [torch.randn(N, K, device="cuda").clamp(-448, 448).to(FP8) for _ in range(3)]
            x = torch.randn(M, K, device="cuda").to(FP8)
            sa = torch.rand(M, 1, device="cuda") * 0.1 + 0.01
            sb = torch.rand(1, N, device="cuda") * 0.1 + 0.01
            for i in range(3):
                ops.cutlass_scaled_mm(x, ws[i].t(), sa, sb, torch.bfloat16)
            torch.cuda.synchronize()
            buf.zero_()
            fn(buf.data_ptr())
            ops.cutlass_scaled_mm(x, ws[0].t(), sa, sb, torch.bfloat16)
            torch.cuda.synchronize()
            fn(None)
            t = buf.view(148, 16).cpu()
            used = t[:, 0] > 0
            t = t[used].double()
            t0 = t[:, 0].min()
            span = (t[:, 10].max() - t0) / 1e3
            rel = (t - t0) / 1e3
            med = rel.median(dim=0).values
            mx = rel.max(dim=0).values
            cols = [i for i in range(11) if (t[:, i] > 1e15).any()]
            print(f"{name} M={M}: ctas={int(used.sum())} span={span:.1f}us")
            print("   median us since first entry: " + " ".join(f"{NAMES[i]}={med[i]:.1f}" for i in cols))
            print("   max    us since first entry: " + " ".join(f"{NAMES[i]}={mx[i]:.1f}" for i in cols))
            lead = t[t[:, 14] > 0]   # leader CTAs: the MMA issuer's own counters
            if len(lead):
                kb = lead[:, 14].median()
                print(f"   MMA issuer (median over {len(lead)} leaders): k-blocks={kb:.0f}, whole loop "
                      f"{lead[:, 15].median():.0f} cycles = {lead[:, 15].median() / kb:.0f} per k-block (first wait included)")


if __name__ == "__main__":
    main()
