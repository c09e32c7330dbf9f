#!/usr/bin/env python3
"""Micro-benchmark of the fp8 w8a8 GEMM at the BASELINE cfg 3 / cfg 5 shapes: TFLOP/s vs the fp8 dense peak and
weight-stream GB/s vs the HBM peak (SURVEY.md 8d reports both), next to torch._scaled_mm (cuBLASLt) as the library
baseline.  Weights rotate through > L2-sized pools so every iteration streams them from HBM."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
FP8 = torch.float8_e4m3fn
SHAPES = {"qkv": (6144, 4096), "o": (4096, 4096), "gate_up": (28672, 4096), "down": (4096, 14336),
          "70b_qkv": (1280, 8192), "70b_o": (8192, 1024), "70b_gate_up": (7168, 8192), "70b_down": (8192, 3584)}


def timeit(fn, iters, warmup=3):
    """Kernel time without host launch overhead: `iters` calls (rotating weight pools) captured in ONE CUDA graph,
    replayed 3 times after a warm-up replay; CUDA events around the replays."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(warmup):
            fn(i)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(iters):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(3):
        g.replay()
    b.record()
    b.synchronize()
    return a.elapsed_time(b) * 1e-3 / (3 * iters)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--M", default="16,64,128,256")
    ap.add_argument("--shapes", default="qkv,o,gate_up,down,70b_qkv,70b_o,70b_gate_up,70b_down")
    ap.add_argument("--iters", type=int, default=30)
    a = ap.parse_args()
    from vllm_b200 import _custom_ops as ops
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from parity_utils import load_ref_cutlass
    ref = load_ref_cutlass()   # the reference's own sm100 CUTLASS kernel (scaled_mm_sm100_fp8_dispatch.cuh), or None
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    fp8_peak = 2 * peaks["bf16_tflops"]  # derived: fp8 dense = 2 x measured bf16 burst
    for name in a.shapes.split(","):
        N, K = SHAPES[name]
        npool = max(2, int(300e6 // (N * K)) + 1)
        ws = [torch.randn(N, K, device="cuda").clamp(-448, 448).to(FP8) for _ in range(npool)]
        for M in [int(x) for x in a.M.split(",")]:
            x = torch.randn(M, K, device="cuda").to(FP8)
            sa = torch.rand(M, 1, device="cuda") * 0.1 + 0.01
            sb = torch.rand(1, N, device="cuda") * 0.1 + 0.01
            t = timeit(lambda i: ops.cutlass_scaled_mm(x, ws[i % npool].t(), sa, sb, torch.bfloat16), a.iters)
            t2 = timeit(lambda i: torch._scaled_mm(x, ws[i % npool].t(), sa, sb, out_dtype=torch.bfloat16), a.iters)
            t3 = None
            if ref is not None:
                o3 = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
                t3 = timeit(lambda i: ref.cutlass_scaled_mm(o3, x, ws[i % npool].t(), sa, sb, None), a.iters)
            flops = 2.0 * M * N * K
            byts = M * K + N * K + 2 * M * N
            print(json.dumps({"shape": name, "M": M, "N": N, "K": K, "us": round(t * 1e6, 1),
                              "TFLOPs": round(flops / t / 1e12, 1), "frac_fp8_peak(2x_bf16_measured)": round(flops / t / 1e12 / fp8_peak, 3),
                              "GBs": round(byts / t / 1e9, 1), "frac_hbm": round(byts / t / 1e9 / peaks["hbm_gbs"], 3),
                              "cublaslt_us": round(t2 * 1e6, 1), "speedup_vs_cublaslt": round(t2 / t, 2),
                              "ref_sm100_cutlass_us": None if t3 is None else round(t3 * 1e6, 1),
                              "speedup_vs_ref_cutlass": None if t3 is None else round(t3 / t, 2)}), flush=True)


if __name__ == "__main__":
    main()
