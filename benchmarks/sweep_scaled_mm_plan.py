#!/usr/bin/env python3
"""Planner check of the fp8 GEMM: for every (shape, M) time EVERY (token tiles, K slices) plan the kernel can run
(debug hook vb200_debug_scaled_mm_plan) next to the planner's own choice -- the cost model in scaled_mm.cu
(plan_cost) is fitted to, and judged by, this table.  Times: 30 back-to-back launches over rotating weight pools
(> L2) inside one CUDA graph, PDL on."""
import argparse
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
from bench_scaled_mm import SHAPES, timeit  # noqa: E402

FP8 = torch.float8_e4m3fn


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--M", default="16,64,128,256")
    ap.add_argument("--shapes", default="qkv,o,gate_up,down,70b_qkv,70b_o,70b_gate_up,70b_down")
    a = ap.parse_args()
    from vllm_b200 import _custom_ops as ops
    from vllm_b200 import _lib
    lib = _lib.load()
    hook = lib.vb200_debug_scaled_mm_plan
    hook.restype, hook.argtypes = None, [ctypes.c_int] * 5 + [ctypes.c_void_p]
    for name in a.shapes.split(","):
        N, K = SHAPES[name]
        npool = max(2, int(300e6 // (N * K)) + 1)
        ws = [torch.randn(N, K, device="cuda").clamp(-448, 448).to(FP8) for _ in range(npool)]
        for M in [int(x) for x in a.M.split(",")]:
            x = torch.randn(M, K, device="cuda").to(FP8)
            sa = torch.rand(M, 1, device="cuda") * 0.1 + 0.01
            sb = torch.rand(1, N, device="cuda") * 0.1 + 0.01
            plan = (ctypes.c_int * 8)()
            hook(0, 0, M, N, K, plan)
            chosen = (plan[1], plan[2])
            res = {}
            for mt in ([1, 2] if 128 < M <= 256 else [max(1, (M + 255) // 256)]):
                for S in range(1, 9):
                    if S > 1 and (K // 128) // S < 2:
                        break
                    hook(S, mt, 0, 0, 0, None)
                    try:
                        t = timeit(lambda i: ops.cutlass_scaled_mm(x, ws[i % npool].t(), sa, sb, torch.bfloat16), 30)
                        res[f"mt{mt}_s{S}"] = round(t * 1e6, 1)
                    except Exception as e:  # noqa: BLE001  (cluster shape not placeable, empty slice, ...)
                        res[f"mt{mt}_s{S}"] = None
                        torch.cuda.synchronize()
            hook(0, 0, 0, 0, 0, None)
            ok = {k: v for k, v in res.items() if v is not None}
            best = min(ok, key=ok.get) if ok else None
            ck = f"mt{chosen[0]}_s{chosen[1]}"
            print(json.dumps({"shape": name, "M": M, "N": N, "K": K, "planner": ck, "planner_us": res.get(ck),
                              "best": best, "best_us": ok.get(best), "all_us": res}), flush=True)


if __name__ == "__main__":
    main()
