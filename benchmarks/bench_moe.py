#!/usr/bin/env python3
"""Microbenchmark of the MoE grouped FP8 GEMM at Mixtral-8x7B decode shapes (HBM-bound on the expert
weights): our ragged-mode `cutlass_moe_mm`, our moe_align-driven `invoke_fused_moe_kernel`, and -- when
oracle/_ref/libvllm_ref_cutlass.so is present -- the reference's own sm100 CUTLASS grouped GEMM on the same inputs.
Prints one JSON line per (GEMM, batch, TP): microseconds, weight GB/s vs the measured HBM copy peak.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

FP8 = torch.float8_e4m3fn


def timeit(fn, flush, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()                       # > L2: every iteration streams the weights from HBM
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", default="16,128,512")
    ap.add_argument("--tps", default="1,8")
    args = ap.parse_args()
    from vllm_b200 import _custom_ops as ops
    from vllm_b200.fused_moe import invoke_fused_moe_kernel, moe_align_block_size
    from vllm_b200.fused_moe.fused_moe import get_block_m
    from parity_utils import load_ref_cutlass
    ref = load_ref_cutlass()
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    dev = "cuda"
    E, H, I, k = 8, 4096, 14336, 2
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    torch.manual_seed(0)
    for tp in [int(x) for x in args.tps.split(",")]:
        for name, N, K in (("w13", 2 * I // tp, H), ("w2", H, I // tp)):
            # signed e4m3 codes (sign bit random, magnitude codes 0..119): outputs centred on zero, so that the
            # difference against the reference kernel reads as a relative error
            w = (torch.randint(0, 120, (E, N, K), dtype=torch.uint8, device=dev) |
                 (torch.randint(0, 2, (E, N, K), dtype=torch.uint8, device=dev) << 7)).view(FP8)
            ws = torch.rand(E, N, device=dev) + 0.5
            for T in [int(x) for x in args.batches.split(",")]:
                ids = torch.stack([torch.randperm(E)[:k] for _ in range(T)]).to(torch.int32).to(dev)
                arows = T if name == "w13" else T * k
                a = (torch.randint(0, 120, (arows, K), dtype=torch.uint8, device=dev) |
                     (torch.randint(0, 2, (arows, K), dtype=torch.uint8, device=dev) << 7)).view(FP8)
                eo = torch.empty(E + 1, dtype=torch.int32, device=dev)
                ps1 = torch.empty(E, 3, dtype=torch.int32, device=dev)
                ps2 = torch.empty(E, 3, dtype=torch.int32, device=dev)
                a_map = torch.empty(T * k, dtype=torch.int32, device=dev)
                c_map = torch.empty(T * k, dtype=torch.int32, device=dev)
                ops.get_cutlass_moe_mm_data(ids, eo, ps1, ps2, a_map, c_map, E, I // tp, H)
                ps = ps1 if name == "w13" else ps2
                ap_ = ops.shuffle_rows(a, a_map) if name == "w13" else a
                sa = torch.rand(T * k, 1, device=dev) + 0.5
                out = torch.empty(T * k, N, dtype=torch.bfloat16, device=dev)
                st = torch.full((E,), K, dtype=torch.int64, device=dev)
                cst = torch.full((E,), N, dtype=torch.int64, device=dev)
                eo_ = eo[:-1]
                t_rag = timeit(lambda: ops.cutlass_moe_mm(out, ap_, w, sa, ws, eo_, ps, st, st, cst, True, True), flush)
                bm = get_block_m(T * k, E)
                sorted_ids, expert_ids, npost = moe_align_block_size(ids, bm, E)
                c = torch.empty(T, k, N, dtype=torch.bfloat16, device=dev)
                sa_tok = torch.rand(a.shape[0], device=dev) + 0.5
                t_al = timeit(lambda: invoke_fused_moe_kernel(a, w, c, sa_tok, ws, None, sorted_ids, expert_ids, npost,
                                                              False, k if name == "w13" else 1, bm, True, True), flush)
                rec = {"gemm": name, "tp": tp, "batch": T, "E": E, "N": N, "K": K, "weight_MB": round(E * N * K / 1e6, 1),
                       "ragged_us": round(t_rag, 1), "ragged_GBs": round(E * N * K / t_rag / 1e3, 1),
                       "ragged_frac_hbm": round(E * N * K / t_rag / 1e3 / peak, 3),
                       "aligned_us": round(t_al, 1), "aligned_frac_hbm": round(E * N * K / t_al / 1e3 / peak, 3)}
                if ref is not None and T * k > 64:
                    out2 = torch.empty_like(out)
                    try:
                        t_ref = timeit(lambda: ref.cutlass_moe_mm(out2, ap_, w, sa, ws, eo_, ps, st, st, cst, True, True),
                                       flush)
                        rec["ref_cutlass_us"] = round(t_ref, 1)
                        rec["speedup_vs_ref"] = round(t_ref / t_rag, 2)
                        d = (out.float() - out2.float()).abs()
                        rec["max_rel_diff_vs_ref"] = float((d / out2.float().abs().clamp_min(float(out2.float().abs().mean()))).max())
                        rec["rms_rel_diff_vs_ref"] = float(d.pow(2).mean().sqrt() / out2.float().pow(2).mean().sqrt())
                    except Exception as e:  # noqa: BLE001
                        rec["ref_cutlass_error"] = str(e)[:120]
                print(json.dumps(rec), flush=True)
            del w, ws
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
