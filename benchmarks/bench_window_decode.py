"""Decode with a sliding window over the flash-layout cache: the HBM-roofline decode kernel with the window
(vb200_paged_attention_flash_window) against the variable-length kernel on the same batch (where such batches ran
before), and against the same decode without a window.  Bytes = the keys / values inside the windows."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / reps


def main():
    from vllm_b200 import _custom_ops as ops
    dev = "cuda"
    S, ctx, H, Hkv, D, bs = 256, 8192, 32, 8, 128, 16
    nb = S * ctx // bs + 64
    for kvd, dt in (("fp8", torch.uint8), ("auto", torch.bfloat16)):
        if kvd == "fp8":
            kc = torch.randint(0, 120, (nb, Hkv, bs, D), dtype=torch.uint8, device=dev).permute(0, 2, 1, 3)   # HND
            vc = torch.randint(0, 120, (nb, Hkv, bs, D), dtype=torch.uint8, device=dev).permute(0, 2, 1, 3)
        else:
            kc = torch.randn(nb, Hkv, bs, D, device=dev).to(dt).permute(0, 2, 1, 3)
            vc = torch.randn(nb, Hkv, bs, D, device=dev).to(dt).permute(0, 2, 1, 3)
        q = (torch.randn(S, H, D, device=dev) * 0.3).bfloat16()
        bt = torch.randperm(nb, device=dev)[:S * ctx // bs].to(torch.int32).view(S, ctx // bs)
        sl = torch.full((S,), ctx, dtype=torch.int32, device=dev)
        cu = torch.arange(S + 1, dtype=torch.int32, device=dev)
        sc = torch.full((1,), 0.02, device=dev)
        out, out2 = torch.empty_like(q), torch.empty_like(q)
        esz = 1 if kvd == "fp8" else 2
        for window in (None, 4096, 1024):
            wl = -1 if window is None else window - 1
            keys = ctx if window is None else window
            nbytes = S * keys * Hkv * D * 2 * esz
            t_dec = timed(lambda: ops.paged_attention_flash(out, q, kc, vc, D ** -0.5, bt, sl, ctx, None, kvd, sc, sc,
                                                            window_left=wl))
            t_var = timed(lambda: ops.flash_attn_varlen_paged(out2, q, kc, vc, cu, sl, bt, D ** -0.5, True, 1, kvd, sc, sc,
                                                              window_left=wl))
            print(json.dumps({"kv": kvd, "batch": S, "ctx": ctx, "window": window, "decode_kernel_us": round(t_dec, 1),
                              "decode_kernel_GBs": round(nbytes / t_dec / 1e3, 1), "varlen_kernel_us": round(t_var, 1),
                              "max_abs_diff": float((out.float() - out2.float()).abs().max())}), flush=True)
        del kc, vc


if __name__ == "__main__":
    main()
