"""Variable-length prefill over the paged flash-layout cache (csrc/paged_prefill.cu) against the library kernel the
reference calls for the same op (vllm.vllm_flash_attn.flash_attn_varlen_func from the installed vllm wheel, FA2 on
this architecture) when it imports.  Causal FLOPs counted as 4 * D * H * sum_i(visible keys of row i).

  python benchmarks/bench_prefill.py
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


def main():
    from vllm_b200 import _custom_ops as ops
    from vllm_b200 import _lib
    _lib.load()
    try:
        from vllm.vllm_flash_attn import flash_attn_varlen_func as lib_fa
    except Exception as e:   # noqa: BLE001
        lib_fa = None
        print(f"# library flash_attn_varlen_func unavailable: {e!r}", file=sys.stderr)
    dev = torch.device("cuda")
    dtype = torch.bfloat16
    H, Hkv, D, bs = 32, 8, 128, 16
    # (name, [(query_len, kv_len)], sliding_window or None, soft_cap or 0)
    cases = [("prefill 8 x 2048", [(2048, 2048)] * 8, None, 0.0), ("prefill 1 x 8192", [(8192, 8192)], None, 0.0),
             ("chunked 16 x 512 @ 4096", [(512, 4096)] * 16, None, 0.0),
             ("mixed 4 x 1024 + 64 decodes @ 2048", [(1024, 1024)] * 4 + [(1, 2048)] * 64, None, 0.0),
             ("prefill 1 x 8192, window 1024", [(8192, 8192)], 1024, 0.0),
             ("prefill 8 x 2048, soft cap 50", [(2048, 2048)] * 8, None, 50.0),
             ("chunked 16 x 512 @ 4096, window 1024 + soft cap 30", [(512, 4096)] * 16, 1024, 30.0)]
    for name, sl, window, cap in cases:
        ql = [x[0] for x in sl]
        kl = [x[1] for x in sl]
        nb = sum((k + bs - 1) // bs for k in kl) + 8
        q = torch.randn(sum(ql), H, D, device=dev, dtype=dtype)
        kc = torch.randn(nb, bs, Hkv, D, device=dev, dtype=dtype)
        vc = torch.randn(nb, bs, Hkv, D, device=dev, dtype=dtype)
        maxb = (max(kl) + bs - 1) // bs
        bt = torch.zeros(len(sl), maxb, dtype=torch.int32, device=dev)
        perm = torch.randperm(nb, device=dev).to(torch.int32)
        o = 0
        for i, k in enumerate(kl):
            n = (k + bs - 1) // bs
            bt[i, :n] = perm[o:o + n]
            o += n
        cu = torch.tensor([0] + ql, dtype=torch.int32).cumsum(0).to(torch.int32).to(dev)
        kv_lens = torch.tensor(kl, dtype=torch.int32, device=dev)
        out = torch.empty_like(q)
        scale = D ** -0.5
        vis = lambda p: p if window is None else min(p, window)   # noqa: E731  keys a query at 1-based position p sees
        flops = 4.0 * D * H * sum(sum(vis(p) for p in range(k - qq + 1, k + 1)) for qq, k in sl)
        wl = -1 if window is None else window - 1
        us = timed(lambda: ops.flash_attn_varlen_paged(out, q, kc, vc, cu, kv_lens, bt, scale, True, max(ql),
                                                       window_left=wl, softcap=cap))
        line = {"case": name, "H": H, "Hkv": Hkv, "D": D, "ours_us": round(us, 1), "ours_TFLOPs": round(flops / us / 1e6, 1)}
        if lib_fa is not None:
            try:
                out2 = torch.empty_like(q)
                f = lambda: lib_fa(q=q, k=kc, v=vc, out=out2, cu_seqlens_q=cu, max_seqlen_q=max(ql), seqused_k=kv_lens,  # noqa: E731
                                   max_seqlen_k=max(kl), softmax_scale=scale, causal=True, block_table=bt, fa_version=2,
                                   window_size=(wl, 0) if window is not None else (-1, -1), softcap=cap)
                lus = timed(f)
                line.update({"library_us": round(lus, 1), "library_TFLOPs": round(flops / lus / 1e6, 1),
                             "max_abs_diff_vs_library": float((out.float() - out2.float()).abs().max())})
            except Exception as e:   # noqa: BLE001
                line["library_error"] = repr(e)[:200]
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
