#!/usr/bin/env python3
"""Micro-benchmark of paged-attention decode at the BASELINE configs: achieved HBM GB/s (algorithmic bytes,
SURVEY.md 8d) vs the measured copy peak, next to the reference's own CUDA kernel (torch.ops._ref_C, test infra).
Inputs are larger than L2 (126 MB) and block tables are random permutations, so every iteration streams from HBM.
`--flash` adds the flash-layout kernel (vb200_paged_attention_flash) on the SAME cache contents re-laid out as
[num_blocks, 16, Hkv, D] in NHD and HND order.
Usage: python benchmarks/bench_paged_attention.py [--cfg 2|3|5|all] [--iters 20] [--partition N] [--no-ref] [--flash]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

CFGS = {
    "2": dict(num_seqs=64, ctx=2048, H=32, Hkv=8, D=128, kvd="auto"),
    "3": dict(num_seqs=256, ctx=4096, H=32, Hkv=8, D=128, kvd="fp8"),
    "5": dict(num_seqs=128, ctx=8192, H=8, Hkv=1, D=128, kvd="fp8"),
    "2r": dict(num_seqs=64, ctx=2048, H=32, Hkv=8, D=128, kvd="auto", ragged=True),
    # per-GPU shards of config 3 under tensor parallelism (not in "all")
    "3tp4": dict(num_seqs=256, ctx=4096, H=8, Hkv=2, D=128, kvd="fp8"),
    "3tp8": dict(num_seqs=256, ctx=4096, H=4, Hkv=1, D=128, kvd="fp8"),
}


def peak_gbs():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


def make(cfg, dev="cuda"):
    torch.manual_seed(0)
    S, ctx, H, Hkv, D = cfg["num_seqs"], cfg["ctx"], cfg["H"], cfg["Hkv"], cfg["D"]
    bs = 16
    fp8 = cfg["kvd"] == "fp8"
    x = 16 if fp8 else 8
    blocks_per_seq = ctx // bs
    nb = S * blocks_per_seq + 1024
    if fp8:
        kc = torch.randint(0, 120, (nb, Hkv, D // x, bs, x), dtype=torch.uint8, device=dev)  # finite e4m3 codes
        vc = torch.randint(0, 120, (nb, Hkv, D, bs), dtype=torch.uint8, device=dev)
        kc |= (torch.randint(0, 2, kc.shape, dtype=torch.uint8, device=dev) << 7)
        vc |= (torch.randint(0, 2, vc.shape, dtype=torch.uint8, device=dev) << 7)
        sc = 0.01
    else:
        kc = torch.empty(nb, Hkv, D // x, bs, x, dtype=torch.bfloat16, device=dev).normal_()
        vc = torch.empty(nb, Hkv, D, bs, dtype=torch.bfloat16, device=dev).normal_()
        sc = 1.0
    q = (torch.randn(S, H, D, device=dev) * 0.3).bfloat16()
    if os.environ.get("PA_SEQ_BLOCKS") == "1":   # experiment: physically consecutive blocks per sequence
        bt = torch.arange(S * blocks_per_seq, device=dev, dtype=torch.int32).reshape(S, blocks_per_seq)
    else:
        bt = torch.randperm(nb, device=dev)[: S * blocks_per_seq].to(torch.int32).reshape(S, blocks_per_seq)
    if cfg.get("ragged"):
        sl = torch.randint(ctx // 2 + 1, ctx + 1, (S,), dtype=torch.int32, device=dev)
    else:
        sl = torch.full((S,), ctx, dtype=torch.int32, device=dev)
    scale_t = torch.full((1,), sc, device=dev)
    return q, kc, vc, bt, sl, scale_t


def algo_bytes(cfg, sl):
    ce = 1 if cfg["kvd"] == "fp8" else 2
    tot = int(sl.sum().item())
    kv = tot * cfg["Hkv"] * cfg["D"] * 2 * ce
    qo = 2 * cfg["num_seqs"] * cfg["H"] * cfg["D"] * 2
    btb = int(((sl + 15) // 16).sum().item()) * 4 + 4 * cfg["num_seqs"]
    return kv + qo + btb


def timeit(fn, iters, warmup=3):
    """(mean, best-batch mean) seconds per call.  Calls are enqueued back to back between two events, `iters` per
    batch, 3 batches: a per-call event pair with a host sync in between would add the host's launch path (ctypes +
    allocator, ~20-40 us) to every sample, which is comparable to the kernel itself at the smaller shapes.  The KV
    working set of every config is larger than L2 (126 MB), so consecutive calls still stream from HBM."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    means = []
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        b.synchronize()
        means.append(a.elapsed_time(b) * 1e-3 / iters)
    means.sort()
    return means[1], means[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", default="all")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--partition", type=int, default=0)
    ap.add_argument("--no-ref", action="store_true")
    ap.add_argument("--flash", action="store_true")
    a = ap.parse_args()
    from vllm_b200 import _custom_ops as ops
    from parity_utils import load_ref_cuda
    ref = None if a.no_ref else load_ref_cuda()
    peak, how = peak_gbs()
    names = [n for n in CFGS if "tp" not in n] if a.cfg == "all" else a.cfg.split(",")
    for name in names:
        cfg = CFGS[name]
        q, kc, vc, bt, sl, sc = make(cfg)
        out = torch.empty_like(q)
        nbytes = algo_bytes(cfg, sl)
        max_len = int(sl.max())
        scale = cfg["D"] ** -0.5

        def ours():
            ops.paged_attention_v1(out, q, kc, vc, cfg["Hkv"], scale, bt, sl, 16, max_len, None, cfg["kvd"], sc, sc,
                                   partition_size=a.partition)
        med, best = timeit(ours, a.iters)
        row = {"cfg": name, **{k: v for k, v in cfg.items()}, "algo_MB": round(nbytes / 1e6, 2),
               "ours_us": round(med * 1e6, 1), "ours_best_us": round(best * 1e6, 1),
               "ours_GBs": round(nbytes / med / 1e9, 1), "frac_of_%s_peak" % how: round(nbytes / med / 1e9 / peak, 3)}
        if ref is not None:
            out_r = torch.empty_like(q)

            def theirs():
                ref.paged_attention_v1(out_r, q, kc, vc, cfg["Hkv"], scale, bt, sl, 16, max_len, None, cfg["kvd"], sc, sc,
                                       0, 0, 0, 64, 0)
            medr, _ = timeit(theirs, max(3, a.iters // 4))
            row.update(ref_v1_us=round(medr * 1e6, 1), ref_v1_GBs=round(nbytes / medr / 1e9, 1),
                       speedup_vs_ref=round(medr / med, 2),
                       max_abs_diff_vs_ref=float((out.float() - out_r.float()).abs().max()))
        if a.flash:
            nbk, Hk, Dx, bsz, xx = kc.shape
            for layout in ("HND", "NHD"):
                kf = kc.permute(0, 3, 1, 2, 4).reshape(nbk, bsz, Hk, Dx * xx)
                vf = vc.permute(0, 3, 1, 2)
                if layout == "NHD":
                    kf, vf = kf.contiguous(), vf.contiguous()
                else:
                    kf = kf.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)
                    vf = vf.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)
                out_f = torch.empty_like(q)

                def flash():
                    ops.paged_attention_flash(out_f, q, kf, vf, scale, bt, sl, max_len, None, cfg["kvd"], sc, sc,
                                              partition_size=a.partition)
                medf, bestf = timeit(flash, a.iters)
                row.update({f"flash_{layout}_us": round(medf * 1e6, 1), f"flash_{layout}_GBs": round(nbytes / medf / 1e9, 1),
                            f"flash_{layout}_frac_of_peak": round(nbytes / medf / 1e9 / peak, 3),
                            f"flash_{layout}_max_abs_diff_vs_paged": float((out_f.float() - out.float()).abs().max())})
                del kf, vf
        print(json.dumps(row), flush=True)
        del q, kc, vc, bt, sl
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
