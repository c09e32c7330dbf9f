"""In-graph cost of the small per-token kernels between the contractions of one decode layer (csrc/layer_glue.cu,
cache_kernels.cu, decode_loop.cu).  ncu's per-launch times are cold-cache and serialised; what the decode step pays is
the kernel inside a CUDA graph with programmatic dependent launch, back to back with its neighbours.  Each kernel is
captured `--layers` times over rotating buffers (so that nothing is L2-hot that would not be in the real step) and the
graph is replayed; prints one JSON line per kernel with the per-launch average and the bytes it has to move.

  python benchmarks/bench_layer_glue.py --tokens 256 --hidden 4096 --inter 14336 --heads 32 --kv-heads 8
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def graph_time(fn, layers, reps=5):
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for l in range(layers):
            fn(l)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for l in range(layers):
            fn(l)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * layers)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, default=256)
    ap.add_argument("--hidden", type=int, default=4096)
    ap.add_argument("--inter", type=int, default=14336)
    ap.add_argument("--heads", type=int, default=32)
    ap.add_argument("--kv-heads", type=int, default=8)
    ap.add_argument("--head-dim", type=int, default=128)
    ap.add_argument("--vocab", type=int, default=128256)
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--ctx", type=int, default=4096)
    a = ap.parse_args()
    from vllm_b200 import _custom_ops as ops
    from vllm_b200 import _lib
    _lib.load()
    dev = torch.device("cuda")
    T, H, I, L = a.tokens, a.hidden, a.inter, a.layers
    bf = torch.bfloat16
    FP8 = torch.float8_e4m3fn
    out = []

    def report(name, us, byts):
        line = {"kernel": name, "tokens": T, "us_per_launch": round(us, 2), "bytes": byts,
                "GBs": round(byts / us / 1e3, 1)}
        out.append(line)
        print(json.dumps(line), flush=True)

    # launch floor: the smallest kernel of the library
    ids = torch.zeros(T, dtype=torch.int64, device=dev)
    nxt = torch.zeros(T, 1, dtype=torch.int64, device=dev)
    pos = torch.full((T,), a.ctx - 1, dtype=torch.int64, device=dev)
    sl = torch.full((T,), 8, dtype=torch.int32, device=dev)
    slots = torch.zeros(T, dtype=torch.int64, device=dev)
    bt = torch.zeros(T, 64, dtype=torch.int32, device=dev)
    report("advance_step (launch floor)", graph_time(
        lambda l: ops.advance_step_flashattn(T, T, 16, ids, nxt, pos, sl, slots, bt), L), 40 * T)
    sl.fill_(8)

    # rms_norm + residual + per-token fp8 quant
    hid = [torch.randn(T, H, device=dev, dtype=bf) for _ in range(L)]
    res = [torch.randn(T, H, device=dev, dtype=bf) for _ in range(L)]
    w = torch.ones(H, device=dev, dtype=bf)
    report("rms_norm_dynamic_per_token_quant(+residual)", graph_time(
        lambda l: ops.rms_norm_dynamic_per_token_quant(hid[l], w, 1e-5, FP8, None, res[l]), L), T * H * 7 + 4 * T)
    report("fused_add_rms_norm", graph_time(lambda l: ops.fused_add_rms_norm(hid[l], res[l], w, 1e-5), L), T * H * 8)
    del hid, res

    # RoPE + reshape_and_cache (fp8 cache)
    nh, nkv, D = a.heads, a.kv_heads, a.head_dim
    qkv = [torch.randn(T, (nh + 2 * nkv) * D, device=dev, dtype=bf) for _ in range(L)]
    from vllm_b200.decode_step import make_cos_sin_cache
    cos_sin = make_cos_sin_cache(D, 8192, 500000.0, bf, dev)
    nb = T * 4 + 8
    kc = [torch.zeros(nb, nkv, D // 16, 16, 16, dtype=torch.uint8, device=dev) for _ in range(L)]
    vc = [torch.zeros(nb, nkv, D, 16, dtype=torch.uint8, device=dev) for _ in range(L)]
    slot = (torch.randperm(nb, device=dev)[:T].to(torch.int64) * 16 + 5)
    one = torch.ones(1, device=dev)

    def rope(l):
        q = qkv[l][:, :nh * D]
        k = qkv[l][:, nh * D:(nh + nkv) * D]
        v = qkv[l][:, (nh + nkv) * D:]
        ops.rotary_embedding_reshape_and_cache(pos, q, k, v, D, cos_sin, True, kc[l], vc[l], slot, "fp8", one, one)
    report("rotary_embedding_reshape_and_cache (fp8 cache)", graph_time(rope, L),
           T * (nh + nkv) * D * 4 + T * nkv * D * 2 + T * nkv * D * 2)
    del qkv, kc, vc

    # silu_and_mul + per-token fp8 quant
    gu = [torch.randn(T, 2 * I, device=dev, dtype=bf) for _ in range(L)]
    report("silu_and_mul_dynamic_per_token_quant", graph_time(
        lambda l: ops.silu_and_mul_dynamic_per_token_quant(gu[l]), L), T * I * 5 + 4 * T)
    act = torch.empty(T, I, device=dev, dtype=bf)
    report("silu_and_mul", graph_time(lambda l: ops.silu_and_mul(act, gu[l]), L), T * I * 6)
    del gu

    # per-token fp8 quant of a [T, hidden] activation
    x = [torch.randn(T, H, device=dev, dtype=bf) for _ in range(L)]
    report("scaled_fp8_quant per token [T, hidden]", graph_time(
        lambda l: ops.scaled_fp8_quant(x[l], use_per_token_if_dynamic=True), L), T * H * 3 + 4 * T)
    del x

    # greedy sampling on bf16 logits vs the torch path it replaced
    logits = [torch.randn(T, a.vocab, device=dev, dtype=bf) for _ in range(4)]
    report("greedy_argmax (bf16 logits)", graph_time(lambda l: ops.greedy_argmax(logits[l % 4]), 8), T * a.vocab * 2)
    report("torch: logits.float().max(dim=-1)", graph_time(lambda l: logits[l % 4].float().max(dim=-1), 8),
           T * a.vocab * 2)


if __name__ == "__main__":
    main()
