#!/usr/bin/env python3
"""Generates tests/golden/*.npz by running the REFERENCE's own CPU kernels (oracle/_ref/libvllm_ref_cpu.so,
compiled from /root/reference/csrc/cpu/{attention,cache}.cpp by oracle/build_ref.py) on seeded inputs.
Run in the build container (needs /root/reference to have been compiled): python tests/golden/make_golden.py
bf16 tensors are stored as their uint16 bit patterns."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
from parity_utils import load_ref_cpu, make_paged_kv, seed_all  # noqa: E402


def bits(t):
    return t.contiguous().view(torch.int16).numpy().view(np.uint16)


def attention_case(name, num_seqs, H, Hkv, D, seq_lens, dist, alibi, dtype=torch.bfloat16):
    ref = load_ref_cpu()
    seed_all(sum(map(ord, name)))
    bs = 16
    per = [(L + bs - 1) // bs for L in seq_lens]
    nb = sum(per)  # ragged allocation: only the blocks that are referenced are stored
    kc, vc = make_paged_kv(nb, bs, Hkv, D, dtype, dist=dist)
    q = torch.empty(num_seqs, H, D, dtype=torch.float32).uniform_(-1, 1).to(dtype) * (1.0 if dist == "uniform" else 0.5)
    perm = torch.randperm(nb).to(torch.int32)
    bt = torch.zeros(num_seqs, max(per), dtype=torch.int32)
    o = 0
    for i, n in enumerate(per):
        bt[i, :n] = perm[o:o + n]
        o += n
    sl = torch.tensor(seq_lens, dtype=torch.int32)
    scale = float(D ** -0.5)
    slopes = torch.randn(H, dtype=torch.float32) * 0.1 if alibi else None
    one = torch.tensor(1.0)
    out1 = torch.zeros_like(q)
    ref.paged_attention_v1(out1, q, kc, vc, Hkv, scale, bt, sl, bs, max(seq_lens), slopes, "auto", one, one,
                           0, 0, 0, 64, 0)
    P = (max(seq_lens) + 511) // 512
    out2 = torch.zeros_like(q)
    es = torch.zeros(num_seqs, H, P, dtype=torch.float32)
    ml = torch.zeros(num_seqs, H, P, dtype=torch.float32)
    tmp = torch.zeros(num_seqs, H, P, D, dtype=dtype)
    ref.paged_attention_v2(out2, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, bs, max(seq_lens), slopes, "auto",
                           one, one, 0, 0, 0, 64, 0)
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"), q=bits(q), key_cache=bits(kc), value_cache=bits(vc),
        block_tables=bt.numpy(), seq_lens=sl.numpy(), scale=np.float32(scale), num_kv_heads=np.int32(Hkv),
        alibi=(slopes.numpy() if alibi else np.zeros(0, np.float32)), out_v1=bits(out1), out_v2=bits(out2))
    print(name, "v1-v2 maxdiff", (out1.float() - out2.float()).abs().max().item())


def cache_case(name, T, H, D, nb):
    ref = load_ref_cpu()
    seed_all(7)
    bs, x = 16, 8
    dtype = torch.bfloat16
    key = torch.randn(T, H, D).to(dtype)
    val = torch.randn(T, H, D).to(dtype)
    kc = torch.randn(nb, H, D // x, bs, x).to(dtype)
    vc = torch.randn(nb, H, D, bs).to(dtype)
    slots = torch.randperm(nb * bs)[:T].to(torch.int64)
    kc0, vc0 = kc.clone(), vc.clone()
    one = torch.tensor(1.0)
    ref.reshape_and_cache(key, val, kc, vc, slots, "auto", one, one)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), key=bits(key), value=bits(val), key_cache_in=bits(kc0),
                        value_cache_in=bits(vc0), slot_mapping=slots.numpy(), key_cache_out=bits(kc),
                        value_cache_out=bits(vc))
    print(name, "ok")


if __name__ == "__main__":
    assert load_ref_cpu() is not None, "build oracle/_ref first: python oracle/build_ref.py cpu"
    attention_case("pa_gqa_uniform", 4, 4, 1, 128, [1, 17, 100, 530], "uniform", False)
    attention_case("pa_mha_normal_alibi", 3, 2, 2, 64, [5, 129, 600], "normal", True)
    cache_case("reshape_and_cache_bf16", 37, 2, 64, 4)
