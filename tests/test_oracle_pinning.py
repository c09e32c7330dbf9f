"""Pins the oracle (CPU restatement) against (1) golden vectors produced by the REFERENCE's own CPU
kernels (tests/golden/*.npz, generator tests/golden/make_golden.py), (2) those kernels live, when
oracle/_ref/libvllm_ref_cpu.so is present, (3) the reference's documented known-answer example for
moe_align_block_size, (4) an independent numpy e4m3 codec.  Runs without a GPU."""
import os

import numpy as np
import pytest
import torch

from oracle import attention as oattn
from oracle import cache as ocache
from oracle import fp8 as ofp8
from oracle import moe as omoe
from oracle import quant as oquant
from parity_utils import load_ref_cpu, make_block_tables, make_paged_kv, seed_all

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _bf16(a):
    return torch.from_numpy(a.view(np.int16).copy()).view(torch.bfloat16)


@pytest.mark.parametrize("name", ["pa_gqa_uniform", "pa_mha_normal_alibi"])
def test_attention_oracle_matches_reference_golden(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    q, kc, vc = _bf16(g["q"]), _bf16(g["key_cache"]), _bf16(g["value_cache"])
    bt, sl = torch.from_numpy(g["block_tables"]), torch.from_numpy(g["seq_lens"])
    alibi = torch.from_numpy(g["alibi"]) if g["alibi"].size else None
    hkv, scale = int(g["num_kv_heads"]), float(g["scale"])
    out = oattn.paged_attention(q, kc, vc, hkv, scale, bt, sl, alibi)
    # the CPU reference keeps fp32 probabilities for PV, the CUDA reference (and the oracle) round them
    # to bf16 first: both are inside 1e-2 rel / 2e-2 abs of each other (north_star tolerance)
    torch.testing.assert_close(out.float(), _bf16(g["out_v1"]).float(), rtol=1e-2, atol=2e-2)
    out2, _, _, _ = oattn.paged_attention_v2(q, kc, vc, hkv, scale, bt, sl, int(sl.max()), alibi)
    torch.testing.assert_close(out2.float(), _bf16(g["out_v2"]).float(), rtol=1e-2, atol=2e-2)
    # much tighter in practice on these inputs
    assert (out.float() - _bf16(g["out_v1"]).float()).abs().max() < 8e-3


def test_reshape_and_cache_oracle_matches_reference_golden_bitexact():
    g = np.load(os.path.join(GOLD, "reshape_and_cache_bf16.npz"))
    kc, vc = _bf16(g["key_cache_in"]), _bf16(g["value_cache_in"])
    ocache.reshape_and_cache(_bf16(g["key"]), _bf16(g["value"]), kc, vc, torch.from_numpy(g["slot_mapping"]))
    assert torch.equal(kc.view(torch.int16), _bf16(g["key_cache_out"]).view(torch.int16))
    assert torch.equal(vc.view(torch.int16), _bf16(g["value_cache_out"]).view(torch.int16))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("dist", ["uniform", "normal"])
def test_attention_oracle_matches_live_reference_cpu(dtype, dist):
    ref = load_ref_cpu()
    if ref is None:
        pytest.skip("oracle/_ref/libvllm_ref_cpu.so not built")
    seed_all(3)
    S, H, Hkv, D, bs = 5, 8, 2, 128, 16
    seq_lens = [1, 16, 33, 513, 700]
    nb = S * ((max(seq_lens) + bs - 1) // bs)
    kc, vc = make_paged_kv(nb, bs, Hkv, D, dtype, dist=dist)
    q = (torch.rand(S, H, D) * 2 - 1).to(dtype)
    bt = make_block_tables(S, seq_lens, bs, nb)
    sl = torch.tensor(seq_lens, dtype=torch.int32)
    one = torch.tensor(1.0)
    out_ref = torch.zeros_like(q)
    ref.paged_attention_v1(out_ref, q, kc, vc, Hkv, D ** -0.5, bt, sl, bs, max(seq_lens), None, "auto", one, one,
                           0, 0, 0, 64, 0)
    out = oattn.paged_attention(q, kc, vc, Hkv, D ** -0.5, bt, sl)
    torch.testing.assert_close(out.float(), out_ref.float(), rtol=1e-2, atol=2e-2)


def test_reshape_and_cache_oracle_matches_live_reference_cpu():
    ref = load_ref_cpu()
    if ref is None:
        pytest.skip("oracle/_ref/libvllm_ref_cpu.so not built")
    seed_all(5)
    T, H, D, nb, bs, x = 50, 4, 128, 8, 16, 8
    key, val = torch.randn(T, H, D).bfloat16(), torch.randn(T, H, D).bfloat16()
    kc, vc = torch.zeros(nb, H, D // x, bs, x).bfloat16(), torch.zeros(nb, H, D, bs).bfloat16()
    slots = torch.randperm(nb * bs)[:T].to(torch.int64)
    kc2, vc2 = kc.clone(), vc.clone()
    one = torch.tensor(1.0)
    ref.reshape_and_cache(key, val, kc, vc, slots, "auto", one, one)
    ocache.reshape_and_cache(key, val, kc2, vc2, slots)
    assert torch.equal(kc.view(torch.int16), kc2.view(torch.int16))
    assert torch.equal(vc.view(torch.int16), vc2.view(torch.int16))


def test_moe_align_known_answer_from_reference_docstring():
    """vllm/model_executor/layers/fused_moe/moe_align_block_size.py:52-66 (experts numbered 1..4 -> E = 5)."""
    ids = torch.tensor([[2, 3, 4], [1, 2, 4], [1, 3, 4], [1, 2, 3]], dtype=torch.int32)
    s, e, n = omoe.moe_align_block_size(ids, 4, 5)
    assert s[:16].tolist() == [3, 6, 9, 12, 0, 4, 10, 12, 1, 7, 11, 12, 2, 5, 8, 12]
    assert e[:4].tolist() == [1, 2, 3, 4] and int(n) == 16
    assert (s[16:] == 12).all()


def test_fp8_codec_matches_torch_and_numpy_model():
    b = torch.arange(256, dtype=torch.uint8)
    dec_t = b.view(torch.float8_e4m3fn).float().numpy()
    dec_n = ofp8.e4m3_decode_np(b.numpy())
    fin = ~np.isnan(dec_n)
    assert np.array_equal(dec_t[fin], dec_n[fin]) and np.isnan(dec_t[~fin]).all()
    seed_all(0)
    x = torch.cat([torch.randn(4096) * 30, torch.randn(4096) * 0.01, torch.tensor([0.0, 448.0, 500.0, -1e9, 2.0 ** -9,
                                                                                  2.0 ** -10, 3 * 2.0 ** -10])])
    enc_n = ofp8.e4m3_encode_satfinite_np(x.numpy())
    enc_t = ofp8.quant_e4m3(x, 1.0).numpy()
    assert np.array_equal(enc_n, enc_t)


def test_topk_softmax_oracle_lowest_index_ties():
    g = torch.tensor([[1.0, 1.0, 0.5, 1.0], [0.0, 2.0, 2.0, -1.0]])
    w, ids, tei = omoe.topk_softmax(g, 2)
    assert ids.tolist() == [[0, 1], [1, 2]]
    assert tei.tolist() == [[0, 2], [1, 3]]  # k_idx * T + row
    assert torch.allclose(w.sum(-1), torch.softmax(g, -1).gather(1, ids.long()).sum(-1))


def test_quant_oracle_matches_reference_test_golden_formulas():
    """ref_dynamic_per_token_quant / ref_dynamic_per_tensor_fp8_quant (tests/kernels/quant_utils.py:22-98)."""
    seed_all(1)
    x = (torch.randn(7, 96) * 3).bfloat16()
    q, s = oquant.dynamic_per_token_scaled_fp8_quant(x)
    amax = x.abs().max(-1).values.float()
    s_ref = (amax / 448.0).clamp(min=1.0 / (448.0 * 512.0))[:, None]
    assert torch.equal(s, s_ref)
    q_ref = (x.float() / s_ref).clamp(-448, 448).to(torch.float8_e4m3fn)
    assert torch.equal(q.view(torch.uint8), q_ref.view(torch.uint8))
    q2, s2 = oquant.dynamic_scaled_fp8_quant(x)
    inv = 1.0 / (x.abs().max().float() / 448.0)
    assert torch.equal(q2.view(torch.uint8), (x.float() * inv).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8))


def test_moe_sum_and_experts_oracle_consistency():
    seed_all(2)
    x = torch.randn(5, 2, 64).bfloat16()
    assert torch.equal(omoe.moe_sum(x), (x[:, 0].float() + x[:, 1].float()).bfloat16())
    T, H, I, E, k = 6, 64, 32, 4, 2
    hid = (torch.randn(T, H) / 4).bfloat16()
    w1 = (torch.randn(E, 2 * I, H) / 8).bfloat16()
    w2 = (torch.randn(E, H, I) / 8).bfloat16()
    w, ids, _ = omoe.topk_softmax(torch.randn(T, E), k)
    out = omoe.fused_experts_bf16(hid, w1, w2, w, ids)
    # torch_experts-style dense reference
    ref = torch.zeros(T, H)
    for t in range(T):
        for j in range(k):
            e = int(ids[t, j])
            h1 = (hid[t].float() @ w1[e].float().t()).bfloat16()
            h2 = oquant.silu_and_mul(h1[None])[0]
            ref[t] += w[t, j] * (h2.float() @ w2[e].float().t())
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2)


def test_cutlass_moe_data_oracle_known_answer():
    """oracle.moe.get_cutlass_moe_mm_data on the routing example of the reference docstring
    (fused_moe/moe_align_block_size.py:52-66, experts numbered 1..4): offsets = exclusive prefix of the per-expert
    counts, a_map/c_map are inverse permutations, swapped problem sizes for <= 64 routed rows (moe_data.cu:9,30-44),
    and the grouped-GEMM oracle equals per-expert dense matmuls of the dequantised operands."""
    import torch
    from oracle import moe as omoe
    ids = torch.tensor([[2, 3, 4], [1, 2, 4], [1, 3, 4], [1, 2, 3]], dtype=torch.int32)
    eo, ps1, ps2, a_map, c_map = omoe.get_cutlass_moe_mm_data(ids, 5, 8, 16)
    assert eo.tolist() == [0, 0, 3, 6, 9, 12]
    assert ps1.tolist() == [[16, 0, 16]] + [[16, 3, 16]] * 4 and ps2.tolist() == [[16, 0, 8]] + [[16, 3, 8]] * 4
    assert c_map.tolist() == [3, 6, 9, 0, 4, 10, 1, 7, 11, 2, 5, 8]
    assert a_map[c_map.long()].tolist() == [i // 3 for i in range(12)]
    big = torch.randint(0, 5, (40, 2), dtype=torch.int32)          # 80 rows > 64: not swapped
    assert omoe.get_cutlass_moe_mm_data(big, 5, 8, 16)[1][:, 1].tolist() == [16] * 5
    # grouped GEMM oracle == dequantised per-expert matmul
    torch.manual_seed(0)
    E, N, K, rows = 3, 32, 64, 10
    a = torch.randn(rows, K).to(torch.float8_e4m3fn)
    b = torch.randn(E, N, K).to(torch.float8_e4m3fn)
    sa, sb = torch.rand(rows) + 0.5, torch.rand(E, N) + 0.5
    off, cnt = torch.tensor([0, 4, 4]), torch.tensor([4, 0, 6])
    got = omoe.cutlass_moe_mm(a, b, sa, sb, off, cnt, torch.float32)
    for e in range(E):
        o, m = int(off[e]), int(cnt[e])
        want = (a[o:o + m].float() * sa[o:o + m, None]) @ (b[e].float() * sb[e][:, None]).t()
        torch.testing.assert_close(got[o:o + m], want, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("name", ["flash_varlen_ref_case", "flash_varlen_prefill_gqa7", "flash_varlen_window256",
                                  "flash_varlen_softcap10_window", "flash_varlen_softcap50"])
def test_flash_varlen_oracle_matches_reference_test_golden(name):
    """oracle.attention.flash_attn_varlen_paged == the reference test suite's own `ref_paged_attn`
    (tests/kernels/attention/test_flash_attn.py:25-77) run in the build container on seeded inputs
    (tests/golden/make_flash_varlen_golden.py).  The reference golden scales the query and runs both einsums in bf16
    (:46,59,72); the oracle keeps the logits and the PV product in fp32 (what the kernels do), so the two agree within
    the tolerance the reference test itself asserts against that golden (atol 1.5e-2, rtol 1e-2, :316-318)."""
    g = np.load(os.path.join(GOLD, name + ".npz"))
    q, kc, vc = _bf16(g["q"]), _bf16(g["key_cache"]), _bf16(g["value_cache"])
    out = oattn.flash_attn_varlen_paged(q, kc, vc, g["query_lens"].tolist(), g["kv_lens"].tolist(),
                                        torch.from_numpy(g["block_tables"]), float(g["scale"]),
                                        sliding_window=int(g["sliding_window"]) if int(g["sliding_window"]) >= 0 else None,
                                        soft_cap=float(g["soft_cap"]) or None)
    want = _bf16(g["out"])
    assert out.shape == want.shape
    torch.testing.assert_close(out.float(), want.float(), rtol=1e-2, atol=1.5e-2)
