"""GPU: the torch.ops registration (vllm_b200/torch_shim.cpp) under the reference's schemas gives the same bits as
the ctypes host path, and passes torch.library.opcheck like the reference ops do
(/root/reference/tests/kernels/utils.py:1184-1198)."""
import pytest
import torch

from parity_utils import make_block_tables, make_paged_kv, seed_all

pytestmark = pytest.mark.gpu
DEV = "cuda"
FP8 = torch.float8_e4m3fn


@pytest.fixture(scope="module")
def tops():
    from vllm_b200.torch_bindings import load_torch_ops
    load_torch_ops()
    return torch.ops


def test_attention_and_cache_ops(tops):
    from vllm_b200 import _custom_ops as ops
    seed_all(0)
    S, H, Hkv, D, bs = 5, 16, 4, 128, 16
    seq_lens = [3, 100, 513, 700, 64]
    nb = S * 44 + 2
    for kvd, ksc in (("auto", 1.0), ("fp8", 0.04)):
        kc, vc = make_paged_kv(nb, bs, Hkv, D, torch.bfloat16, kvd, dist="normal", k_scale=ksc, v_scale=ksc)
        kc, vc = kc.to(DEV), vc.to(DEV)
        bt = make_block_tables(S, seq_lens, bs, nb).to(DEV)
        sl = torch.tensor(seq_lens, dtype=torch.int32, device=DEV)
        ks = torch.tensor(ksc, device=DEV)
        key = torch.randn(S, Hkv, D, device=DEV).bfloat16()
        val = torch.randn(S, Hkv, D, device=DEV).bfloat16()
        slots = torch.tensor([int(bt[i, (L - 1) // bs]) * bs + (L - 1) % bs for i, L in enumerate(seq_lens)], device=DEV)
        kc2, vc2 = kc.clone(), vc.clone()
        ops.reshape_and_cache(key, val, kc, vc, slots, kvd, ks, ks)
        tops._b200_C_cache_ops.reshape_and_cache(key, val, kc2, vc2, slots, kvd, ks, ks)
        assert torch.equal(kc.view(torch.uint8), kc2.view(torch.uint8)) and torch.equal(vc.view(torch.uint8), vc2.view(torch.uint8))
        q = (torch.randn(S, H, D, device=DEV) * 0.5).bfloat16()
        o1, o2 = torch.empty_like(q), torch.empty_like(q)
        ops.paged_attention_v1(o1, q, kc, vc, Hkv, D ** -0.5, bt, sl, bs, max(seq_lens), None, kvd, ks, ks)
        tops._b200_C.paged_attention_v1(o2, q, kc, vc, Hkv, D ** -0.5, bt, sl, bs, max(seq_lens), None, kvd, ks, ks,
                                        0, 0, 0, 64, 0)
        assert torch.equal(o1, o2)
        P = (max(seq_lens) + 511) // 512
        es = torch.empty(S, H, P, device=DEV)
        ml = torch.empty_like(es)
        tmp = torch.empty(S, H, P, D, dtype=q.dtype, device=DEV)
        o3 = torch.empty_like(q)
        tops._b200_C.paged_attention_v2(o3, es, ml, tmp, q, kc, vc, Hkv, D ** -0.5, bt, sl, bs, max(seq_lens), None,
                                        kvd, ks, ks, 0, 0, 0, 64, 0)
        torch.testing.assert_close(o3.float(), o1.float(), rtol=1e-2, atol=2e-2)
        torch.library.opcheck(tops._b200_C.paged_attention_v1,
                              (o2, q, kc, vc, Hkv, D ** -0.5, bt, sl, bs, max(seq_lens), None, kvd, ks, ks, 0, 0, 0, 64, 0),
                              test_utils=("test_schema", "test_faketensor"))
    fk = torch.zeros(nb, bs, Hkv, D, dtype=torch.bfloat16, device=DEV)
    fv, fk2, fv2 = torch.zeros_like(fk), torch.zeros_like(fk), torch.zeros_like(fk)
    one = torch.ones(1, device=DEV)
    ops.reshape_and_cache_flash(key, val, fk, fv, slots, "auto", one, one)
    tops._b200_C_cache_ops.reshape_and_cache_flash(key, val, fk2, fv2, slots, "auto", one, one)
    assert torch.equal(fk, fk2) and torch.equal(fv, fv2)


def test_linear_quant_norm_ops(tops):
    from vllm_b200 import _custom_ops as ops
    seed_all(1)
    M, K, N = 64, 512, 384
    x = torch.randn(M, K, device=DEV).bfloat16()
    q1, s1 = ops.scaled_fp8_quant(x, use_per_token_if_dynamic=True)
    q2, s2 = torch.empty_like(q1), torch.empty_like(s1)
    tops._b200_C.dynamic_per_token_scaled_fp8_quant(q2, x, s2, None)
    assert torch.equal(q1.view(torch.uint8), q2.view(torch.uint8)) and torch.equal(s1, s2)
    w = (torch.randn(N, K, device=DEV) / 4).clamp(-448, 448).to(FP8)
    ws = torch.rand(1, N, device=DEV) * 0.1 + 0.01
    o1 = ops.cutlass_scaled_mm(q1, w.t(), s1, ws, torch.bfloat16)
    o2 = torch.empty_like(o1)
    tops._b200_C.cutlass_scaled_mm(o2, q1, w.t(), s1, ws, None)
    assert torch.equal(o1, o2)
    assert tops._b200_C.cutlass_scaled_mm_supports_fp8(100)
    # (test_schema is skipped for this op: torch's schema checker multiplies its inputs, which fp8 tensors lack)
    torch.library.opcheck(tops._b200_C.cutlass_scaled_mm, (o2, q1, w.t(), s1, ws, None),
                          test_utils=("test_faketensor",))
    wn = torch.randn(K, device=DEV).bfloat16()
    n1, n2 = torch.empty_like(x), torch.empty_like(x)
    ops.rms_norm(n1, x, wn, 1e-5)
    tops._b200_C.rms_norm(n2, x, wn, 1e-5)
    assert torch.equal(n1, n2)
    a, b, r1, r2 = x.clone(), x.clone(), n1.clone(), n1.clone()
    ops.fused_add_rms_norm(a, r1, wn, 1e-5)
    tops._b200_C.fused_add_rms_norm(b, r2, wn, 1e-5)
    assert torch.equal(a, b) and torch.equal(r1, r2)
    act1, act2 = torch.empty(M, K // 2, device=DEV).bfloat16(), torch.empty(M, K // 2, device=DEV).bfloat16()
    ops.silu_and_mul(act1, x)
    tops._b200_C.silu_and_mul(act2, x)
    assert torch.equal(act1, act2)


def test_moe_ops(tops):
    from vllm_b200 import _custom_ops as ops
    seed_all(2)
    T, E, k, blk = 77, 8, 2, 16
    g = torch.randn(T, E, device=DEV)
    w1, i1, t1 = torch.empty(T, k, device=DEV), torch.empty(T, k, dtype=torch.int32, device=DEV), torch.empty(T, k, dtype=torch.int32, device=DEV)
    w2, i2, t2 = torch.empty_like(w1), torch.empty_like(i1), torch.empty_like(t1)
    ops.topk_softmax(w1, i1, t1, g)
    tops._b200_moe_C.topk_softmax(w2, i2, t2, g)
    assert torch.equal(w1, w2) and torch.equal(i1, i2) and torch.equal(t1, t2)
    cap = T * k + E * (blk - 1)
    s1 = torch.empty(cap, dtype=torch.int32, device=DEV)
    e1 = torch.empty((cap + blk - 1) // blk, dtype=torch.int32, device=DEV)
    n1 = torch.empty(1, dtype=torch.int32, device=DEV)
    s2, e2, n2 = torch.empty_like(s1), torch.empty_like(e1), torch.empty_like(n1)
    ops.moe_align_block_size(i1, E, blk, s1, e1, n1)
    tops._b200_moe_C.moe_align_block_size(i1, E, blk, s2, e2, n2)
    assert torch.equal(s1, s2) and torch.equal(e1, e2) and torch.equal(n1, n2)
    x = torch.randn(T, k, 256, device=DEV).bfloat16()
    o1, o2 = torch.empty(T, 256, device=DEV).bfloat16(), torch.empty(T, 256, device=DEV).bfloat16()
    ops.moe_sum(x, o1)
    tops._b200_moe_C.moe_sum(x, o2)
    assert torch.equal(o1, o2)


def test_cutlass_moe_ops(tops):
    """get_cutlass_moe_mm_data / shuffle_rows / cutlass_moe_mm through torch.ops == the ctypes host path, bit for bit."""
    from vllm_b200 import _custom_ops as ops
    seed_all(3)
    T, k, E, N, K = 70, 2, 8, 256, 256
    ids = torch.stack([torch.randperm(E)[:k] for _ in range(T)]).to(torch.int32).to(DEV)

    def bufs():
        return (torch.empty(E + 1, dtype=torch.int32, device=DEV), torch.empty(E, 3, dtype=torch.int32, device=DEV),
                torch.empty(E, 3, dtype=torch.int32, device=DEV), torch.empty(T * k, dtype=torch.int32, device=DEV),
                torch.empty(T * k, dtype=torch.int32, device=DEV))
    b1, b2 = bufs(), bufs()
    ops.get_cutlass_moe_mm_data(ids, *b1, E, N // 2, K)
    tops._b200_C.get_cutlass_moe_mm_data(ids, *b2, E, N // 2, K, None)
    for x, y in zip(b1, b2):
        assert torch.equal(x, y)
    eo, ps1, _, a_map, c_map = b1
    a = (torch.randn(T, K, device=DEV)).to(FP8)
    ap = ops.shuffle_rows(a, a_map)
    ap2 = torch.empty_like(ap)
    tops._b200_moe_C.shuffle_rows(a, a_map, ap2)
    assert torch.equal(ap.view(torch.uint8), ap2.view(torch.uint8))
    w = (torch.randn(E, N, K, device=DEV) / 4).to(FP8)
    sa = torch.rand(T * k, 1, device=DEV) + 0.5
    sb = torch.rand(E, N, device=DEV) + 0.5
    st = torch.full((E,), K, dtype=torch.int64, device=DEV)
    cst = torch.full((E,), N, dtype=torch.int64, device=DEV)
    o1 = torch.zeros(T * k, N, dtype=torch.bfloat16, device=DEV)
    o2 = torch.zeros_like(o1)
    ops.cutlass_moe_mm(o1, ap, w, sa, sb, eo[:-1], ps1, st, st, cst, True, True)
    tops._b200_C.cutlass_moe_mm(o2, ap, w, sa, sb, eo[:-1], ps1, st, st, cst, True, True)
    assert torch.equal(o1, o2) and bool(o1.abs().sum() > 0)
    assert tops._b200_C.cutlass_group_gemm_supported(100)
    torch.library.opcheck(tops._b200_C.get_cutlass_moe_mm_data.default, (ids, *bufs(), E, N // 2, K, None),
                          test_utils=("test_schema", "test_faketensor"))
