"""GPU parity of the flash-layout (NHD / HND) paged decode: vb200_paged_attention_flash through
the C ABI vs the oracle (which reads the same cache contents re-laid out in the reference's paged layout) and vs this
package's paged-layout kernel, plus the V1-style backend (reshape_and_cache_flash -> flash_attn_varlen_func) and its
bit-compatibility with reshape_and_cache_flash.  Reference call sites: vllm/v1/attention/backends/flash_attn.py:79-102
(cache shape / stride order), :448-519 (forward).  Tolerance: 1e-2 rel / 2e-2 abs (fp8: vs the dequantised cache)."""
import pytest
import torch

from oracle import attention as oattn
from parity_utils import make_block_tables, make_paged_kv, seed_all

pytestmark = pytest.mark.gpu
DEV = "cuda"
RTOL, ATOL = 1e-2, 2e-2


def _to_flash(kc, vc, layout):
    """Paged-layout caches (K [nb, H, D/x, bs, x], V [nb, H, D, bs]) -> the same contents as flash-layout
    `[nb, bs, H, D]` views, physically NHD or HND."""
    nb, H, Dx, bs, x = kc.shape
    k = kc.permute(0, 3, 1, 2, 4).reshape(nb, bs, H, Dx * x)
    v = vc.permute(0, 3, 1, 2)
    if layout == "NHD":
        return k.contiguous(), v.contiguous()
    kh = k.permute(0, 2, 1, 3).contiguous()   # physical [nb, H, bs, D]
    vh = v.permute(0, 2, 1, 3).contiguous()
    return kh.permute(0, 2, 1, 3), vh.permute(0, 2, 1, 3)


CASES = [
    # S, H, Hkv, D, seq_lens
    (7, 32, 8, 128, [1, 15, 16, 17, 511, 513, 1300]),
    (3, 8, 8, 64, [33, 700, 2]),
    (2, 40, 1, 128, [77, 640]),
    (2, 6, 2, 128, [50, 530]),
    (2, 4, 2, 256, [40, 300]),
    (2, 4, 2, 192, [40, 300]),
    (5, 8, 1, 128, [2048, 2047, 1, 900, 3000]),
]


@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("kvd", ["auto", "fp8"])
@pytest.mark.parametrize("case", CASES, ids=lambda c: f"S{c[0]}H{c[1]}kv{c[2]}D{c[3]}")
def test_paged_attention_flash(layout, dtype, kvd, case):
    from vllm_b200 import _custom_ops as ops
    S, H, Hkv, D, seq_lens = case
    if kvd == "fp8" and D not in (128, 256):
        pytest.skip("fp8 flash layout: head sizes 128 / 256")
    bs = 16
    seed_all(S * 1000 + D)
    nb = S * ((max(seq_lens) + bs - 1) // bs) + 4
    ksc, vsc = (0.043, 0.029) if kvd == "fp8" else (1.0, 1.0)
    kc, vc = make_paged_kv(nb, bs, Hkv, D, dtype, kvd, dist="normal", k_scale=ksc, v_scale=vsc)
    qkv = (torch.randn(S, H + 2 * Hkv, D) * 0.5).to(dtype)
    q = qkv[:, :H]
    bt = make_block_tables(S, seq_lens, bs, nb)
    sl = torch.tensor(seq_lens, dtype=torch.int32)
    scale = float(D ** -0.5)
    want = oattn.paged_attention(q, kc, vc, Hkv, scale, bt, sl, None, kvd, ksc, vsc)
    kf, vf = _to_flash(kc.to(DEV), vc.to(DEV), layout)
    assert kf.shape == (nb, bs, Hkv, D)
    qd = qkv.to(DEV)[:, :H]
    ks, vs = torch.tensor([ksc], device=DEV), torch.tensor([vsc], device=DEV)
    out = torch.full((S, H, D), float("nan"), dtype=dtype, device=DEV)
    ops.paged_attention_flash(out, qd, kf, vf, scale, bt.to(DEV), sl.to(DEV), max(seq_lens), None, kvd, ks, vs)
    torch.cuda.synchronize()
    torch.testing.assert_close(out.cpu().float(), want.float(), rtol=RTOL, atol=ATOL)
    # and against the paged-layout kernel on the same bytes: same arithmetic up to the summation order inside a tile
    out_p = torch.empty_like(out)
    ops.paged_attention_v1(out_p, qd, kc.to(DEV), vc.to(DEV), Hkv, scale, bt.to(DEV), sl.to(DEV), bs, max(seq_lens), None,
                           kvd, ks, vs)
    torch.testing.assert_close(out.float(), out_p.float(), rtol=4e-3, atol=4e-3)
    # never-split (one CTA per sequence and head group) and a forced 256-token split give the same answer
    for part in (-1, 256):
        o2 = torch.empty_like(out)
        ops.paged_attention_flash(o2, qd, kf, vf, scale, bt.to(DEV), sl.to(DEV), max(seq_lens), None, kvd, ks, vs,
                                  partition_size=part)
        torch.testing.assert_close(o2.cpu().float(), want.float(), rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("layout", ["NHD", "HND"])
def test_paged_attention_flash_alibi_and_graph(layout):
    from vllm_b200 import _custom_ops as ops
    seed_all(11)
    S, H, Hkv, D, bs = 3, 8, 4, 128, 16
    seq_lens = [19, 555, 1024]
    nb = S * 64 + 2
    kc, vc = make_paged_kv(nb, bs, Hkv, D, torch.bfloat16, dist="uniform")
    q = ((torch.rand(S, H, D) * 2 - 1) * D ** -0.5).bfloat16()
    bt = make_block_tables(S, seq_lens, bs, nb)
    sl = torch.tensor(seq_lens, dtype=torch.int32)
    alibi = torch.randn(H) * 0.05
    want = oattn.paged_attention(q, kc, vc, Hkv, 0.2, bt, sl, alibi)
    kf, vf = _to_flash(kc.to(DEV), vc.to(DEV), layout)
    qd, btd, sld, ald = q.to(DEV), bt.to(DEV), sl.to(DEV), alibi.to(DEV)
    out = torch.empty_like(qd)
    ops.paged_attention_flash(out, qd, kf, vf, 0.2, btd, sld, max(seq_lens), ald, "auto", None, None)
    torch.cuda.synchronize()
    torch.testing.assert_close(out.cpu().float(), want.float(), rtol=RTOL, atol=ATOL)
    # capturable: the tensor maps are kernel parameters, the split-K scratch comes from the caching allocator
    g = torch.cuda.CUDAGraph()
    out2 = torch.empty_like(qd)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        ops.paged_attention_flash(out2, qd, kf, vf, 0.2, btd, sld, max(seq_lens), ald, "auto", None, None)
    torch.cuda.current_stream().wait_stream(side)
    with torch.cuda.graph(g):
        ops.paged_attention_flash(out2, qd, kf, vf, 0.2, btd, sld, max(seq_lens), ald, "auto", None, None)
    out2.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out2, out)


@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("kvd", ["auto", "fp8"])
def test_flash_backend_decode_step(layout, kvd):
    """B200FlashAttentionImpl.forward = reshape_and_cache_flash of the new token + decode over the V1 cache, for a
    batch whose caches were filled token by token through the same op (so the kernel reads exactly what
    reshape_and_cache_flash wrote, in NHD and in HND)."""
    from vllm_b200.attention.flash_backend import (B200FlashAttentionBackend, B200FlashAttentionImpl,
                                                   B200FlashAttentionMetadata, set_kv_cache_layout)
    seed_all(3)
    set_kv_cache_layout(layout)
    try:
        S, H, Hkv, D, bs = 4, 16, 4, 128, 16
        seq_lens = [5, 64, 333, 1000]
        nb = S * ((max(seq_lens) + bs - 1) // bs) + 3
        dtype = torch.bfloat16
        cache_dtype = torch.uint8 if kvd == "fp8" else dtype
        kv_cache = B200FlashAttentionBackend.allocate_kv_cache(nb, bs, Hkv, D, cache_dtype, DEV)
        assert kv_cache.shape == (2, nb, bs, Hkv, D)
        if layout == "HND":
            assert kv_cache.stride(3) > kv_cache.stride(2)   # heads outside tokens
        bt = make_block_tables(S, seq_lens, bs, nb)

        class Layer:
            _k_scale = torch.tensor([0.05], device=DEV)
            _v_scale = torch.tensor([0.04], device=DEV)
        impl = B200FlashAttentionImpl(H, D, D ** -0.5, Hkv, kv_cache_dtype=kvd)
        # fill: every earlier token of every sequence goes through reshape_and_cache_flash; keep host copies
        keys = [torch.randn(L, Hkv, D).to(dtype) for L in seq_lens]
        vals = [torch.randn(L, Hkv, D).to(dtype) for L in seq_lens]
        from vllm_b200 import _custom_ops as ops
        kcache, vcache = kv_cache.unbind(0)
        for i, L in enumerate(seq_lens):
            if L > 1:
                slots = torch.tensor([int(bt[i, t // bs]) * bs + t % bs for t in range(L - 1)], dtype=torch.int64)
                ops.reshape_and_cache_flash(keys[i][:L - 1].to(DEV), vals[i][:L - 1].to(DEV), kcache, vcache,
                                            slots.to(DEV), kvd, Layer._k_scale, Layer._v_scale)
        # the decode step: last token of every sequence
        q = (torch.randn(S, H, D) * 0.5).to(dtype)
        k_new = torch.stack([keys[i][-1] for i in range(S)])
        v_new = torch.stack([vals[i][-1] for i in range(S)])
        slots = torch.tensor([int(bt[i, (L - 1) // bs]) * bs + (L - 1) % bs for i, L in enumerate(seq_lens)],
                             dtype=torch.int64)
        meta = B200FlashAttentionMetadata(num_actual_tokens=S, max_query_len=1,
                                          query_start_loc=torch.arange(S + 1, dtype=torch.int32, device=DEV),
                                          max_seq_len=max(seq_lens), seq_lens=torch.tensor(seq_lens, dtype=torch.int32,
                                                                                           device=DEV),
                                          block_table=bt.to(DEV), slot_mapping=slots.to(DEV))
        out = torch.empty(S, H, D, dtype=dtype, device=DEV)
        impl.forward(Layer, q.to(DEV), k_new.to(DEV), v_new.to(DEV), kv_cache, meta, output=out)
        torch.cuda.synchronize()
        # golden: plain attention over the (fp8-round-tripped) keys / values
        from oracle.fp8 import dequant_e4m3, quant_e4m3
        G = H // Hkv
        for i, L in enumerate(seq_lens):
            k, v = keys[i], vals[i]
            if kvd == "fp8":
                k = dequant_e4m3(quant_e4m3(k, 0.05), 0.05, dtype)
                v = dequant_e4m3(quant_e4m3(v, 0.04), 0.04, dtype)
            kk = k.float().repeat_interleave(G, dim=1)
            vv = v.float().repeat_interleave(G, dim=1)
            logits = torch.einsum("hd,lhd->hl", q[i].float(), kk) * D ** -0.5
            p = torch.softmax(logits, -1).to(dtype).float()
            want = torch.einsum("hl,lhd->hd", p, vv)
            torch.testing.assert_close(out[i].cpu().float(), want, rtol=RTOL, atol=ATOL)
    finally:
        set_kv_cache_layout("HND")


def test_flash_attn_varlen_func_rejects_what_is_not_built():
    from vllm_b200.attention.flash_backend import flash_attn_varlen_func
    q = torch.zeros(2, 8, 128, dtype=torch.bfloat16, device=DEV)
    kc = torch.zeros(4, 16, 2, 128, dtype=torch.bfloat16, device=DEV)
    bt = torch.zeros(2, 2, dtype=torch.int32, device=DEV)
    sl = torch.ones(2, dtype=torch.int32, device=DEV)
    cu = torch.arange(3, dtype=torch.int32, device=DEV)
    with pytest.raises(NotImplementedError):                                       # a window needs the causal mask
        flash_attn_varlen_func(q, kc, kc, 1, cu, 16, sl, block_table=bt, window_size=(128, 0), causal=False)
    with pytest.raises(NotImplementedError):                                       # no paged cache
        flash_attn_varlen_func(q, kc, kc, 1, cu, 16, sl)
    with pytest.raises(RuntimeError):
        from vllm_b200 import _custom_ops as ops
        ops.paged_attention_flash(torch.empty_like(q), q, torch.zeros(4, 32, 2, 128, dtype=torch.bfloat16, device=DEV),
                                  torch.zeros(4, 32, 2, 128, dtype=torch.bfloat16, device=DEV), 0.1, bt, sl, 16, None,
                                  "auto", None, None)                               # block size 32


# ---- prefill / chunked prefill / mixed batches over the paged flash-layout cache (csrc/paged_prefill.cu) ----------
# (query_len, kv_len) per sequence; the first two lists are the reference test's own
# (/root/reference/tests/kernels/attention/test_flash_attn.py:192-194), tolerances from there (:316-318).
VARLEN_CASES = [
    [(1, 1328), (5, 18), (129, 463)],
    [(1, 523), (1, 37), (1, 2011)],
    [(64, 64), (65, 65), (200, 200), (1, 1)],            # pure prefill, tile edges
    [(128, 1024), (63, 127), (300, 4000), (17, 17)],     # chunked prefill deep into a long context
    [(700, 700)],
]


def _varlen_inputs(seq_lens, H, Hkv, D, bs, dtype, layout, nb=2048):
    query_lens = [x[0] for x in seq_lens]
    kv_lens = [x[1] for x in seq_lens]
    q = torch.randn(sum(query_lens), H, D).to(dtype)
    kc = torch.randn(nb, bs, Hkv, D).to(dtype)
    vc = torch.randn(nb, bs, Hkv, D).to(dtype)
    maxb = (max(kv_lens) + bs - 1) // bs
    bt = torch.randint(0, nb, (len(seq_lens), maxb), dtype=torch.int32)
    cu = torch.tensor([0] + query_lens, dtype=torch.int32).cumsum(0).to(torch.int32)
    kd, vd = kc.to(DEV), vc.to(DEV)
    if layout == "HND":   # physical [nb, H, bs, D], logical view [nb, bs, H, D]
        kd = kd.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)
        vd = vd.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)
    return q, kc, vc, kd, vd, bt, cu, query_lens, kv_lens


@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("block_size", [16, 32])
@pytest.mark.parametrize("heads", [(4, 4), (8, 2), (16, 2), (14, 2)])
@pytest.mark.parametrize("head_size", [64, 128, 256])
@pytest.mark.parametrize("seq_lens", VARLEN_CASES, ids=lambda c: f"n{len(c)}q{max(x[0] for x in c)}")
def test_flash_attn_varlen_paged(seq_lens, head_size, heads, block_size, dtype, layout):
    from vllm_b200 import _custom_ops as ops
    H, Hkv = heads
    seed_all(len(seq_lens) * 100 + head_size + H)
    q, kc, vc, kd, vd, bt, cu, query_lens, kv_lens = _varlen_inputs(seq_lens, H, Hkv, head_size, block_size, dtype, layout)
    scale = float(head_size ** -0.5)
    out = torch.full_like(q, float("nan"), device=DEV)
    kvl = torch.tensor(kv_lens, dtype=torch.int32, device=DEV)
    ops.flash_attn_varlen_paged(out, q.to(DEV), kd, vd, cu.to(DEV), kvl, bt.to(DEV), scale, True, max(query_lens))
    torch.cuda.synchronize()
    want = oattn.flash_attn_varlen_paged(q, kc, vc, query_lens, kv_lens, bt, scale)
    torch.testing.assert_close(out.cpu().float(), want.float(), atol=1.5e-2, rtol=1e-2)
    if max(query_lens) <= 64:   # the same batch on the 128-row tiles (max_seqlen_q unknown)
        out2 = torch.full_like(q, float("nan"), device=DEV)
        ops.flash_attn_varlen_paged(out2, q.to(DEV), kd, vd, cu.to(DEV), kvl, bt.to(DEV), scale, True)
        torch.testing.assert_close(out2.cpu().float(), want.float(), atol=1.5e-2, rtol=1e-2)


def test_flash_attn_varlen_paged_non_causal_and_backend_prefill():
    """causal=False attends every cached key; the V1-style backend routes a mixed prefill + decode batch through
    flash_attn_varlen_func (reshape_and_cache_flash first, then the variable-length kernel)."""
    from vllm_b200 import _custom_ops as ops
    from vllm_b200.attention.flash_backend import (B200FlashAttentionBackend, B200FlashAttentionImpl,
                                                   B200FlashAttentionMetadata)
    seed_all(5)
    H, Hkv, D, bs, dtype = 8, 2, 128, 16, torch.bfloat16
    seq_lens = [(70, 300), (9, 9)]
    q, kc, vc, kd, vd, bt, cu, query_lens, kv_lens = _varlen_inputs(seq_lens, H, Hkv, D, bs, dtype, "NHD", nb=256)
    scale = float(D ** -0.5)
    out = torch.empty_like(q, device=DEV)
    ops.flash_attn_varlen_paged(out, q.to(DEV), kd, vd, cu.to(DEV), torch.tensor(kv_lens, dtype=torch.int32, device=DEV),
                                bt.to(DEV), scale, False)
    want = oattn.flash_attn_varlen_paged(q, kc, vc, query_lens, kv_lens, bt, scale, causal=False)
    torch.testing.assert_close(out.cpu().float(), want.float(), atol=1.5e-2, rtol=1e-2)

    # backend: sequence 0 prefills 40 new tokens on top of 100 cached ones, sequence 1 decodes one token at 57
    nb = 64
    kv = B200FlashAttentionBackend.allocate_kv_cache(nb, bs, Hkv, D, dtype, DEV)
    kv.copy_(torch.randn(2, nb, bs, Hkv, D).to(dtype))
    lens = [(40, 140), (1, 57)]
    T = sum(x[0] for x in lens)
    perm = torch.randperm(nb)
    bt2 = torch.stack([perm[:16], perm[16:32]]).to(torch.int32)
    qkv = torch.randn(T, H + 2 * Hkv, D).to(dtype)
    slots = []
    for s, (ql, kl) in enumerate(lens):
        for pos in range(kl - ql, kl):
            slots.append(int(bt2[s, pos // bs]) * bs + pos % bs)
    meta = B200FlashAttentionMetadata(
        num_actual_tokens=T, max_query_len=40, query_start_loc=torch.tensor([0, 40, 41], dtype=torch.int32, device=DEV),
        max_seq_len=140, seq_lens=torch.tensor([140, 57], dtype=torch.int32, device=DEV), block_table=bt2.to(DEV),
        slot_mapping=torch.tensor(slots, dtype=torch.int64, device=DEV))
    impl = B200FlashAttentionImpl(H, D, scale, Hkv)

    class _Layer:
        _k_scale = torch.ones(1, device=DEV)
        _v_scale = torch.ones(1, device=DEV)
    qd = qkv.to(DEV)
    out2 = torch.empty(T, H, D, dtype=dtype, device=DEV)
    impl.forward(_Layer(), qd[:, :H], qd[:, H:H + Hkv], qd[:, H + Hkv:], kv, meta, output=out2)
    torch.cuda.synchronize()
    kc2, vc2 = kv[0].cpu(), kv[1].cpu()   # after the write: what the oracle gathers from
    want2 = oattn.flash_attn_varlen_paged(qkv[:, :H].contiguous(), kc2, vc2, [40, 1], [140, 57], bt2, scale)
    torch.testing.assert_close(out2.cpu().float(), want2.float(), atol=1.5e-2, rtol=1e-2)


@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("heads", [(8, 2), (4, 4), (14, 2)])
@pytest.mark.parametrize("head_size", [64, 128, 256])
@pytest.mark.parametrize("seq_lens", VARLEN_CASES[:1] + VARLEN_CASES[2:4], ids=lambda c: f"n{len(c)}q{max(x[0] for x in c)}")
def test_flash_attn_varlen_paged_fp8_cache(seq_lens, head_size, heads, dtype, layout):
    """fp8 (e4m3) cache: the oracle attends over the DEQUANTISED cache (bytes -> half -> x scale -> query dtype, as
    quant_utils.cuh:348-359), the kernel folds the per-tensor scales into the logits / output."""
    from oracle.fp8 import dequant_e4m3
    from vllm_b200 import _custom_ops as ops
    H, Hkv = heads
    bs, nb = 16, 1024
    seed_all(len(seq_lens) * 7 + head_size + H)
    query_lens = [x[0] for x in seq_lens]
    kv_lens = [x[1] for x in seq_lens]
    q = torch.randn(sum(query_lens), H, head_size).to(dtype)
    ksc, vsc = 0.043, 0.029
    k8 = (torch.randn(nb, bs, Hkv, head_size) / ksc).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
    v8 = (torch.randn(nb, bs, Hkv, head_size) / vsc).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
    maxb = (max(kv_lens) + bs - 1) // bs
    bt = torch.randint(0, nb, (len(seq_lens), maxb), dtype=torch.int32)
    cu = torch.tensor([0] + query_lens, dtype=torch.int32).cumsum(0).to(torch.int32)
    kd, vd = k8.to(DEV), v8.to(DEV)
    if layout == "HND":
        kd = kd.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)
        vd = vd.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)
    scale = float(head_size ** -0.5)
    out = torch.full_like(q, float("nan"), device=DEV)
    ops.flash_attn_varlen_paged(out, q.to(DEV), kd, vd, cu.to(DEV), torch.tensor(kv_lens, dtype=torch.int32, device=DEV),
                                bt.to(DEV), scale, True, max(query_lens), "fp8", torch.tensor([ksc], device=DEV),
                                torch.tensor([vsc], device=DEV))
    torch.cuda.synchronize()
    kc = dequant_e4m3(k8, ksc, dtype)
    vc = dequant_e4m3(v8, vsc, dtype)
    want = oattn.flash_attn_varlen_paged(q, kc, vc, query_lens, kv_lens, bt, scale)
    torch.testing.assert_close(out.cpu().float(), want.float(), atol=2e-2, rtol=1e-2)


# ---- sliding window / logits soft-capping (the reference test's grid: sliding_window in {None, 256}, soft_cap in
# {None, 10, 50}, test_flash_attn.py:86-88,201-203; golden :62-69) -------------------------------------------------
@pytest.mark.parametrize("dtype_layout", [(torch.float16, "NHD"), (torch.bfloat16, "HND")], ids=["f16-NHD", "bf16-HND"])
@pytest.mark.parametrize("soft_cap", [None, 10.0, 50.0])
@pytest.mark.parametrize("sliding_window", [None, 256, 70])
@pytest.mark.parametrize("heads", [(8, 2), (14, 2)])
@pytest.mark.parametrize("head_size", [64, 128, 256])
@pytest.mark.parametrize("seq_lens", VARLEN_CASES[:4], ids=lambda c: f"n{len(c)}q{max(x[0] for x in c)}")
def test_flash_attn_varlen_paged_window_softcap(seq_lens, head_size, heads, sliding_window, soft_cap, dtype_layout):
    if sliding_window is None and soft_cap is None:
        pytest.skip("the plain causal case is test_flash_attn_varlen_paged")
    dtype, layout = dtype_layout
    from vllm_b200.attention.flash_backend import flash_attn_varlen_func
    H, Hkv = heads
    seed_all(len(seq_lens) * 31 + head_size + H + (sliding_window or 0))
    q, kc, vc, kd, vd, bt, cu, query_lens, kv_lens = _varlen_inputs(seq_lens, H, Hkv, head_size, 16, dtype, layout)
    scale = float(head_size ** -0.5)
    kvl = torch.tensor(kv_lens, dtype=torch.int32, device=DEV)
    window = (sliding_window - 1, 0) if sliding_window is not None else (-1, -1)      # as test_flash_attn.py:120,237
    out = torch.full_like(q, float("nan"), device=DEV)
    flash_attn_varlen_func(q=q.to(DEV), k=kd, v=vd, out=out, cu_seqlens_q=cu.to(DEV), seqused_k=kvl,
                           max_seqlen_q=max(query_lens), max_seqlen_k=max(kv_lens), softmax_scale=scale, causal=True,
                           window_size=window, block_table=bt.to(DEV), softcap=soft_cap if soft_cap is not None else 0)
    torch.cuda.synchronize()
    want = oattn.flash_attn_varlen_paged(q, kc, vc, query_lens, kv_lens, bt, scale, sliding_window=sliding_window,
                                         soft_cap=soft_cap)
    torch.testing.assert_close(out.cpu().float(), want.float(), atol=1.5e-2, rtol=1e-2)


@pytest.mark.parametrize("name", ["flash_varlen_window256", "flash_varlen_softcap10_window", "flash_varlen_softcap50"])
def test_flash_attn_varlen_window_softcap_vs_reference_test_golden(name):
    """The kernel against outputs of the reference test suite's own `ref_paged_attn` (tests/golden/
    make_flash_varlen_golden.py), tolerance as test_flash_attn.py:316-318."""
    import os
    import numpy as np
    from vllm_b200 import _custom_ops as ops
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
    bf = lambda a: torch.from_numpy(a).view(torch.bfloat16)  # noqa: E731
    q, kc, vc, want = bf(g["q"]), bf(g["key_cache"]), bf(g["value_cache"]), bf(g["out"])
    ql, kl = g["query_lens"].tolist(), g["kv_lens"].tolist()
    cu = torch.tensor([0] + ql, dtype=torch.int32).cumsum(0).to(torch.int32)
    sw, cap = int(g["sliding_window"]), float(g["soft_cap"])
    out = torch.full_like(q, float("nan"), device=DEV)
    ops.flash_attn_varlen_paged(out, q.to(DEV), kc.to(DEV), vc.to(DEV), cu.to(DEV),
                                torch.tensor(kl, dtype=torch.int32, device=DEV),
                                torch.from_numpy(g["block_tables"]).to(DEV), float(g["scale"]), True, max(ql),
                                window_left=sw - 1 if sw >= 0 else -1, softcap=cap)
    torch.testing.assert_close(out.cpu().float(), want.float(), atol=1.5e-2, rtol=1e-2)


def test_flash_backend_decode_with_window_and_softcap_fp8():
    """B200FlashAttentionImpl with sliding_window / logits_soft_cap (flash_attn.py:366-374) on a decode batch over an
    fp8 cache: routed to the variable-length kernel."""
    from oracle.fp8 import dequant_e4m3
    from vllm_b200.attention.flash_backend import B200FlashAttentionImpl, B200FlashAttentionMetadata
    seed_all(77)
    H, Hkv, D, bs, nb, dtype = 8, 2, 128, 16, 128, torch.bfloat16
    kv_lens = [300, 17, 129]
    S = len(kv_lens)
    ksc, vsc = 0.05, 0.03
    kv8 = (torch.randn(2, nb, bs, Hkv, D) / 0.05).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
    perm = torch.randperm(nb)
    bt = perm[:S * 20].view(S, 20).to(torch.int32)
    q = torch.randn(S, H, D).to(dtype)
    meta = B200FlashAttentionMetadata(
        num_actual_tokens=S, max_query_len=1, query_start_loc=torch.arange(S + 1, dtype=torch.int32, device=DEV),
        max_seq_len=max(kv_lens), seq_lens=torch.tensor(kv_lens, dtype=torch.int32, device=DEV), block_table=bt.to(DEV),
        slot_mapping=torch.zeros(S, dtype=torch.int64, device=DEV))
    impl = B200FlashAttentionImpl(H, D, D ** -0.5, Hkv, sliding_window=64, kv_cache_dtype="fp8", logits_soft_cap=30.0,
                                  kv_sharing_target_layer_name="previous layer wrote the cache")

    class _Layer:
        _k_scale = torch.tensor([ksc], device=DEV)
        _v_scale = torch.tensor([vsc], device=DEV)
    out = torch.empty(S, H, D, dtype=dtype, device=DEV)
    impl.forward(_Layer(), q.to(DEV), None, None, kv8.to(DEV), meta, output=out)
    torch.cuda.synchronize()
    kc, vc = dequant_e4m3(kv8[0], ksc, dtype), dequant_e4m3(kv8[1], vsc, dtype)
    want = oattn.flash_attn_varlen_paged(q, kc, vc, [1] * S, kv_lens, bt, D ** -0.5, sliding_window=64, soft_cap=30.0)
    torch.testing.assert_close(out.cpu().float(), want.float(), atol=2e-2, rtol=1e-2)


# ---- cascade attention (shared prefix + per-query suffix merged through the log-sum-exps), modelled on
# /root/reference/tests/kernels/attention/test_cascade_flash_attn.py:70-200 (cases :70-76, tolerance :200) -------------
CASCADE_CASES = [
    ([(129, 871), (18, 280), (37, 988), (1023, 2304), (1, 257)], 256),     # the reference's general case
    ([(1, 1023), (1, 879), (1, 778), (1, 1777)] * 100, 512),               # its flash-decoding case
    ([(5, 40), (1, 33)], 16),
]


@pytest.mark.parametrize("soft_cap", [None, 50])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("head_size", [128, 192, 256])
@pytest.mark.parametrize("heads", [(4, 4), (8, 2), (16, 2)])
@pytest.mark.parametrize("case", CASCADE_CASES, ids=["general", "decode400", "tiny"])
def test_cascade_attention(case, heads, head_size, dtype, soft_cap):
    from vllm_b200.attention.flash_backend import cascade_attention, flash_attn_varlen_func
    seq_lens, common_prefix_len = case
    H, Hkv = heads
    bs, nb = 16, 2048
    seed_all(0)
    query_lens = [x[0] for x in seq_lens]
    kv_lens = [x[1] for x in seq_lens]
    kc = torch.randn(nb, bs, Hkv, head_size).to(dtype)
    vc = torch.randn(nb, bs, Hkv, head_size).to(dtype)
    q = torch.randn(sum(query_lens), H, head_size).to(dtype)
    cu = torch.tensor([0] + query_lens, dtype=torch.int32).cumsum(0).to(torch.int32)
    maxb = (max(kv_lens) + bs - 1) // bs
    bt = torch.randint(0, nb, (len(seq_lens), maxb), dtype=torch.int32)
    ncb = common_prefix_len // bs
    bt[:, :ncb] = bt[0, :ncb]
    scale = float(head_size ** -0.5)
    cap = float(soft_cap or 0)
    qd, kd, vd, btd, cud = q.to(DEV), kc.to(DEV), vc.to(DEV), bt.to(DEV), cu.to(DEV)
    kvl = torch.tensor(kv_lens, dtype=torch.int32, device=DEV)
    ref_out = flash_attn_varlen_func(q=qd, k=kd, v=vd, cu_seqlens_q=cud, seqused_k=kvl, max_seqlen_q=max(query_lens),
                                     max_seqlen_k=max(kv_lens), softmax_scale=scale, causal=True, window_size=(-1, -1),
                                     block_table=btd, softcap=cap)
    out = torch.full_like(qd, float("nan"))
    cascade_attention(output=out, query=qd, key_cache=kd, value_cache=vd, cu_query_lens=cud,
                      max_query_len=max(query_lens),
                      cu_prefix_query_lens=torch.tensor([0, sum(query_lens)], dtype=torch.int32, device=DEV),
                      prefix_kv_lens=torch.tensor([common_prefix_len], dtype=torch.int32, device=DEV),
                      suffix_kv_lens=kvl - common_prefix_len, max_kv_len=max(kv_lens), softmax_scale=scale,
                      alibi_slopes=None, sliding_window=(-1, -1), logits_soft_cap=cap, block_table=btd,
                      common_prefix_len=common_prefix_len)
    torch.cuda.synchronize()
    torch.testing.assert_close(out.float(), ref_out.float(), atol=1e-2, rtol=1e-2)      # the reference test's check
    if len(seq_lens) <= 8:                                                              # and the oracle on the whole sequences
        want = oattn.flash_attn_varlen_paged(q, kc, vc, query_lens, kv_lens, bt, scale, soft_cap=soft_cap)
        torch.testing.assert_close(out.cpu().float(), want.float(), atol=1.5e-2, rtol=1e-2)


def test_flash_attn_varlen_softmax_lse_and_cascade_backend():
    """return_softmax_lse against the oracle's logits; B200FlashAttentionImpl.forward with use_cascade metadata
    (flash_attn.py:520-545) on an fp8 cache."""
    from oracle.fp8 import dequant_e4m3
    from vllm_b200.attention.flash_backend import (B200FlashAttentionImpl, B200FlashAttentionMetadata,
                                                   flash_attn_varlen_func)
    seed_all(9)
    H, Hkv, D, bs, nb, dtype = 8, 2, 128, 16, 256, torch.bfloat16
    seq_lens = [(33, 100), (1, 57), (70, 70)]
    q, kc, vc, kd, vd, bt, cu, query_lens, kv_lens = _varlen_inputs(seq_lens, H, Hkv, D, bs, dtype, "NHD", nb=nb)
    scale = float(D ** -0.5)
    kvl = torch.tensor(kv_lens, dtype=torch.int32, device=DEV)
    out, lse = flash_attn_varlen_func(q=q.to(DEV), k=kd, v=vd, cu_seqlens_q=cu.to(DEV), seqused_k=kvl,
                                      max_seqlen_q=max(query_lens), max_seqlen_k=max(kv_lens), softmax_scale=scale,
                                      causal=True, block_table=bt.to(DEV), return_softmax_lse=True)
    assert lse.shape == (H, sum(query_lens)) and lse.dtype == torch.float32
    want_lse = torch.empty(H, sum(query_lens))
    o = 0
    for i, (ql, kl) in enumerate(seq_lens):
        idx = bt[i, :(kl + bs - 1) // bs].long()
        k = torch.repeat_interleave(kc[idx].reshape(-1, Hkv, D)[:kl].float(), H // Hkv, dim=1)
        logits = torch.einsum("qhd,khd->hqk", q[o:o + ql].float() * scale, k)
        logits.masked_fill_(torch.triu(torch.ones(ql, kl), diagonal=kl - ql + 1).bool(), float("-inf"))
        want_lse[:, o:o + ql] = torch.logsumexp(logits, dim=-1)
        o += ql
    torch.testing.assert_close(lse.cpu(), want_lse, atol=2e-3, rtol=2e-3)

    # backend: two sequences sharing a 32-token prefix, fp8 cache, the cache already written (KV sharing flag)
    ksc, vsc = 0.05, 0.03
    kv8 = (torch.randn(2, nb, bs, Hkv, D) / 0.05).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
    lens = [(4, 90), (1, 45)]
    bt2 = torch.randint(0, nb, (2, 6), dtype=torch.int32)
    bt2[:, :2] = bt2[0, :2]
    T = sum(x[0] for x in lens)
    q2 = torch.randn(T, H, D).to(dtype)
    sl = torch.tensor([x[1] for x in lens], dtype=torch.int32, device=DEV)
    meta = B200FlashAttentionMetadata(
        num_actual_tokens=T, max_query_len=4, query_start_loc=torch.tensor([0, 4, 5], dtype=torch.int32, device=DEV),
        max_seq_len=90, seq_lens=sl, block_table=bt2.to(DEV), slot_mapping=torch.zeros(T, dtype=torch.int64, device=DEV),
        use_cascade=True, common_prefix_len=32, cu_prefix_query_lens=torch.tensor([0, T], dtype=torch.int32, device=DEV),
        prefix_kv_lens=torch.tensor([32], dtype=torch.int32, device=DEV), suffix_kv_lens=sl - 32)
    impl = B200FlashAttentionImpl(H, D, scale, Hkv, kv_cache_dtype="fp8", kv_sharing_target_layer_name="written already")

    class _Layer:
        _k_scale = torch.tensor([ksc], device=DEV)
        _v_scale = torch.tensor([vsc], device=DEV)
    out2 = torch.empty(T, H, D, dtype=dtype, device=DEV)
    impl.forward(_Layer(), q2.to(DEV), None, None, kv8.to(DEV), meta, output=out2)
    torch.cuda.synchronize()
    want2 = oattn.flash_attn_varlen_paged(q2, dequant_e4m3(kv8[0], ksc, dtype), dequant_e4m3(kv8[1], vsc, dtype),
                                          [x[0] for x in lens], [x[1] for x in lens], bt2, scale)
    torch.testing.assert_close(out2.cpu().float(), want2.float(), atol=2e-2, rtol=1e-2)


@pytest.mark.parametrize("soft_cap", [None, 30.0])
@pytest.mark.parametrize("heads", [(8, 2), (6, 6)])
@pytest.mark.parametrize("head_size", [64, 128])
def test_flash_attn_varlen_alibi(head_size, heads, soft_cap):
    """ALiBi on the variable-length kernel: against the oracle (flash-attn's documented bias) on a mixed batch, and, for
    one query token per sequence, against the decode kernel, whose ALiBi is pinned by the reference CPU kernel's golden
    (tests/golden/pa_mha_normal_alibi.npz)."""
    from vllm_b200 import _custom_ops as ops
    from vllm_b200.attention.flash_backend import flash_attn_varlen_func
    H, Hkv = heads
    seed_all(head_size + H)
    dtype, bs = torch.bfloat16, 16
    seq_lens = [(40, 300), (1, 77), (130, 130)]
    q, kc, vc, kd, vd, bt, cu, query_lens, kv_lens = _varlen_inputs(seq_lens, H, Hkv, head_size, bs, dtype, "NHD", nb=256)
    slopes = torch.rand(H) * 0.3
    scale = float(head_size ** -0.5)
    kvl = torch.tensor(kv_lens, dtype=torch.int32, device=DEV)
    out = flash_attn_varlen_func(q=q.to(DEV), k=kd, v=vd, cu_seqlens_q=cu.to(DEV), seqused_k=kvl,
                                 max_seqlen_q=max(query_lens), max_seqlen_k=max(kv_lens), softmax_scale=scale,
                                 causal=True, block_table=bt.to(DEV), alibi_slopes=slopes.to(DEV),
                                 softcap=soft_cap or 0.0)
    want = oattn.flash_attn_varlen_paged(q, kc, vc, query_lens, kv_lens, bt, scale, soft_cap=soft_cap, alibi_slopes=slopes)
    torch.testing.assert_close(out.cpu().float(), want.float(), atol=1.5e-2, rtol=1e-2)
    if soft_cap is None:
        S = 3
        qd = torch.randn(S, H, head_size).to(dtype).to(DEV)
        cu1 = torch.arange(S + 1, dtype=torch.int32, device=DEV)
        out_v = torch.empty_like(qd)
        ops.flash_attn_varlen_paged(out_v, qd, kd, vd, cu1, kvl, bt.to(DEV), scale, True, 1, alibi_slopes=slopes.to(DEV))
        out_d = torch.empty_like(qd)
        ops.paged_attention_flash(out_d, qd, kd, vd, scale, bt.to(DEV), kvl, max(kv_lens), slopes.to(DEV), "auto", None, None)
        torch.testing.assert_close(out_v.float(), out_d.float(), atol=1.5e-2, rtol=1e-2)


@pytest.mark.parametrize("kvd", ["auto", "fp8"])
@pytest.mark.parametrize("partition_size", [0, 128, -1])
@pytest.mark.parametrize("window", [1, 16, 100, 1000, 5000])
def test_paged_attention_flash_decode_with_sliding_window(window, partition_size, kvd):
    """Sliding window on the HBM-roofline decode kernel (vb200_paged_attention_flash_window): blocks left of the window
    are skipped, partitions counted from the window's first block; against the oracle (golden
    test_flash_attn.py:62-67) incl. forced split-K, windows shorter / longer than the sequences, ALiBi, fp8 cache."""
    from oracle.fp8 import dequant_e4m3
    from vllm_b200 import _custom_ops as ops
    seed_all(window + 7)
    H, Hkv, D, bs, nb, dtype = 8, 2, 128, 16, 700, torch.bfloat16
    kv_lens = [1, 15, 16, 17, 130, 1029, 2600]
    S = len(kv_lens)
    ksc, vsc = (0.05, 0.03) if kvd == "fp8" else (1.0, 1.0)
    if kvd == "fp8":
        k8 = (torch.randn(nb, bs, Hkv, D) / ksc).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
        v8 = (torch.randn(nb, bs, Hkv, D) / vsc).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
        kc, vc, kd, vd = dequant_e4m3(k8, ksc, dtype), dequant_e4m3(v8, vsc, dtype), k8.to(DEV), v8.to(DEV)
    else:
        kc, vc = torch.randn(nb, bs, Hkv, D).to(dtype), torch.randn(nb, bs, Hkv, D).to(dtype)
        kd, vd = kc.to(DEV), vc.to(DEV)
    maxb = (max(kv_lens) + bs - 1) // bs
    bt = torch.randint(0, nb, (S, maxb), dtype=torch.int32)
    q = torch.randn(S, H, D).to(dtype)
    slopes = torch.rand(H) * 0.05
    scale = float(D ** -0.5)
    out = torch.full_like(q, float("nan"), device=DEV)
    ops.paged_attention_flash(out, q.to(DEV), kd, vd, scale, bt.to(DEV), torch.tensor(kv_lens, dtype=torch.int32, device=DEV),
                              max(kv_lens), slopes.to(DEV), kvd, torch.tensor([ksc], device=DEV),
                              torch.tensor([vsc], device=DEV), partition_size=partition_size, window_left=window - 1)
    torch.cuda.synchronize()
    want = oattn.flash_attn_varlen_paged(q, kc, vc, [1] * S, kv_lens, bt, scale, sliding_window=window, alibi_slopes=slopes)
    torch.testing.assert_close(out.cpu().float(), want.float(), atol=2e-2 if kvd == "fp8" else 1.5e-2, rtol=1e-2)


@pytest.mark.parametrize("kvd", ["auto", "fp8"])
@pytest.mark.parametrize("window", [None, 200])
@pytest.mark.parametrize("soft_cap", [10.0, 50.0])
def test_paged_attention_flash_decode_softcap(soft_cap, window, kvd):
    """Logits soft-capping (with and without a window, ALiBi on top) on the decode kernel against the oracle
    (golden test_flash_attn.py:68-69), split-K forced."""
    from oracle.fp8 import dequant_e4m3
    from vllm_b200 import _custom_ops as ops
    seed_all(int(soft_cap) + (window or 0))
    H, Hkv, D, bs, nb, dtype = 8, 2, 128, 16, 400, torch.float16
    kv_lens = [1, 17, 333, 1500]
    S = len(kv_lens)
    ksc, vsc = (0.05, 0.03) if kvd == "fp8" else (1.0, 1.0)
    if kvd == "fp8":
        k8 = (torch.randn(nb, bs, Hkv, D) * 3 / ksc).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
        v8 = (torch.randn(nb, bs, Hkv, D) / vsc).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
        kc, vc, kd, vd = dequant_e4m3(k8, ksc, dtype), dequant_e4m3(v8, vsc, dtype), k8.to(DEV), v8.to(DEV)
    else:
        kc, vc = (torch.randn(nb, bs, Hkv, D) * 3).to(dtype), torch.randn(nb, bs, Hkv, D).to(dtype)   # logits past the cap
        kd, vd = kc.to(DEV), vc.to(DEV)
    bt = torch.randint(0, nb, (S, (max(kv_lens) + bs - 1) // bs), dtype=torch.int32)
    q = torch.randn(S, H, D).to(dtype)
    slopes = torch.rand(H) * 0.02
    scale = float(D ** -0.5)
    out = torch.full_like(q, float("nan"), device=DEV)
    ops.paged_attention_flash(out, q.to(DEV), kd, vd, scale, bt.to(DEV), torch.tensor(kv_lens, dtype=torch.int32, device=DEV),
                              max(kv_lens), slopes.to(DEV), kvd, torch.tensor([ksc], device=DEV),
                              torch.tensor([vsc], device=DEV), partition_size=256,
                              window_left=-1 if window is None else window - 1, softcap=soft_cap)
    torch.cuda.synchronize()
    want = oattn.flash_attn_varlen_paged(q, kc, vc, [1] * S, kv_lens, bt, scale, sliding_window=window, soft_cap=soft_cap,
                                         alibi_slopes=slopes)
    torch.testing.assert_close(out.cpu().float(), want.float(), atol=2e-2 if kvd == "fp8" else 1.5e-2, rtol=1e-2)
