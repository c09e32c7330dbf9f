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
    with pytest.raises(NotImplementedError):                                       # ALiBi outside the decode case
        flash_attn_varlen_func(q, kc, kc, 2, cu, 16, sl, block_table=bt, alibi_slopes=torch.ones(8, device=DEV))
    with pytest.raises(NotImplementedError):
        flash_attn_varlen_func(q, kc, kc, 1, cu, 16, sl, block_table=bt, softcap=30.0)
    with pytest.raises(NotImplementedError):
        flash_attn_varlen_func(q, kc, kc, 1, cu, 16, sl, block_table=bt, window_size=(128, 0))
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
