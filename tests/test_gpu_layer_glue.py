"""GPU parity: RMSNorm family, rotary embedding and the fused silu_and_mul + per-token fp8 quant, through the
C ABI vs the oracle (oracle/layernorm.py, oracle/quant.py) and vs the reference's own CUDA kernels
(torch.ops._ref_C).  Modelled on /root/reference/tests/kernels/core/test_layernorm.py,
test_fused_quant_layernorm.py and test_pos_encoding.py (tolerances from there: 1e-2 / 1e-2 for the norms,
get_default_atol/rtol for rope); against the reference CUDA kernels the bf16 outputs differ only where rsqrt
or the reduction order moves a value across a rounding boundary (<= 1 ulp)."""
import pytest
import torch

from oracle import layernorm as oln
from oracle import quant as oq
from parity_utils import load_ref_cuda, seed_all

pytestmark = pytest.mark.gpu
DEV = "cuda"
FP8 = torch.float8_e4m3fn


def _ulp_close(a, b, dtype):
    """|a-b| <= 2 ulp of the 16-bit type (relative 2^-7 per ulp for bf16, 2^-10 for f16): the op rounds twice
    (T(x*inv) then T(n*w)), so a 1-ulp-of-fp32 difference in inv (rsqrtf vs torch.rsqrt, reduction order) can
    flip the first rounding and move the result by up to two steps; and almost every element is identical."""
    rel = 2.0 ** -7 if dtype == torch.bfloat16 else 2.0 ** -10
    torch.testing.assert_close(a.float(), b.float(), rtol=rel * 2.05, atol=1e-6)
    assert (a.float() != b.float()).float().mean().item() < 0.01


def _fp8_close(q, want, frac=0.002):
    """fp8 codes equal except where a <=1-ulp difference upstream crossed a rounding boundary."""
    a, b = q.view(torch.uint8), want.view(torch.uint8)
    neq = (a != b)
    assert neq.float().mean().item() <= frac, f"{neq.float().mean().item():.4f} of fp8 codes differ"
    torch.testing.assert_close(q.float(), want.float(), rtol=0.13, atol=2 ** -9)  # one e4m3 step


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("T,H", [(1, 4096), (7, 768), (256, 4096), (33, 8192), (5, 5120), (3, 16384)])
def test_rms_norm(dtype, T, H):
    from vllm_b200 import _custom_ops as ops
    seed_all(T + H)
    x = (torch.randn(T, H) * 2).to(dtype)
    w = (1 + 0.1 * torch.randn(H)).to(dtype)
    want = oln.rms_norm(x, w, 1e-6)
    out = torch.empty(T, H, dtype=dtype, device=DEV)
    ops.rms_norm(out, x.to(DEV), w.to(DEV), 1e-6)
    _ulp_close(out.cpu(), want, dtype)
    ref = load_ref_cuda()
    if ref is not None and hasattr(ref, "rms_norm"):
        r = torch.empty_like(out)
        ref.rms_norm(r, x.to(DEV), w.to(DEV), 1e-6)
        _ulp_close(out, r, dtype)
        assert (out != r).float().mean().item() < 0.01


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("T,H", [(1, 4096), (7, 768), (256, 4096), (33, 8192)])
@pytest.mark.parametrize("strided", [False, True])
def test_fused_add_rms_norm(dtype, T, H, strided):
    from vllm_b200 import _custom_ops as ops
    seed_all(T * 3 + H)
    full = (torch.randn(T, 2 * H if strided else H)).to(dtype)
    res = torch.randn(T, H).to(dtype)
    w = (1 + 0.1 * torch.randn(H)).to(dtype)
    x = full[:, :H]
    want, want_res = oln.fused_add_rms_norm(x, res, w, 1e-5)
    fd, rd = full.to(DEV), res.to(DEV)
    xd = fd[:, :H]
    ops.fused_add_rms_norm(xd, rd, w.to(DEV), 1e-5)
    assert torch.equal(rd.cpu().view(torch.int16), want_res.view(torch.int16)), "residual must be bit-exact"
    _ulp_close(xd.cpu(), want, dtype)
    if strided:  # the columns outside the view are untouched
        assert torch.equal(fd[:, H:].cpu().view(torch.int16), full[:, H:].view(torch.int16))
    ref = load_ref_cuda()
    if ref is not None and hasattr(ref, "fused_add_rms_norm") and not strided:
        f2, r2 = full.to(DEV), res.to(DEV)
        ref.fused_add_rms_norm(f2, r2, w.to(DEV), 1e-5)
        assert torch.equal(rd.view(torch.int16), r2.view(torch.int16))
        _ulp_close(xd, f2, dtype)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("T,H", [(1, 4096), (9, 1024), (256, 4096), (17, 8192)])
@pytest.mark.parametrize("has_res", [False, True])
@pytest.mark.parametrize("ub", [None, 3.0])
def test_rms_norm_dynamic_per_token_quant(dtype, T, H, has_res, ub):
    from vllm_b200 import _custom_ops as ops
    seed_all(T * 5 + H)
    x = (torch.randn(T, H) * 1.5).to(dtype)
    res = torch.randn(T, H).to(dtype) if has_res else None
    w = (1 + 0.1 * torch.randn(H)).to(dtype)
    ubt = None if ub is None else torch.tensor(ub)
    wq, ws, wres = oln.rms_norm_dynamic_per_token_quant(x, w, 1e-6, ub, res)
    rd = res.to(DEV) if has_res else None
    q, s = ops.rms_norm_dynamic_per_token_quant(x.to(DEV), w.to(DEV), 1e-6, FP8, None if ubt is None else ubt.to(DEV), rd)
    assert q.dtype == FP8 and s.shape == (T, 1)
    if has_res:
        assert torch.equal(rd.cpu().view(torch.int16), wres.view(torch.int16))
    torch.testing.assert_close(s.cpu(), ws, rtol=2.0 ** -7, atol=0)
    _fp8_close(q.cpu(), wq, frac=0.02)
    ref = load_ref_cuda()
    if ref is not None and hasattr(ref, "rms_norm_dynamic_per_token_quant"):
        q2 = torch.empty_like(q)
        s2 = torch.empty_like(s)
        r2 = res.to(DEV) if has_res else None
        ref.rms_norm_dynamic_per_token_quant(q2, x.to(DEV), w.to(DEV), s2, 1e-6, None if ubt is None else ubt.to(DEV), r2)
        torch.testing.assert_close(s, s2, rtol=2.0 ** -7, atol=0)
        _fp8_close(q, q2, frac=0.02)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("neox", [True, False])
@pytest.mark.parametrize("T,H,Hkv,D,rot", [(1, 32, 8, 128, 128), (256, 32, 8, 128, 128), (5, 12, 12, 64, 64),
                                           (9, 8, 2, 80, 32), (4, 8, 1, 256, 128)])
def test_rotary_embedding(dtype, neox, T, H, Hkv, D, rot):
    from vllm_b200 import _custom_ops as ops
    seed_all(T + H + D)
    cache = oln.make_cos_sin_cache(D, rot, 8192, 10000.0, dtype)
    pos = torch.randint(0, 8192, (T,), dtype=torch.int64)
    # q and k are slices of a fused qkv projection output (row stride != H*D)
    qkv = torch.randn(T, (H + 2 * Hkv) * D).to(dtype)
    q, k = qkv[:, :H * D], qkv[:, H * D:(H + Hkv) * D]
    wq, wk = oln.rotary_embedding(pos, q, k, D, cache, neox)
    qkv_d = qkv.to(DEV)
    qd, kd = qkv_d[:, :H * D], qkv_d[:, H * D:(H + Hkv) * D]
    ops.rotary_embedding(pos.to(DEV), qd, kd, D, cache.to(DEV), neox)
    assert torch.equal(qd.cpu().view(torch.int16), wq.contiguous().view(torch.int16)), "rope q must be bit-exact"
    assert torch.equal(kd.cpu().view(torch.int16), wk.contiguous().view(torch.int16)), "rope k must be bit-exact"
    assert torch.equal(qkv_d[:, (H + Hkv) * D:].cpu().view(torch.int16), qkv[:, (H + Hkv) * D:].contiguous().view(torch.int16))
    ref = load_ref_cuda()
    if ref is not None and hasattr(ref, "rotary_embedding"):
        r = qkv.to(DEV)
        ref.rotary_embedding(pos.to(DEV), r[:, :H * D], r[:, H * D:(H + Hkv) * D], D, cache.to(DEV), neox)
        assert torch.equal(r.view(torch.int16), qkv_d.view(torch.int16)), "rope differs from the reference CUDA kernel"
    # key=None
    q2 = qkv.to(DEV)[:, :H * D]
    ops.rotary_embedding(pos.to(DEV), q2, None, D, cache.to(DEV), neox)
    assert torch.equal(q2.cpu().view(torch.int16), wq.contiguous().view(torch.int16))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("T,d", [(1, 14336), (256, 14336), (7, 1792), (5, 128), (3, 16384), (64, 14336), (130, 7168),
                                 (2048, 2048), (33, 3584)])
def test_silu_and_mul_per_token_quant(dtype, T, d):
    """Bit-identical to silu_and_mul -> dynamic_per_token_scaled_fp8_quant run back to back (ours and the oracle)."""
    from vllm_b200 import _custom_ops as ops
    seed_all(T + d)
    x = (torch.randn(T, 2 * d) * 2).to(dtype)
    xd = x.to(DEV)
    q, s = ops.silu_and_mul_dynamic_per_token_quant(xd)
    act = torch.empty(T, d, dtype=dtype, device=DEV)
    ops.silu_and_mul(act, xd)
    q2, s2 = ops.scaled_fp8_quant(act, use_per_token_if_dynamic=True)
    assert torch.equal(s, s2)
    assert torch.equal(q.view(torch.uint8), q2.view(torch.uint8))
    wq, ws = oq.dynamic_per_token_scaled_fp8_quant(oq.silu_and_mul(x))
    torch.testing.assert_close(s.cpu(), ws, rtol=2.0 ** -7, atol=0)
    _fp8_close(q.cpu(), wq, frac=0.02)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("neox", [True, False])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("T,H,Hkv,D,rot,bs", [(1, 32, 8, 128, 128, 16), (256, 32, 8, 128, 128, 16), (37, 12, 12, 64, 64, 32),
                                              (9, 8, 2, 80, 32, 16), (4, 8, 1, 256, 128, 8), (5, 64, 8, 128, 128, 16),
                                              (3, 4, 4, 32, 32, 16), (7, 4, 1, 256, 256, 16), (6, 8, 2, 96, 96, 16)])
def test_rotary_embedding_reshape_and_cache(dtype, neox, kv_cache_dtype, T, H, Hkv, D, rot, bs):
    """The fused RoPE + KV-cache-write launch == rotary_embedding followed by reshape_and_cache, bit for bit, in q, k
    and both caches (ours back to back, and the oracle); padding tokens (slot -1) are rotated but not cached."""
    from oracle import cache as ocache
    from vllm_b200 import _custom_ops as ops
    fp8 = kv_cache_dtype == "fp8"
    seed_all(T + H + D + bs)
    cache = oln.make_cos_sin_cache(D, rot, 8192, 10000.0, dtype)
    pos = torch.randint(0, 8192, (T,), dtype=torch.int64)
    qkv = torch.randn(T, (H + 2 * Hkv) * D).to(dtype)
    x = 16 if fp8 else 8
    nb = (T + bs - 1) // bs + 3
    cdt = torch.uint8 if fp8 else dtype
    kc = torch.zeros(nb, Hkv, D // x, bs, x, dtype=cdt)
    vc = torch.zeros(nb, Hkv, D, bs, dtype=cdt)
    slots = torch.randperm(nb * bs)[:T].to(torch.int64)
    if T > 4:
        slots[2] = -1
    ks, vs = torch.tensor(0.037 if fp8 else 1.0), torch.tensor(0.051 if fp8 else 1.0)
    sl_q, sl_k, sl_v = slice(0, H * D), slice(H * D, (H + Hkv) * D), slice((H + Hkv) * D, (H + 2 * Hkv) * D)
    # two launches
    a = qkv.to(DEV)
    kc_a, vc_a = kc.to(DEV), vc.to(DEV)
    ops.rotary_embedding(pos.to(DEV), a[:, sl_q], a[:, sl_k], D, cache.to(DEV), neox)
    ops.reshape_and_cache(a[:, sl_k].view(T, Hkv, D), a[:, sl_v].view(T, Hkv, D), kc_a, vc_a, slots.to(DEV),
                          kv_cache_dtype, ks.to(DEV), vs.to(DEV))
    # one launch
    b = qkv.to(DEV)
    kc_b, vc_b = kc.to(DEV), vc.to(DEV)
    ops.rotary_embedding_reshape_and_cache(pos.to(DEV), b[:, sl_q], b[:, sl_k], b[:, sl_v], D, cache.to(DEV), neox,
                                           kc_b, vc_b, slots.to(DEV), kv_cache_dtype, ks.to(DEV), vs.to(DEV))
    torch.cuda.synchronize()
    bits = (lambda t: t.view(torch.uint8)) if fp8 else (lambda t: t.view(torch.int16))
    assert torch.equal(a.view(torch.int16), b.view(torch.int16)), "q / k / v after the fused launch differ"
    assert torch.equal(bits(kc_a), bits(kc_b)) and torch.equal(bits(vc_a), bits(vc_b)), "caches differ"
    # oracle: rope then scatter on the CPU
    wq, wk = oln.rotary_embedding(pos, qkv[:, sl_q], qkv[:, sl_k], D, cache, neox)
    kc_o, vc_o = kc.clone(), vc.clone()
    ocache.reshape_and_cache(wk.reshape(T, Hkv, D), qkv[:, sl_v].reshape(T, Hkv, D), kc_o, vc_o, slots, kv_cache_dtype, ks, vs)
    assert torch.equal(b[:, sl_q].cpu().view(torch.int16), wq.contiguous().view(torch.int16))
    assert torch.equal(bits(kc_b.cpu()), bits(kc_o)) and torch.equal(bits(vc_b.cpu()), bits(vc_o))
