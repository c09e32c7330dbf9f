"""GPU parity: the fused-MoE grouped FP8 GEMM (invoke_fused_moe_kernel) and fused_experts / fused_moe through the
C ABI vs the oracle (oracle/moe.py, which follows fused_moe.py:455-468's operation order and torch_experts,
/root/reference/tests/kernels/utils.py:1060-1152).  Tolerance (north_star): 1e-2 rel / 2e-2 abs against the
reference's dequantised result.  Modelled on /root/reference/tests/kernels/moe/test_moe.py:60-160."""
import pytest
import torch

from oracle import moe as omoe
from oracle import quant as oq
from parity_utils import seed_all

pytestmark = pytest.mark.gpu
DEV = "cuda"
FP8 = torch.float8_e4m3fn
RTOL, ATOL = 1e-2, 2e-2


def _fp8_weights(E, N, K, per_channel):
    w = torch.randn(E, N, K) / 10
    if per_channel:
        s = w.abs().amax(dim=2).clamp(min=1e-6) / 448.0          # [E, N]
        q = (w / s[:, :, None]).clamp(-448, 448).to(FP8)
    else:
        s = w.abs().amax(dim=(1, 2)).clamp(min=1e-6) / 448.0      # [E]
        q = (w / s[:, None, None]).clamp(-448, 448).to(FP8)
    return q, s.float()


@pytest.mark.parametrize("T,k,E,N,K,block_m", [
    (1, 2, 8, 256, 128, 16), (33, 2, 8, 512, 256, 16), (128, 2, 8, 1024, 512, 64), (128, 2, 8, 384, 4096, 32),
    (64, 6, 64, 256, 512, 16), (300, 2, 4, 256, 256, 128), (700, 2, 2, 128, 128, 256), (17, 1, 3, 144, 272, 16)])
@pytest.mark.parametrize("per_channel", [False, True])
@pytest.mark.parametrize("mul_w", [False, True])
def test_grouped_gemm(T, k, E, N, K, block_m, per_channel, mul_w):
    from vllm_b200.fused_moe import invoke_fused_moe_kernel, moe_align_block_size
    seed_all(T + N)
    wq, ws = _fp8_weights(E, N, K, per_channel)
    a = torch.randn(T, K)
    if per_channel:
        aq, a_s = oq.dynamic_per_token_scaled_fp8_quant(a.bfloat16())
    else:
        aq, a_s = oq.dynamic_scaled_fp8_quant(a.bfloat16())
        a_s = a_s.reshape(1, 1)
    ids = torch.stack([torch.randperm(E)[:k] for _ in range(T)]).to(torch.int32)
    tw = torch.rand(T, k) + 0.1
    want = omoe.grouped_gemm_fp8(aq, a_s, wq, ws, ids, k, tw if mul_w else None, torch.bfloat16)
    sorted_ids, expert_ids, npost = moe_align_block_size(ids.to(DEV), block_m, E)
    c = torch.full((T, k, N), float("nan"), dtype=torch.bfloat16, device=DEV)
    invoke_fused_moe_kernel(aq.to(DEV), wq.to(DEV), c, a_s.to(DEV).reshape(-1), ws.to(DEV), tw.to(DEV) if mul_w else None,
                            sorted_ids, expert_ids, npost, mul_w, k, block_m, True, per_channel)
    torch.cuda.synchronize()
    got = c.view(T * k, N).cpu().float()
    assert torch.isfinite(got).all(), "every (token, expert) row must be written"
    torch.testing.assert_close(got, want.float(), rtol=RTOL, atol=ATOL)
    # the fp32 accumulation is exact for fp8 x fp8 products of this size: only the final bf16 rounding may differ
    assert (got - want.float()).abs().max() <= 2.0 ** -7 * want.float().abs().max() + 1e-6


def test_grouped_gemm_expert_map_zeros():
    """Blocks whose expert is mapped to -1 (expert parallel) are written as zeros (fused_moe.py:370-377)."""
    from vllm_b200.fused_moe import invoke_fused_moe_kernel, moe_align_block_size
    seed_all(3)
    T, k, E, N, K, bm = 40, 2, 4, 256, 256, 16
    wq, ws = _fp8_weights(E, N, K, False)
    aq, a_s = oq.dynamic_scaled_fp8_quant(torch.randn(T, K).bfloat16())
    ids = torch.stack([torch.randperm(E)[:k] for _ in range(T)]).to(torch.int32)
    emap = torch.tensor([0, -1, 1, -1], dtype=torch.int32)          # experts 0 and 2 live here as local 0, 1
    sorted_ids, expert_ids, npost = moe_align_block_size(ids.to(DEV), bm, E, emap.to(DEV))
    local_w, local_s = wq[[0, 2]], ws[[0, 2]]
    c = torch.full((T, k, N), float("nan"), dtype=torch.bfloat16, device=DEV)
    invoke_fused_moe_kernel(aq.to(DEV), local_w.to(DEV), c, a_s.to(DEV), local_s.to(DEV), None, sorted_ids, expert_ids,
                            npost, False, k, bm)
    got = c.view(T * k, N).cpu().float()
    want = omoe.grouped_gemm_fp8(aq, a_s.reshape(1, 1), wq, ws, ids, k, None, torch.bfloat16).float()
    dead = ((ids.flatten() == 1) | (ids.flatten() == 3))
    assert (got[dead] == 0).all()
    torch.testing.assert_close(got[~dead], want[~dead], rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("T,k,E,H,I", [(1, 2, 8, 256, 512), (128, 2, 8, 512, 1024), (33, 6, 64, 256, 256),
                                        (222, 2, 8, 1024, 896)])
@pytest.mark.parametrize("per_channel", [False, True])
def test_fused_experts_fp8(T, k, E, H, I, per_channel):
    from vllm_b200.fused_moe import fused_experts, fused_topk
    seed_all(T + H)
    w1q, w1s = _fp8_weights(E, 2 * I, H, per_channel)
    w2q, w2s = _fp8_weights(E, H, I, per_channel)
    x = (torch.randn(T, H) / 2).bfloat16()
    gate = torch.randn(T, E)
    tw_o, ids_o, _ = omoe.topk_softmax(gate, k)
    tw_o = tw_o / tw_o.sum(-1, keepdim=True)
    want = omoe.fused_experts_fp8(x, w1q, w2q, w1s, w2s, tw_o, ids_o, per_channel_quant=per_channel)
    tw, ids, _ = fused_topk(x.to(DEV), gate.to(DEV), k, True)
    assert torch.equal(ids.cpu(), ids_o), "top-k ids must be bit-exact"
    w1s_d = w1s.to(DEV).reshape(E, -1, 1) if per_channel else w1s.to(DEV)
    w2s_d = w2s.to(DEV).reshape(E, -1, 1) if per_channel else w2s.to(DEV)
    out = fused_experts(x.to(DEV), w1q.to(DEV), w2q.to(DEV), tw, ids, use_fp8_w8a8=True, per_channel_quant=per_channel,
                        w1_scale=w1s_d, w2_scale=w2s_d)
    torch.cuda.synchronize()
    # intermediate fp8 re-quantisation makes single elements flip by one fp8 step: compare with the reference
    # tests' own bound for fp8 MoE (test_moe.py: atol 2e-2 / rtol 1e-2 scaled by the output magnitude)
    got, w = out.cpu().float(), want.float()
    scale = w.abs().max().item()
    torch.testing.assert_close(got, w, rtol=2e-2, atol=2e-2 * max(scale, 1.0))
    assert (got - w).abs().mean().item() <= 3e-3 * max(scale, 1.0)


def test_fused_experts_cuda_graph():
    """Capturable in a CUDA graph (tests/kernels/moe/test_moe.py:93-107 does the same)."""
    from vllm_b200.fused_moe import fused_moe
    seed_all(1)
    T, k, E, H, I = 64, 2, 8, 256, 512
    w1q, w1s = _fp8_weights(E, 2 * I, H, False)
    w2q, w2s = _fp8_weights(E, H, I, False)
    x = (torch.randn(T, H) / 2).bfloat16().to(DEV)
    gate = torch.randn(T, E).to(DEV)
    args = (x, w1q.to(DEV), w2q.to(DEV), gate, k, True)
    kw = dict(use_fp8_w8a8=True, w1_scale=w1s.to(DEV), w2_scale=w2s.to(DEV))
    eager = fused_moe(*args, **kw)
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fused_moe(*args, **kw)
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        out = fused_moe(*args, **kw)
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, eager)


@pytest.mark.parametrize("T,k,E,H,I", [(1, 2, 8, 256, 512), (33, 2, 8, 128, 1024), (128, 2, 8, 512, 1024),
                                        (33, 6, 64, 256, 256), (222, 2, 8, 1024, 896), (64, 2, 8, 2048, 200)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_fused_experts_unquantised(T, k, E, H, I, dtype):
    """The bf16 / f16-weights variant of fused_experts (tests/kernels/moe/test_moe.py:117-233 compares
    fused_moe with torch_moe on randn / 10 weights at atol 2e-2): same tcgen05 grouped GEMM with kind::f16 MMAs."""
    from vllm_b200.fused_moe import fused_experts, fused_topk
    seed_all(T + H + I)
    w1 = (torch.randn(E, 2 * I, H) / 10).to(dtype)
    w2 = (torch.randn(E, H, I) / 10).to(dtype)
    x = (torch.randn(T, H) / 10).to(dtype)
    gate = torch.randn(T, E)
    tw_o, ids_o, _ = omoe.topk_softmax(gate, k)
    tw_o = tw_o / tw_o.sum(-1, keepdim=True)
    want = omoe.fused_experts_bf16(x, w1, w2, tw_o, ids_o)
    tw, ids, _ = fused_topk(x.to(DEV), gate.to(DEV), k, True)
    assert torch.equal(ids.cpu(), ids_o), "top-k ids must be bit-exact"
    out = fused_experts(x.to(DEV), w1.to(DEV), w2.to(DEV), tw, ids)
    torch.cuda.synchronize()
    assert out.dtype == dtype and out.shape == (T, H)
    torch.testing.assert_close(out.cpu().float(), want.float(), rtol=1e-2, atol=2e-2)


def test_grouped_gemm_16_rejects_mixed_dtypes():
    from vllm_b200.fused_moe.fused_moe import invoke_fused_moe_kernel
    from vllm_b200.fused_moe import moe_align_block_size
    ids = torch.zeros(4, 2, dtype=torch.int32, device=DEV)
    s, e, n = moe_align_block_size(ids, 16, 2)
    a = torch.zeros(4, 64, dtype=torch.bfloat16, device=DEV)
    b = torch.zeros(2, 128, 64, dtype=torch.float16, device=DEV)
    c = torch.zeros(4, 2, 128, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(RuntimeError):
        invoke_fused_moe_kernel(a, b, c, None, None, None, s, e, n, False, 2, 16, False)


@pytest.mark.parametrize("quant", ["fp8", "bf16"])
@pytest.mark.parametrize("mode", ["single", "tp2", "ep2"])
def test_fused_moe_layer(quant, mode):
    """FusedMoE.forward_impl + Fp8MoEMethod.apply / UnquantizedFusedMoEMethod (layer.py:611-1533, fp8.py:918-1010):
    the layer equals fused_experts on its weights; two tensor-parallel (intermediate split) or expert-parallel
    (expert_map) halves, built as the ranks of a TP-2 group would build them, sum to the unsharded layer."""
    from vllm_b200.distributed import communication_op as co
    from vllm_b200.fused_moe import fused_experts, fused_topk
    from vllm_b200.fused_moe.layer import Fp8MoEMethod, FusedMoE
    seed_all(17)
    T, E, k, H, I = 48, 8, 2, 256, 512
    dt = torch.bfloat16
    x = (torch.randn(T, H) / 4).to(dt).to(DEV)
    logits = torch.randn(T, E).to(dt).to(DEV)
    if quant == "fp8":
        w13, s13 = _fp8_weights(E, 2 * I, H, False)
        w2, s2 = _fp8_weights(E, H, I, False)
    else:
        w13, s13 = (torch.randn(E, 2 * I, H) / 10).to(dt), None
        w2, s2 = (torch.randn(E, H, I) / 10).to(dt), None
    tw, ids, _ = fused_topk(x, logits, k, True)
    kw = dict(use_fp8_w8a8=True, w1_scale=s13.to(DEV), w2_scale=s2.to(DEV)) if quant == "fp8" else {}
    want = fused_experts(x, w13.to(DEV), w2.to(DEV), tw, ids, **kw)

    def build(world, rank, ep):
        saved = (co._TP_WORLD, co._TP_RANK)
        co._TP_WORLD, co._TP_RANK = world, rank
        try:
            layer = FusedMoE(E, k, H, I, params_dtype=dt, reduce_results=False, renormalize=True,
                             quant_method=Fp8MoEMethod() if quant == "fp8" else None, ep_size=world if ep else None)
        finally:
            co._TP_WORLD, co._TP_RANK = saved
        layer = layer.to(DEV)
        if ep:
            e0 = rank * (E // world)
            sl_e, i0, i1 = slice(e0, e0 + E // world), 0, I
        else:
            per = I // world
            sl_e, i0, i1 = slice(0, E), rank * per, (rank + 1) * per
        assert layer.w13_weight.shape == (sl_e.stop - sl_e.start, 2 * (i1 - i0), H)
        layer.w13_weight.data.copy_(torch.cat([w13[sl_e, i0:i1], w13[sl_e, I + i0:I + i1]], dim=1).to(DEV))
        layer.w2_weight.data.copy_(w2[sl_e, :, i0:i1].to(DEV))
        if quant == "fp8":
            layer.w13_weight_scale.data.copy_(s13[sl_e].to(DEV))
            layer.w2_weight_scale.data.copy_(s2[sl_e].to(DEV))
        layer.quant_method.process_weights_after_loading(layer)
        return layer

    if mode == "single":
        got = build(1, 0, False)(x, logits)
        assert torch.equal(got, want)
        return
    parts = [build(2, r, mode == "ep2")(x, logits).float() for r in range(2)]
    got = parts[0] + parts[1]
    scale = max(want.float().abs().max().item(), 1.0)
    # ep2: each (token, expert) pair is computed by exactly one rank with the same arithmetic -> only the final bf16
    # sums differ; tp2 (fp8): the second quantisation sees half of the intermediate row, so its per-tensor scale differs
    torch.testing.assert_close(got, want.float(), rtol=2e-2, atol=2e-2 * scale)
    assert (got - want.float()).abs().mean().item() <= 4e-3 * scale


@pytest.mark.parametrize("T,k,E,H,I", [(1, 2, 8, 256, 512), (128, 2, 8, 512, 1024), (33, 6, 64, 256, 256),
                                       (300, 2, 4, 256, 128), (128, 2, 8, 4096, 1792)])
@pytest.mark.parametrize("static_scales", [False, True])
def test_fused_experts_per_tensor_pipeline_is_bit_identical_to_the_unfused_ops(T, k, E, H, I, static_scales):
    """fused_experts with per-tensor activation scales computes only the scales up front (abs-max, fused with
    silu_and_mul for the second GEMM) and quantises while it gathers (vb200_moe_grouped_gemm_fp8_q16): 8 launches
    instead of 12.  Same bytes into the GEMMs: the output must equal, bit for bit, the reference's sequence
    scaled_fp8_quant -> GEMM -> silu_and_mul -> scaled_fp8_quant -> GEMM -> moe_sum (fused_moe.py:1470-1560) run op by op."""
    from vllm_b200 import _custom_ops as ops
    from vllm_b200.fused_moe import fused_experts, invoke_fused_moe_kernel, moe_align_block_size
    from vllm_b200.fused_moe.fused_moe import get_block_m
    seed_all(T + H + I)
    w1, s1 = _fp8_weights(E, 2 * I, H, False)
    w2, s2 = _fp8_weights(E, H, I, False)
    x = (torch.randn(T, H) * 0.7).bfloat16().to(DEV)
    ids = torch.stack([torch.randperm(E)[:k] for _ in range(T)]).to(torch.int32).to(DEV)
    tw = (torch.rand(T, k) + 0.1).to(DEV)
    w1d, s1d, w2d, s2d = w1.to(DEV), s1.to(DEV), w2.to(DEV), s2.to(DEV)
    a1 = torch.tensor([0.011], device=DEV) if static_scales else None
    a2 = torch.tensor([0.023], device=DEV) if static_scales else None
    got = fused_experts(x, w1d, w2d, tw, ids, use_fp8_w8a8=True, w1_scale=s1d, w2_scale=s2d, a1_scale=a1, a2_scale=a2)
    # the same pipeline op by op
    bm = get_block_m(T * k, E)
    sorted_ids, expert_ids, npost = moe_align_block_size(ids, bm, E)
    a1q, a1s = ops.scaled_fp8_quant(x, a1)
    c1 = torch.empty(T, k, 2 * I, dtype=torch.bfloat16, device=DEV)
    invoke_fused_moe_kernel(a1q, w1d, c1, a1s, s1d, None, sorted_ids, expert_ids, npost, False, k, bm, True, False)
    c2 = torch.empty(T * k, I, dtype=torch.bfloat16, device=DEV)
    ops.silu_and_mul(c2, c1.view(T * k, 2 * I))
    a2q, a2s = ops.scaled_fp8_quant(c2, a2)
    c3 = torch.empty(T, k, H, dtype=torch.bfloat16, device=DEV)
    invoke_fused_moe_kernel(a2q, w2d, c3, a2s, s2d, tw, sorted_ids, expert_ids, npost, True, 1, bm, True, False)
    want = torch.empty_like(x)
    ops.moe_sum(c3, want)
    torch.cuda.synchronize()
    assert torch.equal(got.view(torch.int16), want.view(torch.int16))


def test_silu_and_mul_absmax_matches_the_two_ops():
    from vllm_b200 import _custom_ops as ops
    seed_all(12)
    x = (torch.randn(77, 2 * 1792) * 2).bfloat16().to(DEV)
    want = torch.empty(77, 1792, dtype=torch.bfloat16, device=DEV)
    ops.silu_and_mul(want, x)
    _, want_scale = ops.scaled_fp8_quant(want, None)
    out = torch.empty_like(want)
    scale = torch.zeros(1, dtype=torch.float32, device=DEV)
    ops.silu_and_mul_absmax(out, scale, x)
    assert torch.equal(out.view(torch.int16), want.view(torch.int16))
    assert torch.equal(scale.view(-1), want_scale.view(-1))
    s2 = torch.zeros(1, dtype=torch.float32, device=DEV)
    ops.absmax_scale(s2, want)
    assert torch.equal(s2.view(-1), want_scale.view(-1))
