"""Mirror of the decode-relevant part of /root/reference/vllm/model_executor/layers/fused_moe/fused_moe.py:
fused_topk (:882-915), invoke_fused_moe_kernel (:478-647, the Triton kernel :267-475 replaced by the tcgen05
grouped GEMM), fused_experts / fused_experts_impl (:1253-1346, :1349-1563), fused_moe (:1566-1670).
FP8 w8a8 experts (per-tensor or per-channel weight scales; static / dynamic per-tensor / per-token activation scales:
what BASELINE config 4, Mixtral-8x7B FP8, runs) and unquantised f16 / bf16 experts (SURVEY 8d's bf16-weights variant,
tests/kernels/moe/test_moe.py:117-233) on the same tcgen05 grouped GEMM."""
import os
from typing import Optional

import torch

from .. import _custom_ops as ops
from .._lib import check, dtype_enum, load, ptr, stream
from .moe_align_block_size import moe_align_block_size

FP8 = torch.float8_e4m3fn


def fused_topk(hidden_states: torch.Tensor, gating_output: torch.Tensor, topk: int, renormalize: bool,
               indices_type: Optional[torch.dtype] = None) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """fused_moe.py:882-915 (+ vllm_topk_softmax :866-879): fp32 softmax, top-k with lowest-index ties,
    optional renormalisation w /= w.sum(-1)."""
    assert hidden_states.size(0) == gating_output.size(0), "Number of tokens mismatch"
    M = hidden_states.size(0)
    dev = hidden_states.device
    topk_weights = torch.empty(M, topk, dtype=torch.float32, device=dev)
    topk_ids = torch.empty(M, topk, dtype=torch.int32 if indices_type is None else indices_type, device=dev)
    token_expert_indices = torch.empty(M, topk, dtype=torch.int32, device=dev)
    if topk_ids.dtype == torch.int32 and gating_output.is_cuda and gating_output.is_contiguous() and \
            gating_output.dtype in (torch.float32, torch.float16, torch.bfloat16):
        # one launch: logits read in their own dtype (the host-side .float() of fused_moe.py:908 is exact), softmax,
        # top-k and the renormalisation of :869-870
        check(load().vb200_fused_topk(ptr(topk_weights), ptr(topk_ids), ptr(token_expert_indices), ptr(gating_output),
                                      M, gating_output.size(-1), topk, int(bool(renormalize)),
                                      dtype_enum(gating_output.dtype), stream()))
        return topk_weights, topk_ids, token_expert_indices
    ops.topk_softmax(topk_weights, topk_ids, token_expert_indices, gating_output.float().contiguous())
    if renormalize:
        topk_weights = topk_weights / topk_weights.sum(dim=-1, keepdim=True)
    return topk_weights, topk_ids, token_expert_indices


def router_topk(hidden_states: torch.Tensor, w_gate: torch.Tensor, topk: int, renormalize: bool,
                absmax_scale: Optional[torch.Tensor] = None, return_logits: bool = False):
    """The gate Linear + fused_topk in one launch (include/vllm_b200.h vb200_moe_router_topk): router logits =
    hidden @ w_gate^T rounded to the activation dtype, then exactly fused_topk on them.  absmax_scale: optional pre-zeroed
    fp32 [1] that receives max|hidden| / 448 (the experts' dynamic per-tensor a1 scale).  Returns (topk_weights fp32
    [T, k], topk_ids int32 [T, k]) and, with return_logits, the logits [T, E]."""
    T, H = hidden_states.shape
    E = w_gate.shape[0]
    dev = hidden_states.device
    if not (hidden_states.is_cuda and hidden_states.is_contiguous() and w_gate.is_contiguous()) or \
            w_gate.dtype != hidden_states.dtype or w_gate.shape[1] != H:
        raise RuntimeError("router_topk: contiguous CUDA hidden [T, H] and w_gate [E, H] of one dtype")
    topk_weights = torch.empty(T, topk, dtype=torch.float32, device=dev)
    topk_ids = torch.empty(T, topk, dtype=torch.int32, device=dev)
    logits = torch.empty(T, E, dtype=hidden_states.dtype, device=dev) if return_logits else None
    check(load().vb200_moe_router_topk(ptr(topk_weights), ptr(topk_ids), None, ptr(logits), ptr(absmax_scale),
                                       ptr(hidden_states), ptr(w_gate), T, H, E, topk, int(bool(renormalize)),
                                       dtype_enum(hidden_states.dtype), stream()))
    return (topk_weights, topk_ids, logits) if return_logits else (topk_weights, topk_ids)


def get_block_m(num_pairs: int, num_experts: int) -> int:
    """M-block (tokens per grouped-GEMM tile) for the tcgen05 kernel: tokens sit on the UMMA N side, whose size
    the kernel trims to the block's valid rows, so padding costs no math -- pick the block so that an average
    expert fits in one (stands where get_default_config's BLOCK_SIZE_M stands, fused_moe.py:766-824)."""
    avg = max(1, num_pairs // max(num_experts, 1))
    for bm in (16, 32, 64, 128):
        if avg * 3 <= bm * 2:
            return bm
    # above 256 rows per expert every expert has several M-blocks whatever the block; with the M-blocks of a weight tile
    # launched back to back the 128-row block (two CTAs per SM) is the faster one (csrc/moe_gemm.cu, ragged rule)
    return 256 if 192 < avg <= 256 else 128


def _moe_workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    """Scratch for the sorted-order activation gather: taken per call from torch's caching allocator (as
    paged_attention_v1 does for its partials), so it is stream-ordered, never shared between streams, and inside
    a CUDA-graph capture it comes from -- and lives as long as -- the graph's private pool."""
    return torch.empty(max(nbytes, 16), dtype=torch.uint8, device=device)


def invoke_fused_moe_kernel(A: torch.Tensor, B: torch.Tensor, C: torch.Tensor, A_scale: Optional[torch.Tensor],
                            B_scale: Optional[torch.Tensor], topk_weights: Optional[torch.Tensor],
                            sorted_token_ids: torch.Tensor, expert_ids: torch.Tensor,
                            num_tokens_post_padded: torch.Tensor, mul_routed_weight: bool, top_k: int,
                            block_m: int, use_fp8_w8a8: bool = True, per_channel_quant: bool = False) -> None:
    """fused_moe.py:478-647.  A fp8 [rows, K] (row = token_id // top_k); B fp8 [E, N, K]; C [num_pairs, N]
    (any leading shape, contiguous); A_scale [1] or [rows, 1]; B_scale [E] or [E, N(,1)]."""
    assert sorted_token_ids.stride(0) == 1 and B.stride(-1) == 1 and A.stride(-1) == 1
    assert A.is_contiguous() and B.is_contiguous() and C.is_contiguous()
    E, N, K = B.shape
    num_pairs = C.numel() // N
    lib = load()
    if not use_fp8_w8a8:
        # unquantised experts: f16 / bf16 operands, no scales (fused_moe.py:1349-1563 with every use_* flag off)
        if A.dtype not in (torch.float16, torch.bfloat16) or B.dtype != A.dtype or C.dtype != A.dtype:
            raise RuntimeError("invoke_fused_moe_kernel: unquantised experts need f16 / bf16 A, B and C of one dtype")
        ws_bytes = lib.vb200_moe_grouped_gemm_workspace_size(sorted_token_ids.numel(), 2 * K)
        ws = _moe_workspace(A.device, ws_bytes)
        check(lib.vb200_moe_grouped_gemm_16(
            ptr(C), ptr(A), ptr(B), ptr(topk_weights) if mul_routed_weight else None, ptr(sorted_token_ids),
            ptr(expert_ids), ptr(num_tokens_post_padded), num_pairs, A.shape[0], top_k, E, N, K, block_m,
            sorted_token_ids.numel(), int(bool(mul_routed_weight)), dtype_enum(A.dtype), ptr(ws), ws.numel(), stream()))
        return
    if B.dtype != FP8:
        raise RuntimeError("invoke_fused_moe_kernel: use_fp8_w8a8 needs float8_e4m3fn B")
    ws_bytes = lib.vb200_moe_grouped_gemm_workspace_size(sorted_token_ids.numel(), K)
    ws = _moe_workspace(A.device, ws_bytes)
    if A.dtype in (torch.float16, torch.bfloat16):
        # unquantised activations + their per-tensor scale: quantised while they are gathered into expert order
        # (bit-identical to static_scaled_fp8_quant followed by the fp8 call below)
        if A_scale is None or A_scale.numel() != 1:
            raise RuntimeError("invoke_fused_moe_kernel: 16-bit activations need a per-tensor A_scale")
        check(lib.vb200_moe_grouped_gemm_fp8_q16(
            ptr(C), ptr(A), dtype_enum(A.dtype), ptr(B), ptr(A_scale), ptr(B_scale),
            ptr(topk_weights) if mul_routed_weight else None, ptr(sorted_token_ids), ptr(expert_ids),
            ptr(num_tokens_post_padded), num_pairs, A.shape[0], top_k, E, N, K, block_m, sorted_token_ids.numel(),
            1 if B_scale.numel() > E else 0, int(bool(mul_routed_weight)), dtype_enum(C.dtype), ptr(ws), ws.numel(),
            stream()))
        return
    if A.dtype != FP8:
        raise RuntimeError("invoke_fused_moe_kernel: use_fp8_w8a8 needs float8_e4m3fn (or, with a per-tensor scale, "
                           "f16 / bf16) A")
    a_per_token = 1 if A_scale.numel() > 1 else 0
    b_per_channel = 1 if B_scale.numel() > E else 0
    if a_per_token and A_scale.numel() != A.shape[0]:
        raise RuntimeError("invoke_fused_moe_kernel: per-token A_scale needs one entry per row of A")
    check(lib.vb200_moe_grouped_gemm_fp8(
        ptr(C), ptr(A), ptr(B), ptr(A_scale), ptr(B_scale), ptr(topk_weights) if mul_routed_weight else None,
        ptr(sorted_token_ids), ptr(expert_ids), ptr(num_tokens_post_padded), num_pairs, A.shape[0], top_k, E, N, K,
        block_m, sorted_token_ids.numel(), a_per_token, b_per_channel, int(bool(mul_routed_weight)),
        dtype_enum(C.dtype), ptr(ws), ws.numel(), stream()))


def fused_experts(hidden_states: torch.Tensor, w1: torch.Tensor, w2: torch.Tensor, topk_weights: torch.Tensor,
                  topk_ids: torch.Tensor, inplace: bool = False, activation: str = "silu",
                  is_act_and_mul: bool = True, apply_router_weight_on_input: bool = False,
                  use_fp8_w8a8: bool = False, use_int8_w8a8: bool = False, use_int8_w8a16: bool = False,
                  use_int4_w4a16: bool = False, use_mxfp4_w4a4: bool = False, per_channel_quant: bool = False,
                  global_num_experts: int = -1, expert_map: Optional[torch.Tensor] = None,
                  w1_scale: Optional[torch.Tensor] = None, w2_scale: Optional[torch.Tensor] = None,
                  w1_zp=None, w2_zp=None, a1_scale: Optional[torch.Tensor] = None,
                  a2_scale: Optional[torch.Tensor] = None, block_shape=None, allow_deep_gemm: bool = False,
                  allow_cutlass_block_scaled_grouped_gemm: bool = False,
                  dynamic_scales: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fused_moe.py:1253-1346 / fused_experts_impl :1349-1563 for use_fp8_w8a8 (Mixtral FP8):
      quant(hidden) -> align -> GEMM1 [T, k, 2I] -> silu_and_mul -> quant -> GEMM2 (x routed weight) [T, k, H]
      -> moe_sum.  w1 [E, 2I, H], w2 [E, H, I] fp8 with K contiguous.
    dynamic_scales (B200 extension, per-tensor dynamic quantisation only): an fp32 [2] tensor whose [0] ALREADY holds
    the abs-max scale of hidden_states (router_topk computed it) and whose [1] is zero."""
    if use_int8_w8a8 or use_int8_w8a16 or use_int4_w4a16 or use_mxfp4_w4a4:
        raise RuntimeError("fused_experts: fp8 w8a8 and unquantised f16 / bf16 experts are built in vllm_b200")
    if activation != "silu" or not is_act_and_mul or apply_router_weight_on_input or block_shape is not None:
        raise RuntimeError("fused_experts: only silu-and-mul experts with per-tensor / per-channel scales are built")
    assert hidden_states.size(1) == w1.size(2), f"Hidden size mismatch {hidden_states.size(1)} != {w1.size(2)}"
    assert topk_weights.size() == topk_ids.size(), "topk shape mismatch"
    assert hidden_states.is_contiguous(), "Hidden_states must be contiguous"
    assert w1.stride(-1) == 1 and w2.stride(-1) == 1, "Stride of last dimension must be 1"
    assert hidden_states.dtype in (torch.float16, torch.bfloat16)
    assert not use_fp8_w8a8 or (w1_scale is not None and w2_scale is not None)
    T, H = hidden_states.shape
    E, N, _ = w1.shape
    if global_num_experts == -1:
        global_num_experts = E
    k = topk_ids.size(1)
    dev, dt = hidden_states.device, hidden_states.dtype
    if topk_ids.dtype != torch.int32:
        topk_ids = topk_ids.to(torch.int32)
    block_m = get_block_m(T * k, global_num_experts)
    sorted_ids, expert_ids, num_post = moe_align_block_size(topk_ids, block_m, global_num_experts, expert_map)
    cache1 = torch.empty((T, k, N), device=dev, dtype=dt)
    if not use_fp8_w8a8:
        # unquantised experts: GEMM 1 -> silu_and_mul -> GEMM 2 (x routed weight) -> moe_sum, all in the activation dtype
        assert w1.dtype == dt and w2.dtype == dt, "unquantised experts must have the activation dtype"
        invoke_fused_moe_kernel(hidden_states, w1, cache1, None, None, None, sorted_ids, expert_ids, num_post, False, k,
                                block_m, False)
        cache2 = torch.empty((T * k, N // 2), device=dev, dtype=dt)
        ops.silu_and_mul(cache2, cache1.view(T * k, N))
        cache3 = torch.empty((T, k, w2.shape[1]), device=dev, dtype=dt)
        invoke_fused_moe_kernel(cache2, w2, cache3, None, None, topk_weights.contiguous(), sorted_ids, expert_ids,
                                num_post, True, 1, block_m, False)
        out = hidden_states if inplace else torch.empty_like(hidden_states)
        ops.moe_sum(cache3, out)
        return out
    # activation quant: per-token when per_channel_quant, static when a scale is given, else dynamic per-tensor
    # (moe_kernel_quantize_input, fused_moe/utils.py)
    if not per_channel_quant and (N // 2) % 16 == 0 and H % 16 == 0:
        # Per-tensor activation scales (Mixtral-style fp8 checkpoints).  The reference quantises the whole tensor and the
        # GEMM gathers fp8 rows; here only the SCALE is computed up front (dynamic: one abs-max pass, fused with the
        # activation for the second GEMM) and the rows are quantised while they are gathered into expert order: 8
        # launches instead of 12 per layer, same bytes into the GEMMs, bit-identical output.
        scales = dynamic_scales
        if scales is None and (a1_scale is None or a2_scale is None):
            scales = torch.zeros(2, dtype=torch.float32, device=dev)   # one fill for both dynamic scales
        if a1_scale is None:
            a1s = scales[0:1]
            if dynamic_scales is None:
                ops.absmax_scale(a1s, hidden_states)
        else:
            a1s = a1_scale.reshape(-1)[:1].float()
        invoke_fused_moe_kernel(hidden_states, w1, cache1, a1s, w1_scale, None, sorted_ids, expert_ids, num_post, False,
                                k, block_m, True, False)
        cache2 = torch.empty((T * k, N // 2), device=dev, dtype=dt)
        if a2_scale is None:
            a2s = scales[1:2]
            ops.silu_and_mul_absmax(cache2, a2s, cache1.view(T * k, N))
        else:
            a2s = a2_scale.reshape(-1)[:1].float()
            ops.silu_and_mul(cache2, cache1.view(T * k, N))
        cache3 = torch.empty((T, k, w2.shape[1]), device=dev, dtype=dt)
        invoke_fused_moe_kernel(cache2, w2, cache3, a2s, w2_scale, topk_weights.contiguous(), sorted_ids, expert_ids,
                                num_post, True, 1, block_m, True, False)
        out = hidden_states if inplace else torch.empty_like(hidden_states)
        ops.moe_sum(cache3, out)
        return out
    a1q, a1s = ops.scaled_fp8_quant(hidden_states, a1_scale, use_per_token_if_dynamic=per_channel_quant)
    invoke_fused_moe_kernel(a1q, w1, cache1, a1s, w1_scale, None, sorted_ids, expert_ids, num_post, False, k,
                            block_m, True, per_channel_quant)
    if per_channel_quant and a2_scale is None:
        a2q, a2s = ops.silu_and_mul_dynamic_per_token_quant(cache1.view(T * k, N))
    else:
        cache2 = torch.empty((T * k, N // 2), device=dev, dtype=dt)
        ops.silu_and_mul(cache2, cache1.view(T * k, N))
        a2q, a2s = ops.scaled_fp8_quant(cache2, a2_scale, use_per_token_if_dynamic=per_channel_quant)
    cache3 = torch.empty((T, k, w2.shape[1]), device=dev, dtype=dt)
    invoke_fused_moe_kernel(a2q, w2, cache3, a2s, w2_scale, topk_weights.contiguous(), sorted_ids, expert_ids,
                            num_post, True, 1, block_m, True, per_channel_quant)
    out = hidden_states if inplace else torch.empty_like(hidden_states)
    ops.moe_sum(cache3, out)
    return out


def fused_moe(hidden_states: torch.Tensor, w1: torch.Tensor, w2: torch.Tensor, gating_output: torch.Tensor,
              topk: int, renormalize: bool, inplace: bool = False, **kwargs) -> torch.Tensor:
    """fused_moe.py:1566-1670: routing + experts."""
    topk_weights, topk_ids, _ = fused_topk(hidden_states, gating_output, topk, renormalize)
    return fused_experts(hidden_states, w1, w2, topk_weights, topk_ids, inplace=inplace, **kwargs)


# ---- dummy-weight helpers for the decode-step driver ---------------------------------------------------
def make_dummy_experts(E: int, hidden: int, inter: int, device, dtype, quant: str, gen: torch.Generator,
                       inter_range=None):
    """Random fp8 expert weights with per-tensor (per-expert) scales, as Mixtral FP8 checkpoints carry.  `inter` is
    the UNSHARDED intermediate size; `inter_range` = (start, stop) keeps this tensor-parallel rank's slice of it
    (w13 rows of gate and up, w2 columns) of the one seeded weight set, scales taken over the full expert."""
    if quant != "fp8":
        raise RuntimeError("the MoE path is built for fp8 w8a8 experts")
    i0, i1 = (0, inter) if inter_range is None else inter_range

    def mk(n, k, rows=None, cols=None):
        q = torch.empty(E, n if rows is None else sum(b - a for a, b in rows), k if cols is None else cols[1] - cols[0],
                        device=device, dtype=FP8)
        scale = torch.empty(E, device=device, dtype=torch.float32)
        for e in range(E):   # one expert at a time: the unsharded bf16 / fp32 temporaries stay small
            w = torch.empty(n, k, device=device, dtype=torch.bfloat16).normal_(0.0, 0.02, generator=gen)
            sc = w.abs().amax().float().clamp_(min=1e-8) / 448.0
            if rows is not None:
                w = torch.cat([w[a:b] for a, b in rows], dim=0)
            if cols is not None:
                w = w[:, cols[0]:cols[1]]
            q[e] = (w.float() / sc).clamp_(-448.0, 448.0).to(FP8)
            scale[e] = sc
        return q, scale
    w13, s13 = mk(2 * inter, hidden, rows=[(i0, i1), (inter + i0, inter + i1)])
    w2, s2 = mk(hidden, inter, cols=(i0, i1))
    return {"w13": w13, "w13_scale": s13, "w2": w2, "w2_scale": s2}


def moe_forward(hidden: torch.Tensor, w_gate: torch.Tensor, experts: dict, top_k: int, renormalize: bool = True):
    """FusedMoE.forward_impl (fused_moe/layer.py:1482-1533) minus the all-reduce (done by the caller):
    router logits via the library GEMM (replicated, unquantised gate), then fused_moe."""
    E, H = w_gate.shape
    if E <= 32 and H % 16 == 0 and H <= 8192 and hidden.dtype in (torch.float16, torch.bfloat16) and \
            os.environ.get("VB200_FUSED_ROUTER", "1") != "0":
        # gate GEMM + fused_topk + the abs-max of `hidden` (the experts' dynamic a1 scale) in ONE launch
        scales = torch.zeros(2, dtype=torch.float32, device=hidden.device)
        topk_weights, topk_ids = router_topk(hidden, w_gate, top_k, renormalize, absmax_scale=scales[0:1])
        return fused_experts(hidden, experts["w13"], experts["w2"], topk_weights, topk_ids, use_fp8_w8a8=True,
                             w1_scale=experts["w13_scale"], w2_scale=experts["w2_scale"], dynamic_scales=scales)
    router_logits = torch.matmul(hidden, w_gate.t())
    return fused_moe(hidden, experts["w13"], experts["w2"], router_logits, top_k, renormalize,
                     use_fp8_w8a8=True, w1_scale=experts["w13_scale"], w2_scale=experts["w2_scale"])
