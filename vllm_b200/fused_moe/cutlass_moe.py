"""CUTLASS-style fused MoE: the explicit permute -> grouped GEMM -> activation -> quant ->
grouped GEMM -> un-permute -> weight-and-reduce pipeline of
/root/reference/vllm/model_executor/layers/fused_moe/cutlass_moe.py:25-215 (`run_cutlass_moe_fp8`) and :324-409
(`cutlass_moe_fp8`), on this package's ops: `get_cutlass_moe_mm_data`, `shuffle_rows`, `cutlass_moe_mm` (the tcgen05
grouped GEMM in ragged mode), `silu_and_mul`, `scaled_fp8_quant`, `moe_sum`.  Same argument names / meaning /
assertions as the reference; the modular-kernel classes around it (FusedMoEModularKernel, PrepareAndFinalize) are
host orchestration and are folded into `cutlass_moe_fp8`.
"""
from typing import Callable, Optional

import torch

from .. import _custom_ops as ops


def _resize_cache(x: torch.Tensor, v: tuple) -> torch.Tensor:
    """fused_moe/utils.py `_resize_cache`: a view of the first prod(v) elements of a workspace."""
    n = 1
    for d in v:
        n *= d
    assert n <= x.numel(), f"workspace too small: {x.numel()} < {n}"
    return x.flatten()[:n].view(*v)


def run_cutlass_moe_fp8(
    output: torch.Tensor,
    hidden_states: torch.Tensor,
    w1: torch.Tensor,
    w2: torch.Tensor,
    topk_ids: torch.Tensor,
    activation_callable: Callable,
    global_num_experts: int,
    expert_map: Optional[torch.Tensor],
    w1_scale: Optional[torch.Tensor],
    w2_scale: Optional[torch.Tensor],
    a1q_scale: Optional[torch.Tensor],
    a2_scale: Optional[torch.Tensor],
    workspace13: torch.Tensor,
    workspace2: torch.Tensor,
    expert_num_tokens: Optional[torch.Tensor],
    out_dtype: torch.dtype,
    per_act_token: bool,
    per_out_ch: bool,
    use_batched_format: bool,
) -> None:
    """cutlass_moe.py:25-215.  w1 [E, 2N, K], w2 [E, K, N] fp8 (K resp. N contiguous); hidden_states already fp8.
    Non-batched: hidden_states [M, K], output [M*topk, K] in (token, k) order.  Batched: hidden_states
    [E, padded_M, K], output [E, padded_M, K]."""
    a1q = hidden_states
    assert w1_scale is not None
    assert w2_scale is not None
    assert w1.dtype == torch.float8_e4m3fn
    assert w2.dtype == torch.float8_e4m3fn
    assert a1q.size(-1) == w1.size(2), "Hidden size mismatch w1"
    assert w1.size(1) == w2.size(2) * 2, "Hidden size mismatch w2"
    assert w1_scale.dim() == 1 or w1_scale.size(1) == 1 or w1_scale.shape[1] == w1.size(1), "W1 scale shape mismatch"
    assert w2_scale.dim() == 1 or w2_scale.size(1) == 1 or w2_scale.shape[1] == w2.size(1), "W2 scale shape mismatch"
    assert w1.size(0) == w2.size(0), "Expert number mismatch"
    assert a1q_scale is None or a1q_scale.dim() == 0 or a1q_scale.size(0) == 1 or \
        a1q_scale.size(0) == a1q.shape[0], "Input scale shape mismatch"
    assert w1.size(0) == w1_scale.size(0), "w1 scales expert number mismatch"
    assert w1.size(0) == w2_scale.size(0), "w2 scales expert number mismatch"
    assert a2_scale is None or a2_scale.dim() == 0 or a2_scale.size(0) == 1 or \
        a2_scale.size(0) == a1q.shape[0], "Intermediate scale shape mismatch"
    assert out_dtype in [torch.half, torch.bfloat16], "Invalid output dtype"
    if expert_map is not None:
        assert expert_num_tokens is None

    M = a1q.size(0)          # non-batched M
    padded_M = a1q.size(1)   # batched M
    _, K, N = w2.shape
    device = a1q.device
    assert w1.size(2) == K
    assert global_num_experts != -1
    assert a1q_scale is not None

    if expert_map is not None:
        local_topk_ids = torch.where(expert_map[topk_ids] != -1, expert_map[topk_ids], -1)
    else:
        local_topk_ids = topk_ids
    local_topk_ids = local_topk_ids.to(torch.int32).contiguous()
    topk = local_topk_ids.size(1)
    local_E = w1.size(0)
    c_map = None

    if use_batched_format:
        assert expert_num_tokens is not None
        expert_offsets = torch.empty((local_E), dtype=torch.int32, device=device)
        problem_sizes1 = torch.empty((local_E, 3), dtype=torch.int32, device=device)
        problem_sizes2 = torch.empty((local_E, 3), dtype=torch.int32, device=device)
        ops.get_cutlass_pplx_moe_mm_data(expert_offsets, problem_sizes1, problem_sizes2,
                                         expert_num_tokens.to(torch.int32), local_E, padded_M, N, K)
        w1_scale = w1_scale.reshape(w1_scale.size(0), -1)
        w2_scale = w2_scale.reshape(w2_scale.size(0), -1)
        a1q = a1q.reshape(-1, a1q.size(2))
        a1q_scale = a1q_scale.reshape(-1, a1q_scale.size(2)).contiguous()
    else:
        expert_offsets = torch.empty((global_num_experts + 1), dtype=torch.int32, device=device)
        problem_sizes1 = torch.empty((global_num_experts, 3), dtype=torch.int32, device=device)
        problem_sizes2 = torch.empty((global_num_experts, 3), dtype=torch.int32, device=device)
        # with an expert_map not every a_map entry is written: zero them (cutlass_moe.py:132-140)
        a_map = (torch.zeros if expert_map is not None else torch.empty)(
            (local_topk_ids.numel()), dtype=torch.int32, device=device)
        c_map = torch.empty((local_topk_ids.numel()), dtype=torch.int32, device=device)
        ops.get_cutlass_moe_mm_data(local_topk_ids, expert_offsets, problem_sizes1, problem_sizes2, a_map, c_map,
                                    global_num_experts, N, K)
        a1q = ops.shuffle_rows(a1q, a_map)                                    # _fp8_perm(a1q, a_map)
        a1q_scale = ops.shuffle_rows(a1q_scale.view(M, 1), a_map) if per_act_token else a1q_scale
        expert_offsets = expert_offsets[:-1]

    ab_strides1 = torch.full((w1.size(0), ), K, device=device, dtype=torch.int64)
    c_strides1 = torch.full((w1.size(0), ), 2 * N, device=device, dtype=torch.int64)
    ab_strides2 = torch.full((w1.size(0), ), N, device=device, dtype=torch.int64)
    c_strides2 = torch.full((w1.size(0), ), K, device=device, dtype=torch.int64)

    rows = local_E * padded_M if use_batched_format else M * topk
    c1 = _resize_cache(workspace13, (rows, N * 2))
    c2 = _resize_cache(workspace2, (rows, N))
    c3 = _resize_cache(workspace13, (rows, K))

    if not per_act_token and (expert_map is not None or use_batched_format):
        # rows no expert owns would otherwise feed garbage into the per-tensor scale (cutlass_moe.py:180-184)
        c1.fill_(0)

    ops.cutlass_moe_mm(c1, a1q, w1, a1q_scale, w1_scale, expert_offsets, problem_sizes1, ab_strides1, ab_strides1,
                       c_strides1, per_act_token, per_out_ch)
    activation_callable(c2, c1)
    a2q, a2q_scale = ops.scaled_fp8_quant(c2, a2_scale, use_per_token_if_dynamic=per_act_token)
    if expert_map is not None:
        c3.fill_(0)
    ops.cutlass_moe_mm(c3, a2q, w2, a2q_scale, w2_scale, expert_offsets, problem_sizes2, ab_strides2, ab_strides2,
                       c_strides2, per_act_token, per_out_ch)

    if use_batched_format:
        output.copy_(c3.reshape(local_E, padded_M, K), non_blocking=True)
    else:
        # c3[c_map]; ids this rank does not own map to row `total routed rows`, which the fill above zeroed
        # (moe_data.cu:92-99)
        output.copy_(ops.shuffle_rows(c3, c_map).view(M * topk, K), non_blocking=True)


def _activation(activation: str, out: torch.Tensor, inp: torch.Tensor) -> None:
    """FusedMoEPermuteExpertsUnpermute.activation (modular_kernel.py): silu -> torch.ops._C.silu_and_mul."""
    if activation != "silu":
        raise ValueError(f"Unsupported FusedMoe activation: {activation}")
    ops.silu_and_mul(out, inp)


def cutlass_moe_fp8(
    a: torch.Tensor,
    w1_q: torch.Tensor,
    w2_q: torch.Tensor,
    topk_weights: torch.Tensor,
    topk_ids: torch.Tensor,
    w1_scale: torch.Tensor,
    w2_scale: torch.Tensor,
    per_act_token: Optional[bool] = None,
    activation: str = "silu",
    a1_scale: Optional[torch.Tensor] = None,
    a2_scale: Optional[torch.Tensor] = None,
    expert_map: Optional[torch.Tensor] = None,
    apply_router_weight_on_input: bool = False,
    global_num_experts: int = -1,
) -> torch.Tensor:
    """cutlass_moe.py:324-409.  a [M, K] bf16/f16; w1_q [E, 2N, K], w2_q [E, K, N] fp8; returns [M, K] in a.dtype.
    The reference wraps `CutlassExpertsFp8` in a FusedMoEModularKernel with MoEPrepareAndFinalizeNoEP: prepare =
    quantise `a` (static per-tensor when a1_scale is given, else dynamic per token / per tensor), finalize =
    TopKWeightAndReduceContiguous (fp32 weights multiplied in place, then moe_sum)."""
    if per_act_token is None:
        per_act_token = a1_scale.numel() != 1 if a1_scale is not None else (
            a2_scale.numel() != 1 if a2_scale is not None else False)
    per_out_ch = w1_scale.numel() != w1_q.size(0)
    num_experts = global_num_experts if global_num_experts != -1 else w1_q.size(0)
    M, K = a.shape
    N = w2_q.size(2)
    topk = topk_ids.size(1)
    if apply_router_weight_on_input:
        assert topk == 1, "apply_router_weight_on_input is only implemented for topk=1"
        a = a * topk_weights.to(a.dtype)
    a1q, a1q_scale = ops.scaled_fp8_quant(a, a1_scale, use_per_token_if_dynamic=per_act_token)
    workspace13 = torch.empty(M * topk * max(2 * N, K), dtype=a.dtype, device=a.device)
    workspace2 = torch.empty(M * topk * N, dtype=a.dtype, device=a.device)
    fused_out = torch.empty(M * topk, K, dtype=a.dtype, device=a.device)
    run_cutlass_moe_fp8(fused_out, a1q, w1_q, w2_q, topk_ids, lambda o, i: _activation(activation, o, i),
                        num_experts, expert_map, w1_scale, w2_scale, a1q_scale, a2_scale, workspace13, workspace2,
                        None, a.dtype, per_act_token, per_out_ch, False)
    fused_out = fused_out.view(M, topk, K)
    if not apply_router_weight_on_input:
        fused_out.mul_(topk_weights.view(M, -1, 1))   # fp32 product, rounded once (topk_weight_and_reduce.py:89)
    out = torch.empty(M, K, dtype=a.dtype, device=a.device)
    ops.moe_sum(fused_out, out)
    return out
