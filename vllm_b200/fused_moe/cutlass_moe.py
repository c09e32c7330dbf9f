"""Permuted ("CUTLASS-style") fused MoE on B200.

Interface kept from /root/reference/vllm/model_executor/layers/fused_moe/cutlass_moe.py: `cutlass_moe_fp8` (:324-409,
what a model calls) and `run_cutlass_moe_fp8` (:25-215, the 19-argument entry the modular-kernel layer calls with
pre-quantised activations and caller-owned workspaces).  The pipeline behind them is this package's own:

    route     ONE single-CTA stable counting sort -> first row of every expert, the (m_e, N, K) problem list, and the
              two row maps (sorted row -> token, (token, j) -> sorted row)          [get_cutlass_moe_mm_data]
    front     gather + fp8 quantise in one launch (per-token dynamic or static scales)   [moe_gather_quant_fp8]
    GEMM 1    tcgen05 grouped GEMM in ragged mode: A is read by TMA straight from the sorted rows, experts own
              contiguous row ranges; no pointer arrays, no stride tensors, no row padding    [cutlass_moe_mm]
    middle    silu_and_mul + per-token fp8 quantise in one launch          [silu_and_mul_dynamic_per_token_quant]
    GEMM 2    same kernel
    back      un-permute + routing-weight multiply + top-k sum in one launch         [moe_unpermute_reduce]

6 launches where the reference runs ~15 (5 routing kernels, quant, 2 row shuffles, 2 GEMMs with 2 pointer-array
set-up kernels, activation, quant, shuffle, mul, sum).  Rounding points are the reference's (each fused launch is
bit-identical to the ops it replaces), so parity is checked against the same oracle as before.
Granularities without a fused kernel (dynamic PER-TENSOR activation scales need a global amax first) fall back to the
stand-alone ops of this package -- still all on the device, still graph-capturable.
"""
from dataclasses import dataclass
from typing import Callable, Optional

import torch

from .. import _custom_ops as ops

_FP8 = torch.float8_e4m3fn


@dataclass
class _Routing:
    """Device-side routing plan of one call (all int32, written by one kernel)."""
    first_row: torch.Tensor        # [E + 1] first sorted row of every expert; [-1] = number of routed rows
    problems1: torch.Tensor        # [E, 3] GEMM-1 problem sizes
    problems2: torch.Tensor        # [E, 3] GEMM-2 problem sizes
    src_token: Optional[torch.Tensor]   # [rows] sorted row -> token            (None: batched format)
    dst_row: Optional[torch.Tensor]     # [rows] (token, j) flat index -> sorted row
    rows: int                      # capacity of the sorted-row space

    @property
    def num_routed(self) -> torch.Tensor:
        return self.first_row[-1:]


def _local_ids(topk_ids: torch.Tensor, expert_map: Optional[torch.Tensor]) -> torch.Tensor:
    """Global expert ids -> this rank's ids (-1 = not here).  `expert_map[g]` is already -1 for foreign experts."""
    if expert_map is None:
        return topk_ids if topk_ids.dtype == torch.int32 and topk_ids.is_contiguous() else \
            topk_ids.to(torch.int32).contiguous()
    return expert_map.to(torch.int32)[topk_ids.long()].contiguous()


def _route(ids: torch.Tensor, num_experts: int, n: int, k: int) -> _Routing:
    dev, rows = ids.device, ids.numel()
    # one allocation for the five outputs of the routing kernel
    meta = torch.empty(num_experts + 1 + 6 * num_experts + 2 * rows, dtype=torch.int32, device=dev)
    o = num_experts + 1
    first_row = meta[:o]
    p1 = meta[o:o + 3 * num_experts].view(num_experts, 3)
    p2 = meta[o + 3 * num_experts:o + 6 * num_experts].view(num_experts, 3)
    src = meta[o + 6 * num_experts:o + 6 * num_experts + rows]
    dst = meta[o + 6 * num_experts + rows:]
    ops.get_cutlass_moe_mm_data(ids, first_row, p1, p2, src, dst, num_experts, n, k)
    return _Routing(first_row, p1, p2, src, dst, rows)


def _route_batched(expert_num_tokens: torch.Tensor, local_E: int, padded_m: int, n: int, k: int) -> _Routing:
    dev = expert_num_tokens.device
    meta = torch.empty(7 * local_E, dtype=torch.int32, device=dev)
    first_row = meta[:local_E]
    p1 = meta[local_E:4 * local_E].view(local_E, 3)
    p2 = meta[4 * local_E:].view(local_E, 3)
    ops.get_cutlass_pplx_moe_mm_data(first_row, p1, p2, expert_num_tokens.to(torch.int32), local_E, padded_m, n, k)
    return _Routing(first_row, p1, p2, None, None, local_E * padded_m)


def _carve(ws: torch.Tensor, rows: int, cols: int) -> torch.Tensor:
    need = rows * cols
    if ws.numel() < need:
        raise RuntimeError(f"cutlass_moe: workspace holds {ws.numel()} elements, {need} needed")
    return ws.view(-1)[:need].view(rows, cols)


# cutlass_moe_mm keeps the reference schema (three per-expert stride tensors).  The tcgen05 kernel addresses the
# contiguous layouts directly and never reads them; one cached int64 tensor per device stands in for all three.
_STRIDE_STANDIN = {}


def _stride_standin(dev: torch.device, experts: int) -> torch.Tensor:
    t = _STRIDE_STANDIN.get(dev)
    if t is None or t.numel() < experts:
        t = _STRIDE_STANDIN[dev] = torch.zeros(max(experts, 64), dtype=torch.int64, device=dev)
    return t[:experts]


def _grouped_linear(out, a_q, a_scale, w, w_scale, route: _Routing, problems, per_act_token, per_out_ch):
    st = _stride_standin(a_q.device, w.size(0))
    ops.cutlass_moe_mm(out, a_q, w, a_scale, w_scale, route.first_row[:w.size(0)], problems, st, st, st,
                       per_act_token, per_out_ch)


def _validate(a_q, w1, w2, w1_scale, w2_scale, a1_scale, a2_scale, out_dtype, expert_map, expert_num_tokens):
    """Argument contract of run_cutlass_moe_fp8 (cutlass_moe.py:46-76), stated once."""
    def bad(msg):
        raise AssertionError(f"run_cutlass_moe_fp8: {msg}")
    if w1_scale is None or w2_scale is None or a1_scale is None:
        bad("w1_scale, w2_scale and the activation scale are required")
    if w1.dtype != _FP8 or w2.dtype != _FP8:
        bad("expert weights must be float8_e4m3fn")
    E, two_n, k = w1.shape
    if w2.size(0) != E or w1_scale.size(0) != E or w2_scale.size(0) != E:
        bad("expert count differs between w1, w2 and their scales")
    if a_q.size(-1) != k or w2.size(1) != k:
        bad(f"hidden size mismatch: activations {a_q.size(-1)}, w1 {k}, w2 {w2.size(1)}")
    if two_n != 2 * w2.size(2):
        bad(f"w1 holds {two_n} rows per expert, expected 2 x {w2.size(2)}")
    for name, s, w in (("w1_scale", w1_scale, w1), ("w2_scale", w2_scale, w2)):
        if not (s.dim() == 1 or s.size(1) == 1 or s.size(1) == w.size(1)):
            bad(f"{name} must be per expert or per output channel")
    for name, s in (("a1q_scale", a1_scale), ("a2_scale", a2_scale)):
        if s is not None and not (s.dim() == 0 or s.size(0) == 1 or s.size(0) == a_q.size(0)):
            bad(f"{name} must be per tensor or per row")
    if out_dtype not in (torch.half, torch.bfloat16):
        bad("output dtype must be float16 or bfloat16")
    if expert_map is not None and expert_num_tokens is not None:
        bad("expert_map and expert_num_tokens are mutually exclusive")


def _expert_chain(a_q, a_scale, w1, w2, w1_scale, w2_scale, a2_scale, route: _Routing, c1, c3, act_buf,
                  activation: Optional[Callable], per_act_token: bool, per_out_ch: bool, zero_unowned: bool,
                  zero_out: bool = False):
    """GEMM 1 -> activation -> fp8 -> GEMM 2 over the sorted rows.  `activation` None = silu_and_mul (fusable).
    zero_out: rows of c3 no expert owns must read as zero afterwards (c3 may alias c1, so the fill sits between the
    activation and GEMM 2)."""
    if zero_unowned and not per_act_token:
        c1.zero_()   # rows no expert owns must not leak into the per-tensor amax of the second quantisation
    _grouped_linear(c1, a_q, a_scale, w1, w1_scale, route, route.problems1, per_act_token, per_out_ch)
    if activation is None and per_act_token and a2_scale is None:
        h_q, h_scale = ops.silu_and_mul_dynamic_per_token_quant(c1)
    else:
        if activation is None:
            ops.silu_and_mul(act_buf, c1)
        else:
            activation(act_buf, c1)
        h_q, h_scale = ops.scaled_fp8_quant(act_buf, a2_scale, use_per_token_if_dynamic=per_act_token)
    if zero_out:
        c3.zero_()
    _grouped_linear(c3, h_q, h_scale, w2, w2_scale, route, route.problems2, per_act_token, per_out_ch)


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
    """Reference entry point (cutlass_moe.py:25-215), same arguments and results.  hidden_states is ALREADY fp8
    ([M, K], or [E, padded_M, K] in the batched format) with its scale in a1q_scale; w1 [E, 2N, K], w2 [E, K, N];
    `output` receives the un-weighted expert outputs: [M*topk, K] in (token, j) order, or [E, padded_M, K]."""
    _validate(hidden_states, w1, w2, w1_scale, w2_scale, a1q_scale, a2_scale, out_dtype, expert_map,
              expert_num_tokens)
    if global_num_experts == -1:
        raise AssertionError("run_cutlass_moe_fp8: global_num_experts must be given")
    local_E, _, K = w1.shape
    N = w2.size(2)
    if use_batched_format:
        if expert_num_tokens is None:
            raise AssertionError("run_cutlass_moe_fp8: the batched format needs expert_num_tokens")
        padded_M = hidden_states.size(1)
        route = _route_batched(expert_num_tokens, local_E, padded_M, N, K)
        a_q = hidden_states.reshape(-1, K)
        a_scale = a1q_scale.reshape(-1, a1q_scale.size(-1)).contiguous() if a1q_scale.dim() > 1 else a1q_scale
        w1_scale, w2_scale = w1_scale.reshape(local_E, -1), w2_scale.reshape(local_E, -1)
    else:
        M = hidden_states.size(0)
        route = _route(_local_ids(topk_ids, expert_map), global_num_experts, N, K)
        a_q = ops.shuffle_rows(hidden_states, route.src_token)
        a_scale = ops.shuffle_rows(a1q_scale.view(M, 1), route.src_token) if per_act_token else a1q_scale
    c1 = _carve(workspace13, route.rows, 2 * N)
    act = _carve(workspace2, route.rows, N)
    c3 = _carve(workspace13, route.rows, K)   # GEMM 1's output is dead once the activation has run
    sparse = expert_map is not None or use_batched_format
    # the hook this package passes is tagged so that it can be fused with the quantisation; a foreign callable runs as is
    fusable = getattr(activation_callable, "_vb200_silu_and_mul", False)
    # (token, j) pairs of foreign experts map to row `num_routed` of c3, which therefore has to read as zero
    _expert_chain(a_q, a_scale, w1, w2, w1_scale, w2_scale, a2_scale, route, c1, c3, act,
                  None if fusable else activation_callable, per_act_token, per_out_ch, sparse,
                  zero_out=expert_map is not None)
    if use_batched_format:
        output.copy_(c3.view(local_E, -1, K), non_blocking=True)
    else:
        ops.shuffle_rows_into(output.view(route.rows, K), c3, route.dst_row)


def _silu_and_mul_hook(out: torch.Tensor, inp: torch.Tensor) -> None:
    ops.silu_and_mul(out, inp)


_silu_and_mul_hook._vb200_silu_and_mul = True   # lets the expert chain fuse it with the quantisation


def _activation_hook(activation: str) -> Callable:
    if activation != "silu":
        raise ValueError(f"Unsupported FusedMoe activation: {activation}")
    return _silu_and_mul_hook


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
    """cutlass_moe.py:324-409: a [M, K] bf16/f16, w1_q [E, 2N, K], w2_q [E, K, N] fp8 -> [M, K] in a.dtype.
    The reference composes MoEPrepareAndFinalizeNoEP (quantise), CutlassExpertsFp8 (run_cutlass_moe_fp8) and
    TopKWeightAndReduceContiguous; here the six launches of the module docstring."""
    if per_act_token is None:
        per_act_token = next((s.numel() != 1 for s in (a1_scale, a2_scale) if s is not None), False)
    per_out_ch = w1_scale.numel() != w1_q.size(0)
    E = global_num_experts if global_num_experts != -1 else w1_q.size(0)
    act_hook = _activation_hook(activation)
    M, K = a.shape
    N = w2_q.size(2)
    topk = topk_ids.size(1)
    if apply_router_weight_on_input:
        if topk != 1:
            raise AssertionError("apply_router_weight_on_input is only implemented for topk=1")
        a = a * topk_weights.to(a.dtype)
    route = _route(_local_ids(topk_ids, expert_map), E, N, K)
    bound = route.num_routed if expert_map is not None else None
    # ---- front: gather + quantise -------------------------------------------------------------------------------
    if a1_scale is not None and a1_scale.numel() == 1:
        a_q, a_scale = ops.moe_gather_quant_fp8(a, route.src_token, bound, scale=a1_scale.view(1).float())
    elif a1_scale is None and per_act_token:
        a_q, a_scale = ops.moe_gather_quant_fp8(a, route.src_token, bound)
    else:
        # dynamic per-tensor: the scale is a global amax, so quantise first and move rows afterwards
        q, a_scale = ops.scaled_fp8_quant(a, a1_scale, use_per_token_if_dynamic=per_act_token)
        a_q = ops.shuffle_rows(q, route.src_token)
        if per_act_token:
            a_scale = ops.shuffle_rows(a_scale.view(M, 1), route.src_token)
    # ---- experts ------------------------------------------------------------------------------------------------
    ws = torch.empty(route.rows * max(2 * N, K), dtype=a.dtype, device=a.device)
    c1, c3 = _carve(ws, route.rows, 2 * N), _carve(ws, route.rows, K)
    fused_act = per_act_token and a2_scale is None
    act = None if fused_act else torch.empty(route.rows, N, dtype=a.dtype, device=a.device)
    _validate(a_q, w1_q, w2_q, w1_scale, w2_scale, a_scale, a2_scale, a.dtype, expert_map, None)
    _expert_chain(a_q, a_scale, w1_q, w2_q, w1_scale, w2_scale, a2_scale, route, c1, c3, act,
                  None if act_hook is _silu_and_mul_hook else act_hook, per_act_token, per_out_ch,
                  expert_map is not None)
    # ---- back: un-permute, weight, reduce -----------------------------------------------------------------------
    weights = None if apply_router_weight_on_input else topk_weights.float()
    return ops.moe_unpermute_reduce(c3, route.dst_row, weights, route.num_routed, M, topk)
