"""ORACLE (test infrastructure only): fused-MoE routing and experts.

topk_softmax          /root/reference/csrc/moe/topk_softmax_kernels.cu:288-391: fp32 softmax, k rounds of
                      arg-max with LOWEST-index tie-break (:336-358); token_expert_indices = k*T + row (:373)
moe_align_block_size  /root/reference/csrc/moe/moe_align_sum_kernels.cu:17-107,204-275 in the stable order of
                      the reference test's golden (tests/kernels/moe/test_moe_align_block_size.py:89-167)
moe_sum               moe_align_sum_kernels.cu:109-123: accumulate in the tensor dtype, in j order
fused_experts         vllm/model_executor/layers/fused_moe/fused_moe.py:1349-1563 + Triton kernel :267-475
                      (routed weight multiplied BEFORE the scales, :455-468); mirrors torch_experts
                      (tests/kernels/utils.py:1060-1152)
"""
import numpy as np
import torch

from .quant import dynamic_per_token_scaled_fp8_quant, dynamic_scaled_fp8_quant, silu_and_mul, static_scaled_fp8_quant

_f32 = torch.float32


def topk_softmax(gating_output: torch.Tensor, topk: int):
    """-> (weights fp32 [T,k], ids int32 [T,k], token_expert_indices int32 [T,k])."""
    g = gating_output.to(_f32)
    T, E = g.shape
    p = torch.softmax(g, dim=-1).numpy().copy()
    w = np.zeros((T, topk), np.float32)
    ids = np.zeros((T, topk), np.int32)
    tei = np.zeros((T, topk), np.int32)
    for k in range(topk):
        idx = p.argmax(axis=1)  # numpy argmax returns the FIRST maximum -> lowest index
        w[:, k] = p[np.arange(T), idx]
        ids[:, k] = idx
        tei[:, k] = k * T + np.arange(T)
        p[np.arange(T), idx] = -1.0
    return torch.from_numpy(w), torch.from_numpy(ids), torch.from_numpy(tei)


def moe_align_block_size(topk_ids: torch.Tensor, block_size: int, num_experts: int):
    """-> (sorted_token_ids [numel + E*(block-1)], expert_ids [ceil(that/block)], num_tokens_post_pad [1])."""
    flat = topk_ids.flatten().numpy().astype(np.int64)
    numel = flat.size
    max_padded = numel + num_experts * (block_size - 1)
    sorted_ids = np.full(max_padded, numel, np.int32)
    expert_ids = np.zeros((max_padded + block_size - 1) // block_size, np.int32)
    pos = blk = 0
    for e in range(num_experts):
        tok = np.nonzero(flat == e)[0]  # ascending flat index = stable order
        if tok.size == 0:
            continue
        sorted_ids[pos:pos + tok.size] = tok
        nb = (tok.size + block_size - 1) // block_size
        expert_ids[blk:blk + nb] = e
        pos += nb * block_size
        blk += nb
    return torch.from_numpy(sorted_ids), torch.from_numpy(expert_ids), torch.tensor([pos], dtype=torch.int32)


def moe_sum(x: torch.Tensor) -> torch.Tensor:
    """[T, k, H] -> [T, H]; k in {2,3,4}: sequential adds in x.dtype starting from 0; else fp32 sum."""
    T, k, H = x.shape
    if k in (2, 3, 4):
        acc = torch.zeros(T, H, dtype=x.dtype)
        for j in range(k):
            acc = (acc.to(_f32) + x[:, j].to(_f32)).to(x.dtype)
        return acc
    return x.to(_f32).sum(1).to(x.dtype)


def _quant_act(x, a_scale, per_token):
    if a_scale is not None:
        return static_scaled_fp8_quant(x, float(a_scale)), a_scale.to(_f32).reshape(1, 1)
    if per_token:
        return dynamic_per_token_scaled_fp8_quant(x)
    q, s = dynamic_scaled_fp8_quant(x)
    return q, s.reshape(1, 1)


def grouped_gemm_fp8(a_q, a_scale, w_q, w_scale, topk_ids, top_k_of_a, routed_weight=None,
                     out_dtype=torch.bfloat16):
    """One invoke_fused_moe_kernel call (fused_moe.py:478-647).  a_q fp8 [rows, K] where row =
    token_id // top_k_of_a; w_q fp8 [E, N, K]; w_scale [E] or [E, N]; a_scale [1,1] or [rows,1].
    -> c [numel(topk_ids), N]:  cast((acc * routed_weight) * a_scale * b_scale)."""
    flat = topk_ids.flatten().long()
    n_tok = flat.numel()
    E, N, K = w_q.shape
    c = torch.zeros(n_tok, N, dtype=out_dtype)
    rows = torch.arange(n_tok) // top_k_of_a
    for e in range(E):
        sel = torch.nonzero(flat == e).flatten()
        if sel.numel() == 0:
            continue
        acc = (a_q[rows[sel]].to(torch.float64) @ w_q[e].to(torch.float64).t()).to(_f32)
        if routed_weight is not None:
            acc = acc * routed_weight.flatten()[sel].to(_f32)[:, None]
        sa = a_scale if a_scale.numel() == 1 else a_scale[rows[sel]]
        sb = w_scale[e].to(_f32).reshape(1, -1)
        c[sel] = (acc * sa.to(_f32) * sb).to(out_dtype)
    return c


def fused_experts_fp8(hidden, w1_q, w2_q, w1_scale, w2_scale, topk_weights, topk_ids, a1_scale=None,
                      a2_scale=None, per_channel_quant=False):
    """fused_experts(use_fp8_w8a8=True) -> [T, H] in hidden.dtype."""
    T, H = hidden.shape
    k = topk_ids.shape[1]
    a1q, a1s = _quant_act(hidden, a1_scale, per_channel_quant)
    c1 = grouped_gemm_fp8(a1q, a1s, w1_q, w1_scale, topk_ids, k, None, hidden.dtype)        # [T*k, 2I]
    c2 = silu_and_mul(c1)                                                                     # [T*k, I]
    a2q, a2s = _quant_act(c2, a2_scale, per_channel_quant)
    c3 = grouped_gemm_fp8(a2q, a2s, w2_q, w2_scale, topk_ids, 1, topk_weights, hidden.dtype)  # [T*k, H]
    return moe_sum(c3.reshape(T, k, H))


def fused_experts_bf16(hidden, w1, w2, topk_weights, topk_ids):
    """Unquantised path (torch_experts, tests/kernels/utils.py:1103-1106) with the kernel's rounding points."""
    T, H = hidden.shape
    k = topk_ids.shape[1]
    flat = topk_ids.flatten().long()
    rows = torch.arange(T * k) // k
    c3 = torch.zeros(T * k, H, dtype=hidden.dtype)
    for e in range(w1.shape[0]):
        sel = torch.nonzero(flat == e).flatten()
        if sel.numel() == 0:
            continue
        t1 = (hidden[rows[sel]].to(_f32) @ w1[e].to(_f32).t()).to(hidden.dtype)
        t2 = silu_and_mul(t1)
        acc = t2.to(_f32) @ w2[e].to(_f32).t()
        c3[sel] = (acc * topk_weights.flatten()[sel].to(_f32)[:, None]).to(hidden.dtype)
    return moe_sum(c3.reshape(T, k, H))


# ---- CUTLASS-style fused-MoE API ------------------------------------------------------------
def get_cutlass_moe_mm_data(topk_ids: torch.Tensor, num_experts: int, n: int, k: int, blockscale: bool = False):
    """csrc/quantization/cutlass_w8a8/moe/moe_data.cu:11-159.  Returns (expert_offsets [E+1], problem_sizes1 [E,3],
    problem_sizes2 [E,3], input_permutation, output_permutation[, blockscale_offsets]) as int32 tensors; the
    permutation is the stable member of the reference's (atomics-ordered) family; ids == -1 map to row `total`
    (:92-99); the first two problem-size entries are swapped for <= 64 routed rows unless blockscale (:9,121-122)."""
    flat = topk_ids.flatten().long()
    numel, topk = flat.numel(), topk_ids.shape[1]
    swap = (not blockscale) and numel <= 64
    counts = torch.zeros(num_experts, dtype=torch.int64)
    for e in flat.tolist():
        if 0 <= e < num_experts:
            counts[e] += 1
    offs = torch.zeros(num_experts + 1, dtype=torch.int64)
    offs[1:] = torch.cumsum(counts, 0)
    ps1 = torch.zeros(num_experts, 3, dtype=torch.int32)
    ps2 = torch.zeros(num_experts, 3, dtype=torch.int32)
    for e in range(num_experts):
        m = int(counts[e])
        ps1[e] = torch.tensor([2 * n, m, k] if swap else [m, 2 * n, k], dtype=torch.int32)
        ps2[e] = torch.tensor([k, m, n] if swap else [m, k, n], dtype=torch.int32)
    total = int(offs[-1])
    a_map = torch.zeros(numel, dtype=torch.int32)
    c_map = torch.zeros(numel, dtype=torch.int32)
    nxt = offs[:-1].clone()
    for i, e in enumerate(flat.tolist()):
        if 0 <= e < num_experts:
            a_map[int(nxt[e])] = i // topk
            c_map[i] = int(nxt[e])
            nxt[e] += 1
        elif e == -1:
            c_map[i] = total
    out = (offs.to(torch.int32), ps1, ps2, a_map, c_map)
    if blockscale:
        bo = torch.zeros(num_experts + 1, dtype=torch.int64)
        bo[1:] = torch.cumsum((counts + 127) // 128 * 128, 0)
        out = out + (bo.to(torch.int32),)
    return out


def cutlass_moe_mm(a_q, b_q, a_scales, b_scales, expert_offsets, rows_per_expert, out_dtype=torch.bfloat16,
                   out=None):
    """cutlass_moe_mm (scaled_mm_entry.cu:242-270; ScaledEpilogueArray: D = a_scale * (b_scale * acc), the scale
    granularity inferred from the element counts, moe/get_group_starts.cuh:24-31,58-59).  a_q fp8 [rows, K];
    b_q fp8 [E, N, K]; rows of expert e = [expert_offsets[e], +rows_per_expert[e]).  Rows owned by no expert keep
    the contents of `out` (zeros when not given)."""
    E, N, K = b_q.shape
    c = torch.zeros(a_q.shape[0], N, dtype=out_dtype) if out is None else out.clone()
    per_tok = a_scales.numel() != 1
    per_ch = b_scales.numel() != E
    for e in range(E):
        m, o = int(rows_per_expert[e]), int(expert_offsets[e])
        if m == 0:
            continue
        acc = (a_q[o:o + m].to(torch.float64) @ b_q[e].to(torch.float64).t()).to(_f32)
        sa = a_scales.flatten()[o:o + m].to(_f32)[:, None] if per_tok else a_scales.flatten()[0].to(_f32)
        sb = b_scales.reshape(E, -1)[e].to(_f32).reshape(1, -1) if per_ch else b_scales.flatten()[e].to(_f32)
        c[o:o + m] = (sa * (sb * acc)).to(out_dtype)
    return c


def cutlass_moe_fp8(a, w1_q, w2_q, topk_weights, topk_ids, w1_scale, w2_scale, per_act_token, a1_scale=None,
                    a2_scale=None, expert_map=None):
    """cutlass_moe_fp8 (fused_moe/cutlass_moe.py:324-409 = MoEPrepareAndFinalizeNoEP + run_cutlass_moe_fp8 :25-215 +
    TopKWeightAndReduceContiguous): quantise, permute, GEMM1, silu_and_mul, quantise, GEMM2, un-permute, multiply by
    the routing weights (fp32 product rounded once), moe_sum."""
    M, K = a.shape
    E, _, N = w2_q.shape
    topk = topk_ids.shape[1]
    num_experts = expert_map.numel() if expert_map is not None else E
    local = topk_ids.long()
    if expert_map is not None:
        local = torch.where(expert_map[local] != -1, expert_map[local], -1)
    a1q, a1s = _quant_act(a, a1_scale, per_act_token)
    offs, ps1, _, a_map, c_map = get_cutlass_moe_mm_data(local, num_experts, N, K)[:5]
    counts = (offs[1:] - offs[:-1])[:E]
    ap = a1q[a_map.long()]
    asp = a1s.reshape(-1)[a_map.long()] if per_act_token else a1s
    c1 = cutlass_moe_mm(ap, w1_q, asp, w1_scale, offs[:E], counts, a.dtype)
    c2 = silu_and_mul(c1)
    a2q, a2s = _quant_act(c2, a2_scale, per_act_token)
    c3 = cutlass_moe_mm(a2q, w2_q, a2s, w2_scale, offs[:E], counts, a.dtype)
    out = c3[c_map.long()].view(M, topk, K)
    out = (out.to(_f32) * topk_weights.view(M, topk, 1).to(_f32)).to(a.dtype)   # topk_weight_and_reduce.py:89
    return moe_sum(out)
