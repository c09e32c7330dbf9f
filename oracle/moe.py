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
