"""Unquantised (f16 / bf16) dense linear on the tcgen05 grouped GEMM.

Stands where `UnquantizedLinearMethod.apply` stands in the reference
(/root/reference/vllm/model_executor/layers/linear.py:221-230: `F.linear(x, layer.weight, bias)` -> cuBLAS): the
projections of an unquantised model (BASELINE configs[1]) and the lm_head of every model (never quantised).  A dense
`y = x W^T` is the grouped GEMM of `fused_experts` with ONE expert that owns every row, so it runs on the same
kernel (`vb200_moe_grouped_gemm_16`: weight rows on the UMMA M side, tokens on the N side, weights streamed once through
TMA, fp32 accumulation in TMEM) with a trivial routing table that is built once per (token count, device).

Measured (profiles/r02_dense_linear_s76.txt): correct (tests/test_gpu_dense_linear.py) but slower than the library GEMM
at decode batches -- the kernel has no split-K and a dense projection has E times fewer weight-row tiles than a MoE
layer (qkv of Llama-8B: 48 CTAs on 148 SMs): config 2 (bf16, batch 64) 8.66 vs 6.22 ms/step, lm_head at batch 256 ~350
vs 267 us.  The decode step therefore keeps the library GEMM for unquantised projections unless VB200_DENSE16=1.
"""
from typing import Optional

import torch

_ROUTING: dict = {}


def _routing(num_tokens: int, device: torch.device):
    """moe_align_block_size's outputs for one expert owning rows 0 .. num_tokens - 1 (cached: graph-capture safe once a
    warm-up call has created them)."""
    key = (num_tokens, device)
    r = _ROUTING.get(key)
    if r is None:
        block_m = min(256, max(16, (num_tokens + 15) // 16 * 16))
        m_blocks = (num_tokens + block_m - 1) // block_m
        ids = torch.arange(m_blocks * block_m, dtype=torch.int32, device=device)
        ids.clamp_(max=num_tokens)          # ids >= num_tokens are padding
        r = (block_m, ids, torch.zeros(m_blocks, dtype=torch.int32, device=device),
             torch.full((1,), m_blocks * block_m, dtype=torch.int32, device=device))
        _ROUTING[key] = r
    return r


def linear_16(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None,
              out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y = x @ weight.T (+ bias).  x [T, K] contiguous, weight [N, K] contiguous, both f16 or both bf16, K % 8 == 0;
    fp32 accumulation, one rounding to the activation dtype."""
    from ..fused_moe.fused_moe import invoke_fused_moe_kernel
    if x.dim() != 2 or weight.dim() != 2 or x.shape[1] != weight.shape[1]:
        raise RuntimeError(f"linear_16: x [T, K] and weight [N, K] expected, got {tuple(x.shape)} and {tuple(weight.shape)}")
    if x.dtype not in (torch.float16, torch.bfloat16) or weight.dtype != x.dtype:
        raise RuntimeError("linear_16: f16 / bf16 activations and weights of one dtype")
    if not x.is_cuda:
        raise RuntimeError("vllm_b200 ops need CUDA tensors (there is no CPU path in the product)")
    T, N = x.shape[0], weight.shape[0]
    if out is None:
        out = torch.empty((T, N), dtype=x.dtype, device=x.device)
    if T == 0:
        return out
    block_m, ids, expert_ids, num_post = _routing(T, x.device)
    invoke_fused_moe_kernel(x if x.is_contiguous() else x.contiguous(), weight.unsqueeze(0), out, None, None, None, ids,
                            expert_ids, num_post, False, 1, block_m, use_fp8_w8a8=False)
    if bias is not None:
        out.add_(bias)
    return out
