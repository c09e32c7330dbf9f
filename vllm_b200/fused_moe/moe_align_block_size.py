"""Mirror of /root/reference/vllm/model_executor/layers/fused_moe/moe_align_block_size.py:12-87."""
from typing import Optional

import torch

from .. import _custom_ops as ops


def moe_align_block_size(topk_ids: torch.Tensor, block_size: int, num_experts: int,
                         expert_map: Optional[torch.Tensor] = None, pad_sorted_ids: bool = False
                         ) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """Returns (sorted_token_ids, expert_ids, num_tokens_post_padded); see the reference docstring (:19-66).
    Order inside an expert is the stable order (ascending flat index), i.e. deterministic."""
    max_num_tokens_padded = topk_ids.numel() + num_experts * (block_size - 1)
    if pad_sorted_ids:
        max_num_tokens_padded = (max_num_tokens_padded + block_size - 1) // block_size * block_size
    sorted_ids = torch.empty((max_num_tokens_padded,), dtype=torch.int32, device=topk_ids.device)
    max_num_m_blocks = (max_num_tokens_padded + block_size - 1) // block_size
    expert_ids = torch.empty((max_num_m_blocks,), dtype=torch.int32, device=topk_ids.device)
    num_tokens_post_pad = torch.empty((1,), dtype=torch.int32, device=topk_ids.device)
    ops.moe_align_block_size(topk_ids, num_experts, block_size, sorted_ids, expert_ids, num_tokens_post_pad)
    if expert_map is not None:
        expert_ids = expert_map[expert_ids]
    return sorted_ids, expert_ids, num_tokens_post_pad
