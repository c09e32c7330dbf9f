"""Tensor-parallel plumbing of the decode path: the row-parallel all-reduce call site
(/root/reference/vllm/distributed/communication_op.py:12-14, called from
/root/reference/vllm/model_executor/layers/linear.py:1389 and fused_moe/layer.py:1526-1530)."""
from .communication_op import (get_peer_allreduce, get_tensor_model_parallel_rank,
                               get_tensor_model_parallel_world_size, init_peer_allreduce, init_tensor_parallel,
                               tensor_model_parallel_all_reduce)

__all__ = ["init_tensor_parallel", "tensor_model_parallel_all_reduce", "get_tensor_model_parallel_rank",
           "get_tensor_model_parallel_world_size", "init_peer_allreduce", "get_peer_allreduce"]
