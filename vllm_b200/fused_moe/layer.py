"""The MoE layer classes a model instantiates, on this package's kernels:

  FusedMoE                      /root/reference/vllm/model_executor/layers/fused_moe/layer.py:611-900 (__init__),
                                :1355-1440 (select_experts), :1457-1533 (forward / forward_impl, all-reduce :1526-1530)
  UnquantizedFusedMoEMethod     layer.py:255-465 (create_weights :270-300, apply -> forward_cuda :402-465)
  Fp8MoEMethod                  /root/reference/vllm/model_executor/layers/quantization/fp8.py:474-1010
                                (create_weights :520-640, apply :918-1010)

Tensor parallelism splits every expert along the intermediate size (w13 rows of gate and of up, w2 columns), the
reference's default without --enable-expert-parallel; expert parallelism is expressed through `expert_map`
(determine_expert_map, layer.py:568-608) exactly as fused_experts consumes it.  What is not carried over: EPLB,
grouped top-k / custom routing functions, DP dispatch-combine, the modular-kernel plumbing -- control plane around the
same kernels.  Weights are created empty (checkpoint loading is out of scope); `process_weights_after_loading`
stands where the reference re-quantises / transposes, and here only normalises scale shapes.
"""
from typing import Callable, Optional

import torch
from torch import nn

from ..distributed.communication_op import (get_tensor_model_parallel_rank, get_tensor_model_parallel_world_size,
                                            tensor_model_parallel_all_reduce)
from .fused_moe import fused_experts, fused_topk

FP8 = torch.float8_e4m3fn


def determine_expert_map(ep_size: int, ep_rank: int, global_num_experts: int):
    """layer.py:568-608: experts are dealt out in contiguous ranges, the first `remainder` ranks take one more.
    Returns (local_num_experts, expert_map or None); expert_map[g] = local index or -1."""
    if ep_size == 1:
        return global_num_experts, None
    base, rem = divmod(global_num_experts, ep_size)
    local = base + (1 if ep_rank < rem else 0)
    start = ep_rank * base + min(ep_rank, rem)
    emap = torch.full((global_num_experts,), -1, dtype=torch.int32)
    emap[start:start + local] = torch.arange(local, dtype=torch.int32)
    return local, emap


class UnquantizedFusedMoEMethod:
    """f16 / bf16 experts (layer.py:255-465)."""

    def create_weights(self, layer: nn.Module, num_experts: int, hidden_size: int,
                       intermediate_size_per_partition: int, params_dtype: torch.dtype, **_) -> None:
        layer.w13_weight = nn.Parameter(torch.empty(num_experts, 2 * intermediate_size_per_partition, hidden_size,
                                                    dtype=params_dtype), requires_grad=False)
        layer.w2_weight = nn.Parameter(torch.empty(num_experts, hidden_size, intermediate_size_per_partition,
                                                   dtype=params_dtype), requires_grad=False)

    def process_weights_after_loading(self, layer: nn.Module) -> None:
        pass

    def apply(self, layer: nn.Module, x: torch.Tensor, router_logits: torch.Tensor, top_k: int, renormalize: bool,
              use_grouped_topk: bool = False, topk_group: Optional[int] = None, num_expert_group: Optional[int] = None,
              global_num_experts: int = -1, expert_map: Optional[torch.Tensor] = None,
              custom_routing_function: Optional[Callable] = None, scoring_func: str = "softmax",
              e_score_correction_bias: Optional[torch.Tensor] = None, apply_router_weight_on_input: bool = False,
              activation: str = "silu", **_) -> torch.Tensor:
        topk_weights, topk_ids = FusedMoE.select_experts(
            hidden_states=x, router_logits=router_logits, use_grouped_topk=use_grouped_topk, top_k=top_k,
            renormalize=renormalize, topk_group=topk_group, num_expert_group=num_expert_group,
            custom_routing_function=custom_routing_function, scoring_func=scoring_func,
            e_score_correction_bias=e_score_correction_bias)
        return fused_experts(x, layer.w13_weight, layer.w2_weight, topk_weights, topk_ids, inplace=False,
                             activation=activation, apply_router_weight_on_input=apply_router_weight_on_input,
                             global_num_experts=global_num_experts, expert_map=expert_map)


class Fp8MoEMethod:
    """FP8 w8a8 experts with per-tensor (per-expert) weight scales and dynamic or static activation scales
    (fp8.py:474-1010; the block-quantised / Marlin / AITER / FlashInfer branches are other back ends)."""

    def __init__(self, activation_scheme: str = "dynamic", per_channel: bool = False) -> None:
        if activation_scheme not in ("dynamic", "static"):
            raise ValueError(f"Unsupported activation scheme {activation_scheme}")
        self.activation_scheme = activation_scheme
        self.per_channel = per_channel   # per-output-channel weight scales + per-token activation scales

    def create_weights(self, layer: nn.Module, num_experts: int, hidden_size: int,
                       intermediate_size_per_partition: int, params_dtype: torch.dtype, **_) -> None:
        inter = intermediate_size_per_partition
        layer.w13_weight = nn.Parameter(torch.empty(num_experts, 2 * inter, hidden_size, dtype=FP8), requires_grad=False)
        layer.w2_weight = nn.Parameter(torch.empty(num_experts, hidden_size, inter, dtype=FP8), requires_grad=False)
        sshape13 = (num_experts, 2 * inter, 1) if self.per_channel else (num_experts,)
        sshape2 = (num_experts, hidden_size, 1) if self.per_channel else (num_experts,)
        layer.w13_weight_scale = nn.Parameter(torch.ones(sshape13, dtype=torch.float32), requires_grad=False)
        layer.w2_weight_scale = nn.Parameter(torch.ones(sshape2, dtype=torch.float32), requires_grad=False)
        if self.activation_scheme == "static":
            layer.w13_input_scale = nn.Parameter(torch.ones(num_experts, dtype=torch.float32), requires_grad=False)
            layer.w2_input_scale = nn.Parameter(torch.ones(num_experts, dtype=torch.float32), requires_grad=False)
        else:
            layer.w13_input_scale = None
            layer.w2_input_scale = None

    def process_weights_after_loading(self, layer: nn.Module) -> None:
        """fp8.py:700-800: a static checkpoint carries one input scale per expert; the kernel takes ONE per tensor,
        the maximum (fp8.py:760-772)."""
        if self.activation_scheme == "static":
            layer.w13_input_scale = nn.Parameter(layer.w13_input_scale.max().reshape(1), requires_grad=False)
            layer.w2_input_scale = nn.Parameter(layer.w2_input_scale.max().reshape(1), requires_grad=False)

    def apply(self, layer: nn.Module, x: torch.Tensor, router_logits: torch.Tensor, top_k: int, renormalize: bool,
              use_grouped_topk: bool = False, topk_group: Optional[int] = None, num_expert_group: Optional[int] = None,
              global_num_experts: int = -1, expert_map: Optional[torch.Tensor] = None,
              custom_routing_function: Optional[Callable] = None, scoring_func: str = "softmax",
              e_score_correction_bias: Optional[torch.Tensor] = None, apply_router_weight_on_input: bool = False,
              activation: str = "silu", **_) -> torch.Tensor:
        topk_weights, topk_ids = FusedMoE.select_experts(
            hidden_states=x, router_logits=router_logits, use_grouped_topk=use_grouped_topk, top_k=top_k,
            renormalize=renormalize, topk_group=topk_group, num_expert_group=num_expert_group,
            custom_routing_function=custom_routing_function, scoring_func=scoring_func,
            e_score_correction_bias=e_score_correction_bias)
        return fused_experts(x, layer.w13_weight, layer.w2_weight, topk_weights, topk_ids, inplace=False,
                             activation=activation, apply_router_weight_on_input=apply_router_weight_on_input,
                             use_fp8_w8a8=True, per_channel_quant=self.per_channel,
                             global_num_experts=global_num_experts, expert_map=expert_map,
                             w1_scale=layer.w13_weight_scale, w2_scale=layer.w2_weight_scale,
                             a1_scale=layer.w13_input_scale, a2_scale=layer.w2_input_scale)


class FusedMoE(nn.Module):
    """layer.py:611-1533.  `quant_method`: an Fp8MoEMethod / UnquantizedFusedMoEMethod instance (the reference
    derives it from `quant_config`; passing None gives unquantised experts)."""

    def __init__(self, num_experts: int, top_k: int, hidden_size: int, intermediate_size: int,
                 params_dtype: Optional[torch.dtype] = None, reduce_results: bool = False, renormalize: bool = True,
                 use_grouped_topk: bool = False, num_expert_group: Optional[int] = None,
                 topk_group: Optional[int] = None, quant_method=None, tp_size: Optional[int] = None,
                 ep_size: Optional[int] = None, prefix: str = "", custom_routing_function: Optional[Callable] = None,
                 scoring_func: str = "softmax", e_score_correction_bias: Optional[torch.Tensor] = None,
                 apply_router_weight_on_input: bool = False, activation: str = "silu") -> None:
        super().__init__()
        self.params_dtype = params_dtype if params_dtype is not None else torch.get_default_dtype()
        world = get_tensor_model_parallel_world_size()
        rank = get_tensor_model_parallel_rank()
        self.ep_size = ep_size if ep_size is not None else 1
        if self.ep_size > 1:
            # expert parallel: whole experts per rank (the TP group is the EP group, tensor parallel degree 1 inside)
            if self.ep_size != world:
                raise ValueError("FusedMoE: ep_size must equal the tensor-parallel world size")
            self.tp_size, self.tp_rank, self.ep_rank = 1, 0, rank
        else:
            self.tp_size = tp_size if tp_size is not None else world
            self.tp_rank, self.ep_rank = (rank if self.tp_size > 1 else 0), 0
        self.global_num_experts = num_experts
        self.local_num_experts, self.expert_map = determine_expert_map(self.ep_size, self.ep_rank, num_experts)
        self.top_k = top_k
        if intermediate_size % self.tp_size != 0:
            raise ValueError("FusedMoE: intermediate_size must be divisible by the tensor-parallel size")
        self.hidden_size = hidden_size
        self.intermediate_size_per_partition = intermediate_size // self.tp_size
        self.reduce_results = reduce_results
        self.renormalize = renormalize
        self.use_grouped_topk = use_grouped_topk
        self.num_expert_group, self.topk_group = num_expert_group, topk_group
        self.custom_routing_function = custom_routing_function
        self.scoring_func = scoring_func
        self.e_score_correction_bias = e_score_correction_bias
        self.apply_router_weight_on_input = apply_router_weight_on_input
        self.activation = activation
        self.layer_name = prefix
        if scoring_func != "softmax" and not use_grouped_topk:
            raise ValueError("Only softmax scoring function is supported for non-grouped topk.")
        self.quant_method = quant_method if quant_method is not None else UnquantizedFusedMoEMethod()
        self.quant_method.create_weights(layer=self, num_experts=self.local_num_experts, hidden_size=hidden_size,
                                         intermediate_size_per_partition=self.intermediate_size_per_partition,
                                         params_dtype=self.params_dtype)

    @property
    def use_ep(self) -> bool:
        return self.ep_size > 1

    @staticmethod
    def select_experts(hidden_states: torch.Tensor, router_logits: torch.Tensor, top_k: int, use_grouped_topk: bool,
                       renormalize: bool, topk_group: Optional[int] = None, num_expert_group: Optional[int] = None,
                       custom_routing_function: Optional[Callable] = None, scoring_func: str = "softmax",
                       e_score_correction_bias: Optional[torch.Tensor] = None,
                       indices_type: Optional[torch.dtype] = None, **_):
        """layer.py:1355-1440.  Returns (topk_weights fp32 [T, k], topk_ids [T, k])."""
        if use_grouped_topk:
            raise NotImplementedError("grouped top-k (DeepSeek routing) is not built")
        if custom_routing_function is not None:
            topk_weights, topk_ids = custom_routing_function(hidden_states=hidden_states, gating_output=router_logits,
                                                             topk=top_k, renormalize=renormalize)
            return topk_weights, topk_ids if indices_type is None else topk_ids.to(indices_type)
        topk_weights, topk_ids, _ = fused_topk(hidden_states, router_logits, top_k, renormalize, indices_type)
        return topk_weights, topk_ids

    def maybe_all_reduce_tensor_model_parallel(self, final_hidden_states: torch.Tensor) -> torch.Tensor:
        return tensor_model_parallel_all_reduce(final_hidden_states)

    def forward(self, hidden_states: torch.Tensor, router_logits: torch.Tensor) -> torch.Tensor:
        return self.forward_impl(hidden_states, router_logits)

    def forward_impl(self, hidden_states: torch.Tensor, router_logits: torch.Tensor) -> torch.Tensor:
        """layer.py:1482-1533 without the DP dispatch / combine branches."""
        emap = self.expert_map
        if emap is not None and emap.device != hidden_states.device:
            emap = self.expert_map = emap.to(hidden_states.device)
        out = self.quant_method.apply(
            layer=self, x=hidden_states, router_logits=router_logits, top_k=self.top_k, renormalize=self.renormalize,
            use_grouped_topk=self.use_grouped_topk, global_num_experts=self.global_num_experts, expert_map=emap,
            topk_group=self.topk_group, num_expert_group=self.num_expert_group,
            custom_routing_function=self.custom_routing_function, scoring_func=self.scoring_func,
            e_score_correction_bias=self.e_score_correction_bias, activation=self.activation,
            apply_router_weight_on_input=self.apply_router_weight_on_input)
        if self.reduce_results and (self.tp_size > 1 or self.ep_size > 1):
            out = self.maybe_all_reduce_tensor_model_parallel(out)
        return out
