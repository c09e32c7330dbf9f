"""fused-MoE layer glue of the B200 decode path (mirror of vllm/model_executor/layers/fused_moe)."""
from .fused_moe import (fused_experts, fused_moe, fused_topk, invoke_fused_moe_kernel, make_dummy_experts,
                        moe_forward, router_topk)
from .cutlass_moe import cutlass_moe_fp8, run_cutlass_moe_fp8
from .layer import Fp8MoEMethod, FusedMoE, UnquantizedFusedMoEMethod, determine_expert_map
from .moe_align_block_size import moe_align_block_size

__all__ = ["fused_topk", "fused_experts", "fused_moe", "invoke_fused_moe_kernel", "moe_align_block_size",
           "make_dummy_experts", "moe_forward", "router_topk", "cutlass_moe_fp8", "run_cutlass_moe_fp8", "FusedMoE", "Fp8MoEMethod",
           "UnquantizedFusedMoEMethod", "determine_expert_map"]
