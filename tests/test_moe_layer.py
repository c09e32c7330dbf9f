"""CPU checks of the MoE layer classes (host logic only; the kernels run in tests/test_gpu_fused_moe.py)."""
import inspect

import torch


def test_determine_expert_map_matches_reference_dealing():
    """layer.py:568-608: contiguous ranges, the first `remainder` ranks take one expert more."""
    from vllm_b200.fused_moe import determine_expert_map
    assert determine_expert_map(1, 0, 8) == (8, None)
    seen = torch.zeros(10, dtype=torch.int32)
    sizes = []
    for r in range(4):
        n, emap = determine_expert_map(4, r, 10)
        sizes.append(n)
        own = emap >= 0
        assert int(own.sum()) == n
        assert torch.equal(emap[own], torch.arange(n, dtype=torch.int32))
        seen += own.int()
    assert sizes == [3, 3, 2, 2] and bool((seen == 1).all())


def test_fused_moe_layer_shapes_and_interface():
    from vllm_b200.distributed import communication_op as co
    from vllm_b200.fused_moe import Fp8MoEMethod, FusedMoE, UnquantizedFusedMoEMethod
    saved = (co._TP_WORLD, co._TP_RANK)
    try:
        co._TP_WORLD, co._TP_RANK = 4, 1
        tp = FusedMoE(8, 2, 64, 256, params_dtype=torch.bfloat16, quant_method=Fp8MoEMethod("static"))
        assert tp.w13_weight.shape == (8, 128, 64) and tp.w13_weight.dtype == torch.float8_e4m3fn
        assert tp.w2_weight.shape == (8, 64, 64) and tp.w13_input_scale.shape == (8,)
        tp.quant_method.process_weights_after_loading(tp)
        assert tp.w13_input_scale.shape == (1,)
        ep = FusedMoE(8, 2, 64, 256, params_dtype=torch.bfloat16, ep_size=4)
        assert isinstance(ep.quant_method, UnquantizedFusedMoEMethod)
        assert ep.w13_weight.shape == (2, 512, 64) and ep.use_ep and ep.expert_map.tolist() == [-1, -1, 0, 1, -1, -1, -1, -1]
    finally:
        co._TP_WORLD, co._TP_RANK = saved
    # argument names of the reference's call sites (layer.py:1504-1524, fp8.py:918-940)
    ap = list(inspect.signature(Fp8MoEMethod.apply).parameters)
    assert ap[:6] == ["self", "layer", "x", "router_logits", "top_k", "renormalize"]
    for name in ("use_grouped_topk", "global_num_experts", "expert_map", "custom_routing_function", "scoring_func",
                 "e_score_correction_bias", "apply_router_weight_on_input", "activation"):
        assert name in ap
    init = list(inspect.signature(FusedMoE.__init__).parameters)
    assert init[1:7] == ["num_experts", "top_k", "hidden_size", "intermediate_size", "params_dtype", "reduce_results"]
