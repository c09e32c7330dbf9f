"""CPU-side checks of the drop-in boundary: libvllm_b200.so loads and exports every symbol that
include/vllm_b200.h declares, and the Python host mirrors the reference wrapper names."""
import ctypes
import inspect
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    txt = open(os.path.join(ROOT, "include", "vllm_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(vb200_[a-z0-9_]+)\s*\(", txt)))


def test_header_declares_symbols():
    syms = _declared_symbols()
    assert "vb200_paged_attention_v1" in syms and "vb200_reshape_and_cache" in syms
    assert len(syms) >= 10


def test_library_exports_every_declared_symbol():
    from vllm_b200 import _lib
    assert os.path.exists(_lib.LIB_PATH), "build with python vllm_b200/build.py"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [s for s in _declared_symbols() if not hasattr(lib, s)]
    assert not missing, f"symbols declared in include/vllm_b200.h but not exported: {missing}"


def test_ctypes_table_matches_header():
    from vllm_b200 import _lib
    assert sorted(_lib.SIGNATURES) == _declared_symbols()
    lib = _lib.load()
    assert lib.vb200_version().decode().startswith("vllm_b200")


def test_missing_library_fails_loudly(monkeypatch):
    from vllm_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libvllm_b200.so")
    with pytest.raises(RuntimeError, match="no fallback"):
        _lib.load()


def test_custom_ops_mirror_reference_signatures():
    """Argument names/order of the reference wrappers (vllm/_custom_ops.py:37-97, 1617-1660)."""
    from vllm_b200 import _custom_ops as ops
    v1 = list(inspect.signature(ops.paged_attention_v1).parameters)
    assert v1[:19] == ["out", "query", "key_cache", "value_cache", "num_kv_heads", "scale", "block_tables",
                       "seq_lens", "block_size", "max_seq_len", "alibi_slopes", "kv_cache_dtype", "k_scale",
                       "v_scale", "tp_rank", "blocksparse_local_blocks", "blocksparse_vert_stride",
                       "blocksparse_block_size", "blocksparse_head_sliding_step"]
    v2 = list(inspect.signature(ops.paged_attention_v2).parameters)
    assert v2[:4] == ["out", "exp_sum", "max_logits", "tmp_out"] and v2[4:] == v1[1:19]
    rc = list(inspect.signature(ops.reshape_and_cache).parameters)
    assert rc == ["key", "value", "key_cache", "value_cache", "slot_mapping", "kv_cache_dtype", "k_scale",
                  "v_scale"]
    assert list(inspect.signature(ops.reshape_and_cache_flash).parameters) == rc


def test_cpu_tensors_are_rejected():
    import torch
    from vllm_b200 import _custom_ops as ops
    k = torch.zeros(2, 1, 16, dtype=torch.bfloat16)
    kc = torch.zeros(4, 1, 2, 16, 8, dtype=torch.bfloat16)
    vc = torch.zeros(4, 1, 16, 16, dtype=torch.bfloat16)
    one = torch.ones(1)
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.reshape_and_cache(k, k, kc, vc, torch.zeros(2, dtype=torch.int64), "auto", one, one)


def test_torch_shim_registers_reference_schemas():
    """The TORCH_LIBRARY shim loads without a GPU and registers the ops under the reference's schema strings
    (csrc/torch_bindings.cpp:45-69, 422-426, 676-691; csrc/moe/torch_bindings.cpp:6-22)."""
    import torch
    from vllm_b200 import torch_bindings
    if not os.path.exists(torch_bindings.SHIM_PATH):
        pytest.skip("libvllm_b200_torch.so not built")
    torch_bindings.load_torch_ops()
    s = str(torch.ops._b200_C.paged_attention_v1.default._schema)
    assert "!" in s.split("out")[0] and "str kv_cache_dtype" in s and "int blocksparse_head_sliding_step" in s
    s = str(torch.ops._b200_C.cutlass_scaled_mm.default._schema)
    assert "out, Tensor a, Tensor b, Tensor a_scales, Tensor b_scales, Tensor? bias) -> ()" in s
    s = str(torch.ops._b200_C_cache_ops.reshape_and_cache_flash.default._schema)
    assert s.count("!") == 2 and "key_cache" in s and "Tensor slot_mapping" in s
    s = str(torch.ops._b200_moe_C.moe_align_block_size.default._schema)
    assert "int num_experts, int block_size" in s
    assert torch.ops._b200_C.cutlass_scaled_mm_supports_fp8(100) and not torch.ops._b200_C.cutlass_scaled_mm_supports_fp8(89)
    # CUTLASS-style MoE API (csrc/torch_bindings.cpp:446-488, csrc/moe/torch_bindings.cpp:75-78)
    s = str(torch.ops._b200_C.cutlass_moe_mm.default._schema)
    assert "out_tensors, Tensor a_tensors, Tensor b_tensors, Tensor a_scales, Tensor b_scales, Tensor expert_offsets, " \
           "Tensor problem_sizes, Tensor a_strides, Tensor b_strides, Tensor c_strides, bool per_act_token, " \
           "bool per_out_ch) -> ()" in s
    s = str(torch.ops._b200_C.get_cutlass_moe_mm_data.default._schema)
    assert s.count("!") == 5 and "int num_experts, int n, int k, Tensor? blockscale_offsets) -> ()" in s
    s = str(torch.ops._b200_C.get_cutlass_pplx_moe_mm_data.default._schema)
    assert s.count("!") == 3 and "int num_local_experts, int padded_m, int n, int k) -> ()" in s
    s = str(torch.ops._b200_moe_C.shuffle_rows.default._schema)
    assert "Tensor input_tensor, Tensor dst2src_map, Tensor(" in s and s.count("!") == 1 and "output_tensor) -> ()" in s
    assert torch.ops._b200_C.cutlass_group_gemm_supported(100) and not torch.ops._b200_C.cutlass_group_gemm_supported(90)


def test_glue_mirrors_reference_names():
    """fused_moe / attention / quantization glue keeps the reference's callables and argument names."""
    from vllm_b200.attention import B200PagedAttentionBackend, PagedAttention
    from vllm_b200.fused_moe import fused_experts, fused_topk, moe_align_block_size
    from vllm_b200.quantization import Fp8LinearOp
    assert list(inspect.signature(fused_topk).parameters) == ["hidden_states", "gating_output", "topk", "renormalize",
                                                              "indices_type"]
    fe = list(inspect.signature(fused_experts).parameters)
    assert fe[:6] == ["hidden_states", "w1", "w2", "topk_weights", "topk_ids", "inplace"]
    for name in ("use_fp8_w8a8", "per_channel_quant", "global_num_experts", "expert_map", "w1_scale", "w2_scale",
                 "a1_scale", "a2_scale"):
        assert name in fe
    assert list(inspect.signature(moe_align_block_size).parameters) == ["topk_ids", "block_size", "num_experts",
                                                                        "expert_map", "pad_sorted_ids"]
    assert list(inspect.signature(Fp8LinearOp.apply).parameters)[1:] == ["input", "weight", "weight_scale", "out_dtype",
                                                                         "input_scale", "input_scale_ub", "bias"]
    assert PagedAttention.get_kv_cache_shape(4, 16, 8, 128) == (2, 4, 16 * 8 * 128)
    # CUTLASS-style MoE (vllm/_custom_ops.py:871-969, fused_moe/cutlass_moe.py:25-45, 324-339)
    from vllm_b200 import _custom_ops as ops
    from vllm_b200.fused_moe import cutlass_moe_fp8, run_cutlass_moe_fp8
    assert list(inspect.signature(ops.get_cutlass_moe_mm_data).parameters) == [
        "topk_ids", "expert_offsets", "problem_sizes1", "problem_sizes2", "input_permutation", "output_permutation",
        "num_experts", "n", "k", "blockscale_offsets"]
    assert list(inspect.signature(ops.cutlass_moe_mm).parameters) == [
        "out_tensors", "a_tensors", "b_tensors", "a_scales", "b_scales", "expert_offsets", "problem_sizes", "a_strides",
        "b_strides", "c_strides", "per_act_token", "per_out_ch"]
    assert list(inspect.signature(ops.shuffle_rows).parameters) == ["input_tensor", "dst2src_map"]
    assert list(inspect.signature(cutlass_moe_fp8).parameters) == [
        "a", "w1_q", "w2_q", "topk_weights", "topk_ids", "w1_scale", "w2_scale", "per_act_token", "activation",
        "a1_scale", "a2_scale", "expert_map", "apply_router_weight_on_input", "global_num_experts"]
    assert list(inspect.signature(run_cutlass_moe_fp8).parameters)[:8] == [
        "output", "hidden_states", "w1", "w2", "topk_ids", "activation_callable", "global_num_experts", "expert_map"]
    for m in ("get_name", "get_impl_cls", "get_metadata_cls", "get_builder_cls", "get_kv_cache_shape", "swap_blocks",
              "copy_blocks", "get_supported_head_sizes"):
        assert hasattr(B200PagedAttentionBackend, m)
    fwd = list(inspect.signature(B200PagedAttentionBackend.get_impl_cls().forward).parameters)
    assert fwd == ["self", "layer", "query", "key", "value", "kv_cache", "attn_metadata", "output", "output_scale"]
