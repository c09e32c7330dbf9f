"""torch.ops registration of the hot-path ops under the reference's schemas
(/root/reference/csrc/torch_bindings.cpp, /root/reference/csrc/moe/torch_bindings.cpp): loads
lib/libvllm_b200_torch.so (vllm_b200/torch_shim.cpp over the C ABI).  After `load_torch_ops()`:
  torch.ops._b200_C.{paged_attention_v1, paged_attention_v2, cutlass_scaled_mm, cutlass_scaled_mm_supports_fp8,
                     static_scaled_fp8_quant, dynamic_scaled_fp8_quant, dynamic_per_token_scaled_fp8_quant,
                     silu_and_mul, rms_norm, fused_add_rms_norm, rms_norm_dynamic_per_token_quant, rotary_embedding}
  torch.ops._b200_C_cache_ops.{reshape_and_cache, reshape_and_cache_flash, convert_fp8}
  torch.ops._b200_moe_C.{topk_softmax, moe_sum, moe_align_block_size}
A drop-in build registers the same ops as _C / _C_cache_ops / _moe_C (build.build_torch_shim(names=...))."""
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
SHIM_PATH = os.environ.get("VLLM_B200_TORCH_LIB", os.path.join(_HERE, "lib", "libvllm_b200_torch.so"))
_loaded = False


def load_torch_ops() -> None:
    global _loaded
    if _loaded:
        return
    if not os.path.exists(SHIM_PATH):
        raise RuntimeError(f"vllm_b200: {SHIM_PATH} not found; build it with `python vllm_b200/build.py --shim`")
    torch.ops.load_library(SHIM_PATH)
    _loaded = True
