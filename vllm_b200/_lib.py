"""ctypes binding of libvllm_b200.so (C ABI declared in include/vllm_b200.h).

The product path has NO fallback: if the shared library is missing or a symbol is absent this
module raises at import / first use.
"""
import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_size_t, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VLLM_B200_LIB", os.path.join(_HERE, "lib", "libvllm_b200.so"))

# dtype enums (include/vllm_b200.h)
F32, F16, BF16, FP8_E4M3, I32, I64, U32 = range(7)
KV_AUTO, KV_FP8_E4M3 = 0, 1

_DTYPES = {
    torch.float32: F32, torch.float16: F16, torch.bfloat16: BF16, torch.float8_e4m3fn: FP8_E4M3,
    torch.uint8: FP8_E4M3,  # fp8 caches are carried as uint8 tensors (vllm/utils: STR_DTYPE_TO_TORCH_DTYPE)
    torch.int32: I32, torch.int64: I64, torch.uint32: U32,
}


def dtype_enum(dt: torch.dtype) -> int:
    try:
        return _DTYPES[dt]
    except KeyError:
        raise RuntimeError(f"vllm_b200: unsupported dtype {dt}") from None


def kv_dtype_enum(kv_cache_dtype: str) -> int:
    """Reference strings: csrc/quantization/fp8/nvidia/quant_utils.cuh:532-569."""
    if kv_cache_dtype == "auto":
        return KV_AUTO
    if kv_cache_dtype in ("fp8", "fp8_e4m3"):
        return KV_FP8_E4M3
    raise RuntimeError(f"Unsupported data type of kv cache: {kv_cache_dtype}")


# name -> (restype, argtypes); must list EVERY symbol include/vllm_b200.h declares
_P = c_void_p
SIGNATURES = {
    "vb200_last_error": (c_char_p, []),
    "vb200_version": (c_char_p, []),
    "vb200_reshape_and_cache": (c_int, [_P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int64,
                                        c_int64, c_int, c_int, _P, _P, _P]),
    "vb200_reshape_and_cache_flash": (c_int, [_P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int64,
                                              c_int64, c_int64, c_int64, c_int64, c_int, c_int, _P, _P, _P]),
    "vb200_rotary_embedding_reshape_and_cache": (c_int, [_P, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int,
                                                         c_int, c_int, c_int, c_int, c_int64, c_int64, c_int64, c_int,
                                                         c_int, _P, _P, _P]),
    "vb200_convert_fp8": (c_int, [_P, _P, c_int64, c_int, c_int, c_float, _P]),
    "vb200_paged_attention_v2": (c_int, [_P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_float,
                                         _P, _P, c_int, c_int, c_int, _P, c_int, c_int, _P, _P, c_int64,
                                         c_int64, c_int64, c_int, _P]),
    "vb200_paged_attention_v1_workspace_size": (c_size_t, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "vb200_paged_attention_v1_partition_size": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int, c_int]),
    "vb200_paged_attention_v1": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, c_int, c_float, _P, _P,
                                         c_int, c_int, c_int, _P, c_int, c_int, _P, _P, c_int64, c_int64,
                                         c_int64, c_int, _P, c_size_t, _P]),
    "vb200_paged_attention_v1_static_max_bytes": (c_size_t, []),
    "vb200_paged_attention_v1_static": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, c_int, c_float, _P, _P,
                                                c_int, c_int, c_int, _P, c_int, c_int, _P, _P, c_int64, c_int64,
                                                c_int64, c_int, _P]),
    "vb200_paged_attention_v1_quant_workspace_size": (c_size_t, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "vb200_paged_attention_v1_quant": (c_int, [_P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_float, _P, _P, c_int,
                                               c_int, c_int, _P, c_int, c_int, _P, _P, c_int64, c_int64, c_int64, c_int,
                                               _P, c_size_t, _P]),
    "vb200_paged_attention_flash_workspace_size": (c_size_t, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "vb200_paged_attention_flash": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, c_int, c_float, _P, _P, c_int, c_int,
                                            c_int, c_int, _P, c_int, c_int, _P, _P, c_int64, c_int64, c_int64, c_int64,
                                            c_int, _P, c_size_t, _P]),
    "vb200_paged_attention_flash_window": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, c_int, c_float, _P, _P, c_int,
                                                   c_int, c_int, c_int, _P, c_int, c_int, _P, _P, c_int64, c_int64,
                                                   c_int64, c_int64, c_int, _P, c_size_t, c_int, c_float, _P]),
    "vb200_static_scaled_fp8_quant": (c_int, [_P, _P, _P, c_int64, c_int, c_int, _P]),
    "vb200_dynamic_scaled_fp8_quant": (c_int, [_P, _P, _P, c_int64, c_int, c_int, _P]),
    "vb200_dynamic_per_token_scaled_fp8_quant": (c_int, [_P, _P, _P, _P, c_int64, c_int, c_int, _P]),
    "vb200_silu_and_mul": (c_int, [_P, _P, c_int64, c_int, c_int, _P]),
    "vb200_absmax_scale": (c_int, [_P, _P, c_int64, c_int, _P]),
    "vb200_silu_and_mul_absmax": (c_int, [_P, _P, _P, c_int64, c_int, c_int, _P]),
    "vb200_rms_norm": (c_int, [_P, _P, _P, c_float, c_int, c_int, c_int64, c_int, _P]),
    "vb200_fused_add_rms_norm": (c_int, [_P, _P, _P, c_float, c_int, c_int, c_int64, c_int, _P]),
    "vb200_rms_norm_dynamic_per_token_quant": (c_int, [_P, _P, _P, _P, c_float, _P, _P, c_int, c_int, c_int, _P]),
    "vb200_rotary_embedding": (c_int, [_P, _P, _P, c_int, c_int, c_int, c_int, c_int, _P, c_int, c_int64,
                                       c_int64, c_int64, c_int, _P]),
    "vb200_silu_and_mul_dynamic_per_token_quant": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, _P]),
    "vb200_cutlass_scaled_mm_workspace_size": (c_size_t, [c_int, c_int, c_int]),
    "vb200_cutlass_scaled_mm": (c_int, [_P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int64, c_int64,
                                        c_int64, c_int, c_int, c_int, _P, c_size_t, _P]),
    "vb200_cutlass_scaled_mm_supports_fp8": (c_int, [c_int]),
    "vb200_topk_softmax": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, c_int, _P]),
    "vb200_fused_topk": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, _P]),
    "vb200_moe_align_workspace_size": (c_size_t, [c_int, c_int]),
    "vb200_moe_align_block_size": (c_int, [_P, c_int, c_int, c_int, _P, c_int, _P, c_int, _P, _P,
                                           c_size_t, _P]),
    "vb200_moe_sum": (c_int, [_P, _P, c_int, c_int, c_int, c_int, _P]),
    "vb200_moe_grouped_gemm_fp8": (c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int,
                                           c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, _P, c_size_t,
                                           _P]),
    "vb200_moe_router_topk": (c_int, [_P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, _P]),
    "vb200_moe_grouped_gemm_fp8_q16": (c_int, [_P, _P, c_int, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int,
                                               c_int, c_int, c_int, c_int, c_int, c_int, c_int, _P, c_size_t, _P]),
    "vb200_moe_grouped_gemm_16": (c_int, [_P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                          c_int, c_int, c_int, _P, c_size_t, _P]),
    "vb200_moe_grouped_gemm_workspace_size": (c_size_t, [c_int, c_int]),
    "vb200_get_cutlass_moe_mm_data": (c_int, [_P, c_int, c_int, _P, _P, _P, _P, _P, c_int, c_int, c_int, _P, _P]),
    "vb200_get_cutlass_pplx_moe_mm_data": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, c_int, _P]),
    "vb200_shuffle_rows": (c_int, [_P, _P, _P, c_int, c_int, c_int64, _P]),
    "vb200_moe_gather_quant_fp8": (c_int, [_P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, _P]),
    "vb200_moe_unpermute_reduce": (c_int, [_P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, _P]),
    "vb200_flash_attn_varlen_paged": (c_int, [_P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                              c_int, c_float, c_int, c_int, c_int, _P, _P, c_int64, c_int64, c_int64, c_int64,
                                              c_int64, _P]),
    "vb200_flash_attn_varlen_paged_opts": (c_int, [_P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int,
                                                   c_int, c_int, c_float, c_int, c_int, c_int, _P, _P, c_int64, c_int64,
                                                   c_int64, c_int64, c_int64, c_int, c_float, _P, c_int64, _P, _P]),
    "vb200_greedy_argmax": (c_int, [_P, _P, _P, c_int, c_int, c_int64, c_int, _P]),
    "vb200_advance_step_flashattn": (c_int, [c_int, c_int, c_int, _P, _P, _P, _P, _P, _P, c_int64, _P]),
    "vb200_merge_attn_states": (c_int, [_P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, _P]),
    "vb200_peer_allreduce_signal_bytes": (c_size_t, []),
    "vb200_peer_buffer_alloc": (c_int, [c_size_t, _P, _P]),
    "vb200_peer_buffer_open": (c_int, [_P, _P]),
    "vb200_peer_buffer_close": (c_int, [_P]),
    "vb200_peer_buffer_free": (c_int, [_P]),
    "vb200_peer_allreduce": (c_int, [c_int, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, _P, _P, c_float, _P, _P,
                                     _P]),
    "vb200_peer_allreduce_2shot": (c_int, [c_int, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, _P, _P, c_float, _P,
                                           _P, _P]),
    "vb200_peer_allreduce_split": (c_int, [c_int, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, _P, _P, c_float, _P,
                                           _P, _P]),
    "vb200_cutlass_moe_mm": (c_int, [_P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int,
                                     c_int, _P]),
}

_lib = None


def load() -> ctypes.CDLL:
    """Load libvllm_b200.so and bind every declared symbol; raises if anything is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"vllm_b200: native library not found at {LIB_PATH}. Build it with "
            "`python vllm_b200/build.py` (or __graft_entry__.build()). There is no fallback path.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name, None)
        if fn is None:
            raise RuntimeError(f"vllm_b200: symbol {name} missing from {LIB_PATH}")
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    """Non-zero return -> RuntimeError, like the reference's TORCH_CHECK -> c10::Error."""
    if rc != 0:
        raise RuntimeError(load().vb200_last_error().decode())


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def require_cuda(*tensors) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("vllm_b200 ops need CUDA tensors (there is no CPU path in the product)")
