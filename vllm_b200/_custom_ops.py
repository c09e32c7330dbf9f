"""Host-side mirror of /root/reference/vllm/_custom_ops.py for the decode hot path.

Same function names, argument order, defaults and error behaviour (RuntimeError) as the
reference wrappers; each one unwraps its tensors into the C ABI of libvllm_b200.so
(include/vllm_b200.h).  There is no Triton / CPU fallback: CUDA tensors on a B200 only.
"""
import contextlib
from typing import Optional

import torch

from . import _lib
from ._lib import check, dtype_enum, kv_dtype_enum, load, ptr, require_cuda, stream


@contextlib.contextmanager
def _device_of(t: torch.Tensor):
    """OptionalCUDAGuard(device_of(tensor)) of the reference launchers (paged_attention_v1.cu:88)."""
    if t.is_cuda and t.device.index != torch.cuda.current_device():
        with torch.cuda.device(t.device):
            yield
    else:
        yield


def _check_blocksparse(blocksparse_vert_stride: int) -> None:
    # block-sparse attention (attention_kernels.cuh:204-251) is outside the north_star scope
    if blocksparse_vert_stride is not None and blocksparse_vert_stride > 1:
        raise RuntimeError("vllm_b200: block-sparse paged attention is not supported")


# ------------------------------------------------------------------------------------------------
# paged attention  (reference: vllm/_custom_ops.py:37-97)
# ------------------------------------------------------------------------------------------------
_PA_WORKSPACE = {}


def _pa_common(query, key_cache, value_cache, block_tables, seq_lens, kv_cache_dtype):
    require_cuda(query, key_cache, value_cache, block_tables, seq_lens)
    if query.dim() != 3:
        raise RuntimeError("query must be [num_seqs, num_heads, head_size]")
    if query.stride(2) != 1 or query.stride(1) != query.shape[2]:
        raise RuntimeError("query heads must be contiguous (stride(1) == head_size)")
    if block_tables.dtype != torch.int32 or seq_lens.dtype != torch.int32:
        raise RuntimeError("block_tables and seq_lens must be int32")
    kvd = kv_dtype_enum(kv_cache_dtype)
    if kvd == _lib.KV_AUTO and key_cache.dtype != query.dtype:
        raise RuntimeError(f"kv_cache_dtype 'auto' needs cache dtype == query dtype "
                           f"({key_cache.dtype} vs {query.dtype})")
    if kvd == _lib.KV_FP8_E4M3 and key_cache.element_size() != 1:
        raise RuntimeError("fp8 kv cache must be a 1-byte tensor (uint8 / float8_e4m3fn)")
    esz = key_cache.element_size()
    if (query.data_ptr() % 8) or (query.stride(0) % 4) or (key_cache.data_ptr() % 16) or (
            value_cache.data_ptr() % 16) or (key_cache.stride(0) * esz % 16) or (
            key_cache.stride(1) * esz % 16):
        raise RuntimeError("paged attention: query must be 8-byte and caches 16-byte aligned")
    return kvd


def paged_attention_v1(
    out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    num_kv_heads: int,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    block_size: int,
    max_seq_len: int,
    alibi_slopes: Optional[torch.Tensor],
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
    tp_rank: int = 0,
    blocksparse_local_blocks: int = 0,
    blocksparse_vert_stride: int = 0,
    blocksparse_block_size: int = 64,
    blocksparse_head_sliding_step: int = 0,
    partition_size: int = 0,
) -> None:
    """torch.ops._C.paged_attention_v1 (csrc/torch_bindings.cpp:45-54).  The trailing
    `partition_size` (0 = library heuristic, <0 = never split) is a B200 tuning extension."""
    _check_blocksparse(blocksparse_vert_stride)
    kvd = _pa_common(query, key_cache, value_cache, block_tables, seq_lens, kv_cache_dtype)
    lib = load()
    num_seqs, num_heads, head_size = query.shape
    qd = dtype_enum(query.dtype)
    with _device_of(query):
        ws, ws_bytes = None, 0
        if partition_size >= 0:
            if partition_size == 0:
                partition_size = lib.vb200_paged_attention_v1_partition_size(
                    num_seqs, num_heads, num_kv_heads, max_seq_len, block_size)
            ws_bytes = lib.vb200_paged_attention_v1_workspace_size(
                num_seqs, num_heads, head_size, max_seq_len, qd, partition_size)
            if ws_bytes > 0:
                # scratch through the torch caching allocator (graph-capture safe), like the
                # reference ops that allocate workspaces (topk_softmax_kernels.cu:515)
                ws = torch.empty(ws_bytes, dtype=torch.uint8, device=query.device)
        check(lib.vb200_paged_attention_v1(
            ptr(out), ptr(query), ptr(key_cache), ptr(value_cache), num_seqs, num_heads, head_size,
            num_kv_heads, float(scale), ptr(block_tables), ptr(seq_lens), block_size, max_seq_len,
            block_tables.stride(0), ptr(alibi_slopes), qd, kvd, ptr(k_scale), ptr(v_scale),
            query.stride(0), key_cache.stride(0), key_cache.stride(1),
            max(partition_size, 0), ptr(ws), ws_bytes, stream()))


def paged_attention_v2(
    out: torch.Tensor,
    exp_sum: torch.Tensor,
    max_logits: torch.Tensor,
    tmp_out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    num_kv_heads: int,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    block_size: int,
    max_seq_len: int,
    alibi_slopes: Optional[torch.Tensor],
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
    tp_rank: int = 0,
    blocksparse_local_blocks: int = 0,
    blocksparse_vert_stride: int = 0,
    blocksparse_block_size: int = 64,
    blocksparse_head_sliding_step: int = 0,
) -> None:
    """torch.ops._C.paged_attention_v2 (csrc/torch_bindings.cpp:58-68); PARTITION_SIZE = 512 is
    implied by the caller-allocated partial buffers (vllm/attention/ops/paged_attn.py:26,120)."""
    _check_blocksparse(blocksparse_vert_stride)
    kvd = _pa_common(query, key_cache, value_cache, block_tables, seq_lens, kv_cache_dtype)
    lib = load()
    num_seqs, num_heads, head_size = query.shape
    max_parts = exp_sum.shape[2]
    partition_size = 512
    if max_parts * partition_size < max_seq_len:
        raise RuntimeError("paged_attention_v2: exp_sums has too few partitions for max_seq_len")
    if not (exp_sum.is_contiguous() and max_logits.is_contiguous() and tmp_out.is_contiguous()):
        raise RuntimeError("paged_attention_v2: partial buffers must be contiguous")
    # the kernel derives max_num_partitions from max_seq_len; the buffers must match it
    if (max_seq_len + partition_size - 1) // partition_size != max_parts:
        raise RuntimeError("paged_attention_v2: partial buffers must have ceil(max_seq_len/512) partitions")
    with _device_of(query):
        check(lib.vb200_paged_attention_v2(
            ptr(out), ptr(exp_sum), ptr(max_logits), ptr(tmp_out), ptr(query), ptr(key_cache),
            ptr(value_cache), num_seqs, num_heads, head_size, num_kv_heads, float(scale),
            ptr(block_tables), ptr(seq_lens), block_size, max_seq_len, block_tables.stride(0),
            ptr(alibi_slopes), dtype_enum(query.dtype), kvd, ptr(k_scale), ptr(v_scale),
            query.stride(0), key_cache.stride(0), key_cache.stride(1), partition_size, stream()))


# ------------------------------------------------------------------------------------------------
# cache ops  (reference: vllm/_custom_ops.py:1617-1660)
# ------------------------------------------------------------------------------------------------
def reshape_and_cache(
    key: torch.Tensor,
    value: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    """torch.ops._C_cache_ops.reshape_and_cache (csrc/torch_bindings.cpp:676-681)."""
    require_cuda(key, value, key_cache, value_cache, slot_mapping)
    if slot_mapping.dtype != torch.int64:
        raise RuntimeError("slot_mapping must be int64")
    kvd = kv_dtype_enum(kv_cache_dtype)
    num_tokens = slot_mapping.shape[0]
    num_heads, head_size = key.shape[1], key.shape[2]
    block_size, x = key_cache.shape[3], key_cache.shape[4]
    if not (key_cache.is_contiguous() and value_cache.is_contiguous()):
        raise RuntimeError("reshape_and_cache: caches must be contiguous")
    if key.stride(2) != 1 or key.stride(1) != head_size or value.stride(2) != 1 or value.stride(1) != head_size:
        raise RuntimeError("reshape_and_cache: key/value heads must be contiguous")
    with _device_of(key):
        check(load().vb200_reshape_and_cache(
            ptr(key), ptr(value), ptr(key_cache), ptr(value_cache), ptr(slot_mapping), num_tokens,
            num_heads, head_size, block_size, x, key.stride(0), value.stride(0),
            dtype_enum(key.dtype), kvd, ptr(k_scale), ptr(v_scale), stream()))


def reshape_and_cache_flash(
    key: torch.Tensor,
    value: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    """torch.ops._C_cache_ops.reshape_and_cache_flash (csrc/torch_bindings.cpp:685-691).  The
    cache is addressed through key_cache's strides, so NHD and HND-permuted views both work
    (csrc/cache_kernels.cu:438-440)."""
    require_cuda(key, value, key_cache, value_cache, slot_mapping)
    if slot_mapping.dtype != torch.int64:
        raise RuntimeError("slot_mapping must be int64")
    kvd = kv_dtype_enum(kv_cache_dtype)
    num_tokens = slot_mapping.shape[0]  # may be < key.shape[0] (cache_kernels.cu:421-431)
    num_heads, head_size = key.shape[1], key.shape[2]
    block_size = key_cache.shape[1]
    if key_cache.stride(3) != 1 or value_cache.stride() != key_cache.stride():
        raise RuntimeError("reshape_and_cache_flash: caches need unit head-dim stride and equal strides")
    if key.stride(2) != 1 or key.stride(1) != head_size or value.stride(2) != 1 or value.stride(1) != head_size:
        raise RuntimeError("reshape_and_cache_flash: key/value heads must be contiguous")
    with _device_of(key):
        check(load().vb200_reshape_and_cache_flash(
            ptr(key), ptr(value), ptr(key_cache), ptr(value_cache), ptr(slot_mapping), num_tokens,
            num_heads, head_size, block_size, key_cache.stride(0), key_cache.stride(1),
            key_cache.stride(2), key.stride(0), value.stride(0), dtype_enum(key.dtype), kvd,
            ptr(k_scale), ptr(v_scale), stream()))


def convert_fp8(output: torch.Tensor, input: torch.Tensor, scale: float = 1.0,
                kv_dtype: str = "fp8") -> None:
    """torch.ops._C_cache_ops.convert_fp8 (csrc/torch_bindings.cpp:705-707; test helper)."""
    require_cuda(output, input)
    kv_dtype_enum(kv_dtype)
    if not (output.is_contiguous() and input.is_contiguous()) or output.numel() != input.numel():
        raise RuntimeError("convert_fp8: tensors must be contiguous with equal numel")
    with _device_of(input):
        check(load().vb200_convert_fp8(ptr(output), ptr(input), input.numel(),
                                       dtype_enum(output.dtype), dtype_enum(input.dtype),
                                       float(scale), stream()))
