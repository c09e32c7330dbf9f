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
                    num_seqs, num_heads, num_kv_heads, head_size, max_seq_len, block_size, kvd)
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


def flash_attn_varlen_paged(out: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                            cu_seqlens_q: torch.Tensor, seqused_k: torch.Tensor, block_table: torch.Tensor,
                            scale: float, causal: bool = True, max_seqlen_q: int = 0, kv_cache_dtype: str = "auto",
                            k_scale: Optional[torch.Tensor] = None, v_scale: Optional[torch.Tensor] = None,
                            window_left: int = -1, softcap: float = 0.0,
                            softmax_lse: Optional[torch.Tensor] = None,
                            alibi_slopes: Optional[torch.Tensor] = None) -> None:
    """Variable-length causal attention over the flash cache layout (include/vllm_b200.h
    vb200_flash_attn_varlen_paged): prefill / chunked prefill / mixed batches of the V1 backend's
    flash_attn_varlen_func call (vllm/v1/attention/backends/flash_attn.py:497-519).  query / out [total_q, num_heads,
    head_size]; key_cache / value_cache the `[num_blocks, block_size, num_kv_heads, head_size]` halves (NHD or HND
    strides) holding every sequence's tokens, the new ones included; cu_seqlens_q int32 [num_seqs + 1]; seqused_k
    int32 [num_seqs]; block_table int32 [num_seqs, max_blocks]; max_seqlen_q the longest query length when the caller
    knows it (0 = unknown); kv_cache_dtype "auto" or "fp8" (e4m3 bytes + fp32 per-tensor k_scale / v_scale);
    window_left >= 0: sliding window (flash-attn's window_size = (window_left, 0), i.e. sliding_window - 1); softcap > 0:
    logits soft-capping (flash_attn.py:342-350); softmax_lse: optional fp32 [num_heads, total_q] output, the log-sum-exp
    of every query row (return_softmax_lse=True of the reference call, used by cascade attention); alibi_slopes: optional
    fp32 [num_heads], bias -slope * |key position - query position|."""
    require_cuda(out, query, key_cache, value_cache, cu_seqlens_q, seqused_k, block_table)
    if query.dim() != 3 or query.stride(2) != 1 or query.stride(1) != query.shape[2]:
        raise RuntimeError("query must be [total_q, num_heads, head_size] with contiguous heads")
    if out.shape != query.shape or out.dtype != query.dtype or out.stride(2) != 1 or out.stride(1) != out.shape[2]:
        raise RuntimeError("out must have query's shape and dtype, with contiguous heads")
    if key_cache.dim() != 4 or key_cache.shape != value_cache.shape or key_cache.stride(3) != 1 or any(
            key_cache.size(i) > 1 and key_cache.stride(i) != value_cache.stride(i) for i in range(4)):
        raise RuntimeError("flash_attn_varlen_paged: key_cache / value_cache must be [num_blocks, block_size, "
                           "num_kv_heads, head_size] views with equal strides and a contiguous head dimension")
    kvd = kv_dtype_enum(kv_cache_dtype)
    require_cuda(k_scale, v_scale)
    if kvd == _lib.KV_AUTO and key_cache.dtype != query.dtype:
        raise RuntimeError(f"kv_cache_dtype 'auto' needs cache dtype == query dtype ({key_cache.dtype} vs {query.dtype})")
    if kvd == _lib.KV_FP8_E4M3 and (key_cache.element_size() != 1 or k_scale is None or v_scale is None):
        raise RuntimeError("fp8 kv cache must be a 1-byte tensor (uint8 / float8_e4m3fn) with k_scale and v_scale")
    for name, t in (("cu_seqlens_q", cu_seqlens_q), ("seqused_k", seqused_k)):
        if t.dtype != torch.int32 or not t.is_contiguous():
            raise RuntimeError(f"{name} must be a contiguous int32 tensor")
    # block_table may be a row / column slice of the full table (cascade attention: block_table[:1] and
    # block_table[:, num_common_kv_blocks:], flash_attn.py:722,749)
    if block_table.dtype != torch.int32 or block_table.dim() != 2 or (block_table.shape[1] > 1 and block_table.stride(1) != 1):
        raise RuntimeError("block_table must be an int32 [num_seqs, max_blocks] tensor with contiguous rows")
    require_cuda(softmax_lse, alibi_slopes)
    if alibi_slopes is not None and (alibi_slopes.dtype != torch.float32 or alibi_slopes.numel() != query.shape[1]
                                     or not alibi_slopes.is_contiguous()):
        raise RuntimeError("alibi_slopes must be a contiguous fp32 [num_heads] tensor")
    if softmax_lse is not None and (softmax_lse.dtype != torch.float32 or tuple(softmax_lse.shape) != (query.shape[1], query.shape[0])
                                    or not softmax_lse.is_contiguous()):
        raise RuntimeError("softmax_lse must be a contiguous fp32 [num_heads, total_q] tensor")
    num_seqs = seqused_k.shape[0]
    if cu_seqlens_q.numel() != num_seqs + 1 or block_table.dim() != 2 or block_table.shape[0] != num_seqs:
        raise RuntimeError("cu_seqlens_q must hold num_seqs + 1 offsets and block_table num_seqs rows")
    num_blocks, block_size, num_kv_heads, head_size = key_cache.shape
    total_q, num_heads, hs = query.shape
    if hs != head_size:
        raise RuntimeError(f"head size mismatch: query {hs}, cache {head_size}")
    strides = [key_cache.stride(i) if key_cache.size(i) > 1 else head_size for i in range(3)]
    with _device_of(query):
        check(load().vb200_flash_attn_varlen_paged_opts(
            ptr(out), ptr(query), ptr(key_cache), ptr(value_cache), ptr(cu_seqlens_q), ptr(seqused_k), ptr(block_table),
            num_seqs, total_q, num_heads, num_kv_heads, head_size, block_size, block_table.stride(0), int(max_seqlen_q),
            float(scale), 1 if causal else 0, dtype_enum(query.dtype), kvd, ptr(k_scale), ptr(v_scale),
            query.stride(0) if total_q > 1 else num_heads * head_size,
            out.stride(0) if total_q > 1 else num_heads * head_size, strides[0], strides[1], strides[2],
            int(window_left), float(softcap or 0.0), ptr(softmax_lse), total_q, ptr(alibi_slopes), stream()))


def paged_attention_v1_quant(
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
    scale_ub: Optional[torch.Tensor] = None,
) -> tuple[torch.Tensor, torch.Tensor]:
    """paged_attention_v1 followed by scaled_fp8_quant(use_per_token_if_dynamic=True) of its output, with the
    quantisation fused into the split-K reduce kernel: returns (fp8 [num_seqs, num_heads *
    head_size], fp32 scales [num_seqs, 1]), bit-identical to the two ops run back to back.  Block size 16,
    head_size % 32 == 0."""
    kvd = _pa_common(query, key_cache, value_cache, block_tables, seq_lens, kv_cache_dtype)
    require_cuda(alibi_slopes, k_scale, v_scale, scale_ub)
    lib = load()
    num_seqs, num_heads, head_size = query.shape
    qd = dtype_enum(query.dtype)
    out_q = torch.empty((num_seqs, num_heads * head_size), dtype=FP8_DTYPE, device=query.device)
    scales = torch.empty((num_seqs, 1), dtype=torch.float32, device=query.device)
    with _device_of(query):
        part = lib.vb200_paged_attention_v1_partition_size(num_seqs, num_heads, num_kv_heads, head_size, max_seq_len,
                                                           block_size, kvd)
        ws_bytes = lib.vb200_paged_attention_v1_quant_workspace_size(num_seqs, num_heads, head_size, max_seq_len, qd, part)
        ws = torch.empty(max(ws_bytes, 64), dtype=torch.uint8, device=query.device)
        check(lib.vb200_paged_attention_v1_quant(
            ptr(out_q), ptr(scales), ptr(scale_ub), ptr(query), ptr(key_cache), ptr(value_cache), num_seqs, num_heads,
            head_size, num_kv_heads, float(scale), ptr(block_tables), ptr(seq_lens), block_size, max_seq_len,
            block_tables.stride(0), ptr(alibi_slopes), qd, kvd, ptr(k_scale), ptr(v_scale), query.stride(0),
            key_cache.stride(0), key_cache.stride(1), part, ptr(ws), ws.numel(), stream()))
    return out_q, scales


def paged_attention_flash(
    out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    max_seq_len: int,
    alibi_slopes: Optional[torch.Tensor],
    kv_cache_dtype: str,
    k_scale: Optional[torch.Tensor],
    v_scale: Optional[torch.Tensor],
    partition_size: int = 0,
    window_left: int = -1,
    softcap: float = 0.0,
) -> None:
    """Decode attention over the flash cache layout (include/vllm_b200.h
    vb200_paged_attention_flash): key_cache / value_cache are the `[num_blocks, block_size, num_kv_heads, head_size]`
    halves of the V1 cache (`kv_cache.unbind(0)`, vllm/v1/attention/backends/flash_attn.py:465), NHD-contiguous or
    the HND-permuted view -- addressed through their strides like reshape_and_cache_flash.  query / out
    [num_seqs, num_heads, head_size]; one query token per sequence; seq_lens int32 = context length.  window_left >= 0:
    sliding window, the query attends its last window_left + 1 keys only (flash-attn's window_size = (window_left, 0));
    blocks left of the window are never read.  softcap > 0: logits = softcap * tanh(logits / softcap)."""
    kvd = kv_dtype_enum(kv_cache_dtype)
    require_cuda(out, query, key_cache, value_cache, block_tables, seq_lens, alibi_slopes, k_scale, v_scale)
    if query.dim() != 3 or query.stride(2) != 1 or query.stride(1) != query.shape[2]:
        raise RuntimeError("query must be [num_seqs, num_heads, head_size] with contiguous heads")
    if key_cache.dim() != 4 or key_cache.shape != value_cache.shape or any(
            key_cache.size(i) > 1 and key_cache.stride(i) != value_cache.stride(i) for i in range(4)):
        raise RuntimeError("paged_attention_flash: key_cache / value_cache must be [num_blocks, block_size, num_kv_heads, "
                           "head_size] views with equal strides")
    if key_cache.stride(3) != 1:
        raise RuntimeError("paged_attention_flash: the head dimension of the cache must be contiguous")
    if block_tables.dtype != torch.int32 or seq_lens.dtype != torch.int32:
        raise RuntimeError("block_tables and seq_lens must be int32")
    if kvd == _lib.KV_AUTO and key_cache.dtype != query.dtype:
        raise RuntimeError(f"kv_cache_dtype 'auto' needs cache dtype == query dtype ({key_cache.dtype} vs {query.dtype})")
    if kvd == _lib.KV_FP8_E4M3 and key_cache.element_size() != 1:
        raise RuntimeError("fp8 kv cache must be a 1-byte tensor (uint8 / float8_e4m3fn)")
    if out.shape != query.shape or out.dtype != query.dtype or not out.is_contiguous():
        raise RuntimeError("out must be a contiguous tensor of query's shape and dtype")
    num_blocks, block_size, num_kv_heads, head_size = key_cache.shape
    num_seqs, num_heads, hs = query.shape
    if hs != head_size:
        raise RuntimeError(f"head size mismatch: query {hs}, cache {head_size}")
    # the stride of a size-1 dimension is arbitrary in torch; the tensor map wants a well-formed one
    strides = [key_cache.stride(i) if key_cache.size(i) > 1 else head_size for i in range(3)]
    lib = load()
    qd = dtype_enum(query.dtype)
    window_left = int(window_left)
    if window_left >= 0:   # the split-K plan is made for what a sequence can contribute: the window + its first block's slack
        max_seq_len = min(max_seq_len, window_left + 16)
    with _device_of(query):
        ws, ws_bytes = None, 0
        if partition_size >= 0:
            if partition_size == 0:
                partition_size = lib.vb200_paged_attention_v1_partition_size(
                    num_seqs, num_heads, num_kv_heads, head_size, max_seq_len, block_size, kvd)
            ws_bytes = lib.vb200_paged_attention_flash_workspace_size(num_seqs, num_heads, head_size, max_seq_len, qd,
                                                                      partition_size)
            if ws_bytes > 0:
                ws = torch.empty(ws_bytes, dtype=torch.uint8, device=query.device)
        check(lib.vb200_paged_attention_flash_window(
            ptr(out), ptr(query), ptr(key_cache), ptr(value_cache), num_seqs, num_heads, head_size, num_kv_heads,
            float(scale), ptr(block_tables), ptr(seq_lens), block_size, num_blocks, max_seq_len, block_tables.stride(0),
            ptr(alibi_slopes), qd, kvd, ptr(k_scale), ptr(v_scale), query.stride(0), *strides,
            max(partition_size, 0), ptr(ws), ws_bytes, window_left, float(softcap or 0.0), stream()))


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


# ------------------------------------------------------------------------------------------------
# fp8 activation quantisation + activation  (reference: vllm/_custom_ops.py:1244-1304, :111)
# ------------------------------------------------------------------------------------------------
FP8_DTYPE = torch.float8_e4m3fn


def _quant_in(input: torch.Tensor, output: torch.Tensor):
    require_cuda(input, output)
    if input.dim() != 2 or not input.is_contiguous() or not output.is_contiguous():
        raise RuntimeError("fp8 quant: input and output must be contiguous 2-D tensors")
    if output.element_size() != 1 or output.shape[1] != input.shape[1] or output.shape[0] < input.shape[0]:
        raise RuntimeError("fp8 quant: output must be a 1-byte tensor of (at least) the input's shape")
    return input.shape[0], input.shape[1], dtype_enum(input.dtype)


def static_scaled_fp8_quant(result: torch.Tensor, input: torch.Tensor, scale: torch.Tensor) -> None:
    """torch.ops._C.static_scaled_fp8_quant (csrc/torch_bindings.cpp:575-577)."""
    m, k, dt = _quant_in(input, result)
    require_cuda(scale)
    with _device_of(input):
        check(load().vb200_static_scaled_fp8_quant(ptr(result), ptr(input), ptr(scale), m, k, dt, stream()))


def dynamic_scaled_fp8_quant(result: torch.Tensor, input: torch.Tensor, scale: torch.Tensor) -> None:
    """torch.ops._C.dynamic_scaled_fp8_quant (csrc/torch_bindings.cpp:581-584); `scale` must be
    pre-zeroed by the caller (vllm/_custom_ops.py:1297)."""
    m, k, dt = _quant_in(input, result)
    require_cuda(scale)
    with _device_of(input):
        check(load().vb200_dynamic_scaled_fp8_quant(ptr(result), ptr(input), ptr(scale), m, k, dt, stream()))


def dynamic_per_token_scaled_fp8_quant(result: torch.Tensor, input: torch.Tensor, scale: torch.Tensor,
                                       scale_ub: Optional[torch.Tensor] = None) -> None:
    """torch.ops._C.dynamic_per_token_scaled_fp8_quant (csrc/torch_bindings.cpp:588-591)."""
    m, k, dt = _quant_in(input, result)
    require_cuda(scale, scale_ub)
    if scale.numel() < m or not scale.is_contiguous():
        raise RuntimeError("per-token quant: scale must be a contiguous fp32 tensor with one entry per row")
    with _device_of(input):
        check(load().vb200_dynamic_per_token_scaled_fp8_quant(ptr(result), ptr(input), ptr(scale),
                                                              ptr(scale_ub), m, k, dt, stream()))


def scaled_fp8_quant(
    input: torch.Tensor,
    scale: Optional[torch.Tensor] = None,
    num_token_padding: Optional[int] = None,
    scale_ub: Optional[torch.Tensor] = None,
    use_per_token_if_dynamic: bool = False,
    output: Optional[torch.Tensor] = None,
) -> tuple[torch.Tensor, torch.Tensor]:
    """vllm/_custom_ops.py:1244-1304, same allocation and dispatch rules."""
    assert input.ndim == 2
    shape = input.shape
    if num_token_padding:
        shape = (max(num_token_padding, input.shape[0]), shape[1])
    if output is None:
        output = torch.empty(shape, device=input.device, dtype=FP8_DTYPE)
    else:
        assert num_token_padding is None, "padding not supported if output passed in"
        assert output.dtype == FP8_DTYPE
    if scale is None:
        if use_per_token_if_dynamic:
            scale = torch.empty((shape[0], 1), device=input.device, dtype=torch.float32)
            dynamic_per_token_scaled_fp8_quant(output, input.contiguous(), scale, scale_ub)
        else:
            scale = torch.zeros(1, device=input.device, dtype=torch.float32)
            dynamic_scaled_fp8_quant(output, input.contiguous(), scale)
    else:
        assert scale.numel() == 1, f"{scale.shape}"
        static_scaled_fp8_quant(output, input.contiguous(), scale)
    return output, scale


def silu_and_mul(out: torch.Tensor, x: torch.Tensor) -> None:
    """torch.ops._C.silu_and_mul (csrc/torch_bindings.cpp:111): x [..., 2d] -> out [..., d]."""
    require_cuda(out, x)
    d = x.shape[-1] // 2
    if not (x.is_contiguous() and out.is_contiguous()) or out.shape[-1] != d or out.dtype != x.dtype:
        raise RuntimeError("silu_and_mul: contiguous x [..., 2d] and out [..., d] of one dtype")
    n = x.numel() // max(x.shape[-1], 1)
    with _device_of(x):
        check(load().vb200_silu_and_mul(ptr(out), ptr(x), n, d, dtype_enum(x.dtype), stream()))


def absmax_scale(scale: torch.Tensor, x: torch.Tensor) -> None:
    """First pass of the dynamic per-tensor scaled_fp8_quant (fp8/common.cu:113-137) on its own:
    scale[0] = max(scale[0], max|x| / 448); `scale` is an fp32 tensor the caller zeroed."""
    require_cuda(scale, x)
    if scale.dtype != torch.float32 or not x.is_contiguous():
        raise RuntimeError("absmax_scale: fp32 scale and a contiguous input")
    with _device_of(x):
        check(load().vb200_absmax_scale(ptr(scale), ptr(x), x.numel(), dtype_enum(x.dtype), stream()))


def silu_and_mul_absmax(out: torch.Tensor, scale: torch.Tensor, x: torch.Tensor) -> None:
    """silu_and_mul that also leaves max|out| / 448 in scale[0] (pre-zeroed): activation + first pass of the dynamic
    per-tensor quantisation that follows it in the fp8 MoE pipeline, one launch."""
    require_cuda(out, scale, x)
    d = x.shape[-1] // 2
    if not (x.is_contiguous() and out.is_contiguous()) or out.shape[-1] != d or out.dtype != x.dtype:
        raise RuntimeError("silu_and_mul_absmax: contiguous x [..., 2d] and out [..., d] of one dtype")
    if scale.dtype != torch.float32:
        raise RuntimeError("silu_and_mul_absmax: fp32 scale")
    n = x.numel() // max(x.shape[-1], 1)
    with _device_of(x):
        check(load().vb200_silu_and_mul_absmax(ptr(out), ptr(scale), ptr(x), n, d, dtype_enum(x.dtype), stream()))


# ------------------------------------------------------------------------------------------------
# fp8 w8a8 linear  (reference: vllm/_custom_ops.py:688-741)
# ------------------------------------------------------------------------------------------------
def cutlass_scaled_mm_supports_fp8(cuda_device_capability: int) -> bool:
    return bool(load().vb200_cutlass_scaled_mm_supports_fp8(int(cuda_device_capability)))


def cutlass_scaled_mm_out(out: torch.Tensor, a: torch.Tensor, b: torch.Tensor, scale_a: torch.Tensor,
                          scale_b: torch.Tensor, bias: Optional[torch.Tensor] = None) -> None:
    """torch.ops._C.cutlass_scaled_mm (csrc/torch_bindings.cpp:422-426) with the checks of
    csrc/quantization/cutlass_w8a8/scaled_mm_entry.cu:169-187."""
    require_cuda(out, a, b, scale_a, scale_b, bias)
    if a.dim() != 2 or b.dim() != 2 or out.dim() != 2:
        raise RuntimeError("cutlass_scaled_mm: a, b, out must be 2-D")
    m, k = a.shape
    n = b.shape[1]
    if out.shape[0] != m or b.shape[0] != k or out.shape[1] != n:
        raise RuntimeError("cutlass_scaled_mm: shape mismatch")
    if a.stride(1) != 1 or out.stride(1) != 1 or b.stride(0) != 1:
        raise RuntimeError("cutlass_scaled_mm: a/out must be row-major and b column-major")
    if out.stride(0) % 16 or b.stride(1) % 16:
        raise RuntimeError("cutlass_scaled_mm: out.stride(0) and b.stride(1) must be multiples of 16")
    if a.dtype != FP8_DTYPE or b.dtype != FP8_DTYPE:
        raise RuntimeError("cutlass_scaled_mm: only the fp8 e4m3 path is built (a and b must be float8_e4m3fn)")
    if scale_a.dtype != torch.float32 or scale_b.dtype != torch.float32:
        raise RuntimeError("cutlass_scaled_mm: scales must be fp32")
    if scale_a.numel() not in (1, m) or scale_b.numel() not in (1, n):
        raise RuntimeError("cutlass_scaled_mm: scale_a must have 1 or M entries, scale_b 1 or N")
    if not (scale_a.is_contiguous() and scale_b.is_contiguous()):
        raise RuntimeError("cutlass_scaled_mm: scales must be contiguous")
    if bias is not None and (bias.numel() != n or not bias.is_contiguous() or bias.dtype != out.dtype):
        raise RuntimeError("cutlass_scaled_mm: bias must be a contiguous [N] tensor of the output dtype")
    with _device_of(a):
        # no scratch: split-K reduces through distributed shared memory inside the kernel's clusters
        check(load().vb200_cutlass_scaled_mm(
            ptr(out), ptr(a), ptr(b), ptr(scale_a), ptr(scale_b), ptr(bias), m, n, k, a.stride(0),
            b.stride(1), out.stride(0), scale_a.numel(), scale_b.numel(), dtype_enum(out.dtype),
            None, 0, stream()))


def cutlass_scaled_mm(a: torch.Tensor, b: torch.Tensor, scale_a: torch.Tensor, scale_b: torch.Tensor,
                      out_dtype: torch.dtype, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """vllm/_custom_ops.py:697-741 (no Triton fallback: incompatible shapes raise)."""
    assert out_dtype is torch.bfloat16 or out_dtype is torch.float16
    assert bias is None or bias.shape[0] == b.shape[1] and bias.dtype == out_dtype
    if b.shape[0] % 16 or b.shape[1] % 16:
        raise RuntimeError("cutlass_scaled_mm: K and N must be multiples of 16")
    out = torch.empty((a.shape[0], b.shape[1]), dtype=out_dtype, device=a.device)
    cutlass_scaled_mm_out(out, a, b, scale_a, scale_b, bias)
    return out


# ------------------------------------------------------------------------------------------------
# MoE routing  (reference: vllm/_custom_ops.py:1512-1548)
# ------------------------------------------------------------------------------------------------
def topk_softmax(topk_weights: torch.Tensor, topk_ids: torch.Tensor, token_expert_indices: torch.Tensor,
                 gating_output: torch.Tensor) -> None:
    """torch.ops._moe_C.topk_softmax (csrc/moe/torch_bindings.cpp:6-8)."""
    require_cuda(topk_weights, topk_ids, token_expert_indices, gating_output)
    if gating_output.dtype != torch.float32 or not gating_output.is_contiguous():
        raise RuntimeError("topk_softmax: gating_output must be contiguous fp32")
    if topk_weights.dtype != torch.float32 or token_expert_indices.dtype != torch.int32:
        raise RuntimeError("topk_softmax: topk_weights fp32, token_expert_indices int32")
    num_experts = gating_output.shape[-1]
    num_tokens = gating_output.numel() // num_experts
    topk = topk_weights.shape[-1]
    with _device_of(gating_output):
        check(load().vb200_topk_softmax(ptr(topk_weights), ptr(topk_ids), ptr(token_expert_indices),
                                        ptr(gating_output), num_tokens, num_experts, topk,
                                        dtype_enum(topk_ids.dtype), stream()))


def moe_align_block_size(topk_ids: torch.Tensor, num_experts: int, block_size: int,
                         sorted_token_ids: torch.Tensor, experts_ids: torch.Tensor,
                         num_tokens_post_pad: torch.Tensor) -> None:
    """torch.ops._moe_C.moe_align_block_size (csrc/moe/torch_bindings.cpp:18-22)."""
    require_cuda(topk_ids, sorted_token_ids, experts_ids, num_tokens_post_pad)
    for t in (topk_ids, sorted_token_ids, experts_ids, num_tokens_post_pad):
        if t.dtype != torch.int32 or not t.is_contiguous():
            raise RuntimeError("moe_align_block_size: all tensors must be contiguous int32")
    with _device_of(topk_ids):
        check(load().vb200_moe_align_block_size(
            ptr(topk_ids), topk_ids.numel(), num_experts, block_size, ptr(sorted_token_ids),
            sorted_token_ids.numel(), ptr(experts_ids), experts_ids.numel(), ptr(num_tokens_post_pad),
            None, 0, stream()))


def moe_sum(input: torch.Tensor, output: torch.Tensor) -> None:
    """torch.ops._moe_C.moe_sum (csrc/moe/torch_bindings.cpp:13): input [T, k, H] -> output [T, H]."""
    require_cuda(input, output)
    if input.dim() != 3 or not (input.is_contiguous() and output.is_contiguous()):
        raise RuntimeError("moe_sum: contiguous input [T, k, H] and output [T, H]")
    t, k, h = input.shape
    with _device_of(input):
        check(load().vb200_moe_sum(ptr(output), ptr(input), t, k, h, dtype_enum(input.dtype), stream()))


# ---------------------------------------------------------------------------------------------------
# CUTLASS-style fused-MoE API (vllm/_custom_ops.py:858-969)
# ---------------------------------------------------------------------------------------------------
def cutlass_group_gemm_supported(cuda_device_capability: int) -> bool:
    """torch.ops._C.cutlass_group_gemm_supported (csrc/quantization/cutlass_w8a8/scaled_mm_entry.cu:131-159):
    this build carries exactly one target, sm_100."""
    return int(cuda_device_capability) == 100


def _i32c(name: str, *ts) -> None:
    for t in ts:
        if t is not None and (t.dtype != torch.int32 or not t.is_contiguous()):
            raise RuntimeError(f"{name}: index tensors must be contiguous int32")


def get_cutlass_moe_mm_data(topk_ids: torch.Tensor, expert_offsets: torch.Tensor, problem_sizes1: torch.Tensor,
                            problem_sizes2: torch.Tensor, input_permutation: torch.Tensor,
                            output_permutation: torch.Tensor, num_experts: int, n: int, k: int,
                            blockscale_offsets: Optional[torch.Tensor] = None) -> None:
    """torch.ops._C.get_cutlass_moe_mm_data (vllm/_custom_ops.py:871-909; csrc/.../moe/moe_data.cu:103-159)."""
    require_cuda(topk_ids, expert_offsets, problem_sizes1, problem_sizes2, input_permutation, output_permutation,
                 blockscale_offsets)
    _i32c("get_cutlass_moe_mm_data", topk_ids, expert_offsets, problem_sizes1, problem_sizes2, input_permutation,
          output_permutation, blockscale_offsets)
    if topk_ids.dim() != 2:
        raise RuntimeError("get_cutlass_moe_mm_data: topk_ids must be [num_tokens, topk]")
    if expert_offsets.numel() < num_experts + 1 or problem_sizes1.numel() < 3 * num_experts or \
            problem_sizes2.numel() < 3 * num_experts or input_permutation.numel() < topk_ids.numel() or \
            output_permutation.numel() < topk_ids.numel():
        raise RuntimeError("get_cutlass_moe_mm_data: output tensors too small")
    if blockscale_offsets is not None and blockscale_offsets.numel() < num_experts + 1:
        raise RuntimeError("get_cutlass_moe_mm_data: blockscale_offsets too small")
    with _device_of(topk_ids):
        check(load().vb200_get_cutlass_moe_mm_data(
            ptr(topk_ids), topk_ids.size(0), topk_ids.size(1), ptr(expert_offsets), ptr(problem_sizes1),
            ptr(problem_sizes2), ptr(input_permutation), ptr(output_permutation), num_experts, n, k,
            ptr(blockscale_offsets), stream()))


def get_cutlass_pplx_moe_mm_data(expert_offsets: torch.Tensor, problem_sizes1: torch.Tensor,
                                 problem_sizes2: torch.Tensor, expert_num_tokens: torch.Tensor,
                                 num_local_experts: int, padded_m: int, n: int, k: int) -> None:
    """torch.ops._C.get_cutlass_pplx_moe_mm_data (vllm/_custom_ops.py:925-946; moe_data.cu:161-195)."""
    require_cuda(expert_offsets, problem_sizes1, problem_sizes2, expert_num_tokens)
    _i32c("get_cutlass_pplx_moe_mm_data", expert_offsets, problem_sizes1, problem_sizes2, expert_num_tokens)
    with _device_of(expert_offsets):
        check(load().vb200_get_cutlass_pplx_moe_mm_data(
            ptr(expert_offsets), ptr(problem_sizes1), ptr(problem_sizes2), ptr(expert_num_tokens),
            num_local_experts, padded_m, n, k, stream()))


def shuffle_rows_into(out: torch.Tensor, input_tensor: torch.Tensor, dst2src_map: torch.Tensor) -> torch.Tensor:
    """shuffle_rows into a caller-owned contiguous [len(dst2src_map), cols] tensor (no allocation)."""
    require_cuda(input_tensor, dst2src_map, out)
    _i32c("shuffle_rows", dst2src_map)
    if input_tensor.dim() != 2 or not input_tensor.is_contiguous():
        raise RuntimeError("shuffle_rows: input must be a contiguous 2-D tensor")
    if out.shape != (dst2src_map.shape[0], input_tensor.shape[1]) or out.dtype != input_tensor.dtype or \
            not out.is_contiguous():
        raise RuntimeError("shuffle_rows: out must be a contiguous [len(dst2src_map), cols] tensor of the input dtype")
    with _device_of(input_tensor):
        check(load().vb200_shuffle_rows(ptr(out), ptr(input_tensor), ptr(dst2src_map), out.size(0),
                                        input_tensor.size(0), input_tensor.size(1) * input_tensor.element_size(),
                                        stream()))
    return out


def shuffle_rows(input_tensor: torch.Tensor, dst2src_map: torch.Tensor) -> torch.Tensor:
    """vllm/_custom_ops.py:912-922: output[r] = input[dst2src_map[r]] (allocates the output like the reference)."""
    require_cuda(input_tensor, dst2src_map)
    out = torch.empty((dst2src_map.shape[0], input_tensor.shape[1]), device=input_tensor.device,
                      dtype=input_tensor.dtype)
    return shuffle_rows_into(out, input_tensor, dst2src_map)


def cutlass_moe_mm(out_tensors: torch.Tensor, a_tensors: torch.Tensor, b_tensors: torch.Tensor,
                   a_scales: torch.Tensor, b_scales: torch.Tensor, expert_offsets: torch.Tensor,
                   problem_sizes: torch.Tensor, a_strides: torch.Tensor, b_strides: torch.Tensor,
                   c_strides: torch.Tensor, per_act_token: bool, per_out_ch: bool) -> None:
    """torch.ops._C.cutlass_moe_mm (vllm/_custom_ops.py:949-969; scaled_mm_entry.cu:242-270): for each expert e,
    out[off_e:off_e+m_e] = (a[off_e:off_e+m_e] @ b[e].T) * a_scale * b_scale.  b_tensors [E, N, K] (K contiguous)."""
    require_cuda(out_tensors, a_tensors, b_tensors, a_scales, b_scales, expert_offsets, problem_sizes)
    if a_tensors.size(0) == 0:
        raise RuntimeError("No input A tensors provided.")
    if b_tensors.size(0) == 0:
        raise RuntimeError("No input B tensors provided.")
    if out_tensors.size(0) == 0:
        raise RuntimeError("No output tensors provided.")
    if a_tensors.dtype != torch.float8_e4m3fn:
        raise RuntimeError("A tensors must be of type float8_e4m3fn.")
    if b_tensors.dtype != torch.float8_e4m3fn:
        raise RuntimeError("B tensors must be of type float8_e4m3fn.")
    if out_tensors.dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError("Invalid output type (must be float16 or bfloat16)")
    if a_scales.dtype != torch.float32 or b_scales.dtype != torch.float32:
        raise RuntimeError("cutlass_moe_mm: scales must be float32")
    _i32c("cutlass_moe_mm", expert_offsets, problem_sizes)
    E, N, K = b_tensors.shape
    if a_tensors.dim() != 2 or a_tensors.size(1) != K or out_tensors.size(1) != N or \
            out_tensors.size(0) != a_tensors.size(0):
        raise RuntimeError("cutlass_moe_mm: shape mismatch between a, b and out")
    if not (a_tensors.is_contiguous() and b_tensors.is_contiguous() and out_tensors.is_contiguous()):
        raise RuntimeError("cutlass_moe_mm: a, b ([E, N, K]) and out must be contiguous")
    if expert_offsets.numel() < E or problem_sizes.numel() < 3 * E:
        raise RuntimeError("cutlass_moe_mm: expert_offsets / problem_sizes too small")
    # get_group_starts.cuh:58-59: scale granularity is inferred from the element counts
    per_act_token = a_scales.numel() != 1
    per_out_ch = b_scales.numel() != E
    if per_act_token and a_scales.numel() < a_tensors.size(0):
        raise RuntimeError("cutlass_moe_mm: per-token a_scales need one entry per row of a")
    if per_out_ch and b_scales.numel() != E * N:
        raise RuntimeError("cutlass_moe_mm: per-channel b_scales must be [E, N]")
    with _device_of(a_tensors):
        check(load().vb200_cutlass_moe_mm(
            ptr(out_tensors), ptr(a_tensors), ptr(b_tensors), ptr(a_scales.contiguous()), ptr(b_scales.contiguous()),
            ptr(expert_offsets), ptr(problem_sizes), a_tensors.size(0), E, N, K, int(per_act_token),
            int(per_out_ch), dtype_enum(out_tensors.dtype), stream()))


def moe_gather_quant_fp8(input: torch.Tensor, src_row: torch.Tensor, num_routed: Optional[torch.Tensor] = None,
                         scale: Optional[torch.Tensor] = None, scale_ub: Optional[torch.Tensor] = None):
    """B200 path of the CUTLASS-style MoE front end: ONE launch for `scaled_fp8_quant` followed by `shuffle_rows`
    (of the codes and of the per-token scales; fused_moe/cutlass_moe.py:142-150).  Row r of the result is row
    `src_row[r]` of `input` quantised to fp8 -- with `scale` (static per-tensor, [1] fp32) or, when `scale` is None,
    with the row's own dynamic scale, returned as `[rows, 1]` fp32.  `num_routed` (int32 [1] on the device, e.g.
    `expert_offsets[-1:]`) bounds the rows that are written.  Returns (codes, scales)."""
    require_cuda(input, src_row, num_routed, scale, scale_ub)
    _i32c("moe_gather_quant_fp8", src_row, num_routed)
    if input.dim() != 2 or not input.is_contiguous() or input.dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError("moe_gather_quant_fp8: input must be a contiguous 2-D bf16 / f16 tensor")
    rows = src_row.numel()
    out = torch.empty((rows, input.size(1)), device=input.device, dtype=torch.float8_e4m3fn)
    scales = scale if scale is not None else torch.empty((rows, 1), device=input.device, dtype=torch.float32)
    with _device_of(input):
        check(load().vb200_moe_gather_quant_fp8(
            ptr(out), None if scale is not None else ptr(scales), ptr(input), ptr(src_row), ptr(num_routed),
            ptr(scale), ptr(scale_ub), rows, input.size(0), input.size(1), dtype_enum(input.dtype), stream()))
    return out, scales


def moe_unpermute_reduce(c: torch.Tensor, dst_row: torch.Tensor, topk_weights: Optional[torch.Tensor],
                         num_routed: torch.Tensor, num_tokens: int, topk: int,
                         out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """B200 path of the CUTLASS-style MoE back end: ONE launch for `shuffle_rows(c, c_map)`, the in-place multiply by
    the routing weights and `moe_sum` (cutlass_moe.py:205-214, topk_weight_and_reduce.py:60-100) with the same
    rounding points.  c [rows, hidden]; dst_row int32 [num_tokens*topk] (flat (token, j) -> row of c; rows >=
    *num_routed count as zero); topk_weights fp32 [num_tokens, topk] or None."""
    require_cuda(c, dst_row, topk_weights, num_routed, out)
    _i32c("moe_unpermute_reduce", dst_row, num_routed)
    if c.dim() != 2 or not c.is_contiguous() or c.dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError("moe_unpermute_reduce: c must be a contiguous 2-D bf16 / f16 tensor")
    if dst_row.numel() < num_tokens * topk:
        raise RuntimeError("moe_unpermute_reduce: dst_row needs num_tokens * topk entries")
    if topk_weights is not None:
        if topk_weights.dtype != torch.float32 or topk_weights.numel() != num_tokens * topk:
            raise RuntimeError("moe_unpermute_reduce: topk_weights must be fp32 [num_tokens, topk]")
        topk_weights = topk_weights.contiguous()
    if out is None:
        out = torch.empty((num_tokens, c.size(1)), device=c.device, dtype=c.dtype)
    elif out.dtype != c.dtype or not out.is_contiguous() or out.numel() != num_tokens * c.size(1):
        raise RuntimeError("moe_unpermute_reduce: out must be a contiguous [num_tokens, hidden] tensor of c's dtype")
    with _device_of(c):
        check(load().vb200_moe_unpermute_reduce(ptr(out), ptr(c), ptr(dst_row), ptr(topk_weights), ptr(num_routed),
                                                num_tokens, topk, c.size(1), dtype_enum(c.dtype), stream()))
    return out


# ------------------------------------------------------------------------------------------------
# per-token glue between the contractions  (reference: vllm/_custom_ops.py:248-281, 368-384)
# ------------------------------------------------------------------------------------------------
def rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: Optional[torch.Tensor], head_size: int,
                     cos_sin_cache: torch.Tensor, is_neox: bool) -> None:
    """torch.ops._C.rotary_embedding (csrc/torch_bindings.cpp:216-219), launcher checks of
    csrc/pos_encoding_kernels.cu:133-185.  In place on query / key ([T, H*D] or [T, H, D])."""
    require_cuda(positions, query, key, cos_sin_cache)
    if positions.dtype != torch.int64 or positions.dim() != 1 or not positions.is_contiguous():
        raise RuntimeError("positions must have shape [num_tokens] (contiguous int64)")
    num_tokens = positions.numel()
    if query.shape[0] != num_tokens or (key is not None and key.shape[0] != num_tokens):
        raise RuntimeError("query, key and positions must have the same number of tokens")
    qh = query.numel() // max(num_tokens, 1)
    kh = key.numel() // max(num_tokens, 1) if key is not None else 0
    if qh % head_size or kh % head_size:
        raise RuntimeError("rotary_embedding: hidden sizes must be multiples of head_size")
    num_heads = qh // head_size
    num_kv_heads = kh // head_size if key is not None else num_heads
    if num_heads % num_kv_heads:
        raise RuntimeError("rotary_embedding: num_heads must be a multiple of num_kv_heads")
    if cos_sin_cache.dtype != query.dtype or not cos_sin_cache.is_contiguous():
        raise RuntimeError("rotary_embedding: cos_sin_cache must be contiguous and of the query dtype")
    head_stride = query.stride(-2) if query.dim() == 3 else head_size
    if query.stride(-1) != 1 or (key is not None and key.stride(-1) != 1):
        raise RuntimeError("rotary_embedding: head dim must be contiguous")
    with _device_of(query):
        check(load().vb200_rotary_embedding(
            ptr(positions), ptr(query), ptr(key), num_tokens, num_heads, num_kv_heads, head_size,
            cos_sin_cache.shape[1], ptr(cos_sin_cache), int(bool(is_neox)), query.stride(0),
            key.stride(0) if key is not None else 0, head_stride, dtype_enum(query.dtype), stream()))


def rotary_embedding_reshape_and_cache(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                                       value: torch.Tensor, head_size: int, cos_sin_cache: torch.Tensor, is_neox: bool,
                                       key_cache: torch.Tensor, value_cache: torch.Tensor, slot_mapping: torch.Tensor,
                                       kv_cache_dtype: str, k_scale: torch.Tensor, v_scale: torch.Tensor) -> None:
    """One launch for `rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox)` followed by
    `reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale)` -- same
    results bit for bit (reference kernels csrc/pos_encoding_kernels.cu:77-101 and
    csrc/cache_kernels.cu:210-262).  query [T, H*D], key / value [T, Hkv*D] (or [T, heads, D]), heads contiguous."""
    require_cuda(positions, query, key, value, cos_sin_cache, key_cache, value_cache, slot_mapping)
    if positions.dtype != torch.int64 or positions.dim() != 1 or not positions.is_contiguous():
        raise RuntimeError("positions must have shape [num_tokens] (contiguous int64)")
    if slot_mapping.dtype != torch.int64 or slot_mapping.shape[0] != positions.numel():
        raise RuntimeError("slot_mapping must be int64 with one entry per token")
    T = positions.numel()
    if query.shape[0] != T or key.shape[0] != T or value.shape[0] != T:
        raise RuntimeError("query, key, value and positions must have the same number of tokens")
    qh, kh = query.numel() // max(T, 1), key.numel() // max(T, 1)
    if qh % head_size or kh % head_size or value.numel() // max(T, 1) != kh:
        raise RuntimeError("rotary_embedding_reshape_and_cache: hidden sizes must be multiples of head_size")
    num_heads, num_kv_heads = qh // head_size, kh // head_size
    if cos_sin_cache.dtype != query.dtype or not cos_sin_cache.is_contiguous() or key.dtype != query.dtype \
            or value.dtype != query.dtype:
        raise RuntimeError("rotary_embedding_reshape_and_cache: query / key / value / cos_sin_cache dtypes differ")
    for t in (query, key, value):
        if t.stride(-1) != 1 or (t.dim() == 3 and t.stride(1) != head_size):
            raise RuntimeError("rotary_embedding_reshape_and_cache: heads must be contiguous")
    if not (key_cache.is_contiguous() and value_cache.is_contiguous()):
        raise RuntimeError("rotary_embedding_reshape_and_cache: caches must be contiguous")
    if key_cache.shape[1] != num_kv_heads or key_cache.shape[2] * key_cache.shape[4] != head_size:
        raise RuntimeError("rotary_embedding_reshape_and_cache: key_cache shape does not match key")
    block_size, x = key_cache.shape[3], key_cache.shape[4]
    with _device_of(query):
        check(load().vb200_rotary_embedding_reshape_and_cache(
            ptr(positions), ptr(query), ptr(key), ptr(value), ptr(cos_sin_cache), ptr(key_cache), ptr(value_cache),
            ptr(slot_mapping), T, num_heads, num_kv_heads, head_size, cos_sin_cache.shape[1], int(bool(is_neox)),
            block_size, x, query.stride(0), key.stride(0), value.stride(0), dtype_enum(query.dtype),
            kv_dtype_enum(kv_cache_dtype), ptr(k_scale), ptr(v_scale), stream()))


def _norm_rows(op: str, input: torch.Tensor, weight: torch.Tensor):
    require_cuda(input, weight)
    hidden = input.shape[-1]
    if input.stride(-1) != 1 or weight.numel() != hidden or weight.dtype != input.dtype or not weight.is_contiguous():
        raise RuntimeError(f"{op}: contiguous rows and a [hidden] weight of the input dtype are required")
    return input.numel() // max(hidden, 1), hidden


def rms_norm(out: torch.Tensor, input: torch.Tensor, weight: torch.Tensor, epsilon: float) -> None:
    """torch.ops._C.rms_norm (csrc/torch_bindings.cpp:171-173)."""
    input = input.contiguous()
    n, hidden = _norm_rows("rms_norm", input, weight)
    if not out.is_contiguous() or out.shape != input.shape or out.dtype != input.dtype:
        raise RuntimeError("rms_norm: out must be contiguous with the input's shape and dtype")
    with _device_of(input):
        check(load().vb200_rms_norm(ptr(out), ptr(input), ptr(weight), float(epsilon), n, hidden, hidden,
                                    dtype_enum(input.dtype), stream()))


def fused_add_rms_norm(input: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor, epsilon: float) -> None:
    """torch.ops._C.fused_add_rms_norm (csrc/torch_bindings.cpp:177-179); both tensors updated in place."""
    n, hidden = _norm_rows("fused_add_rms_norm", input, weight)
    if input.dim() != 2 or not residual.is_contiguous() or residual.shape != input.shape or residual.dtype != input.dtype:
        raise RuntimeError("fused_add_rms_norm: 2-D input and a contiguous residual of the same shape/dtype")
    with _device_of(input):
        check(load().vb200_fused_add_rms_norm(ptr(input), ptr(residual), ptr(weight), float(epsilon), n, hidden,
                                              input.stride(0), dtype_enum(input.dtype), stream()))


def rms_norm_dynamic_per_token_quant(input: torch.Tensor, weight: torch.Tensor, epsilon: float,
                                     quant_dtype: torch.dtype, scale_ub: Optional[torch.Tensor] = None,
                                     residual: Optional[torch.Tensor] = None) -> tuple[torch.Tensor, torch.Tensor]:
    """vllm/_custom_ops.py:368-384 (fp8 e4m3 output only; the int8 variant is outside the path)."""
    if quant_dtype != FP8_DTYPE:
        raise RuntimeError("rms_norm_dynamic_per_token_quant: only float8_e4m3fn output is built")
    n, hidden = _norm_rows("rms_norm_dynamic_per_token_quant", input, weight)
    if not input.is_contiguous() or (residual is not None and (not residual.is_contiguous() or residual.shape != input.shape)):
        raise RuntimeError("rms_norm_dynamic_per_token_quant: input and residual must be contiguous")
    require_cuda(scale_ub, residual)
    output = torch.empty_like(input, dtype=quant_dtype)
    scales = torch.empty((n, 1), device=input.device, dtype=torch.float32)
    with _device_of(input):
        check(load().vb200_rms_norm_dynamic_per_token_quant(
            ptr(output), ptr(input), ptr(weight), ptr(scales), float(epsilon), ptr(scale_ub), ptr(residual), n,
            hidden, dtype_enum(input.dtype), stream()))
    return output, scales


def silu_and_mul_dynamic_per_token_quant(x: torch.Tensor, scale_ub: Optional[torch.Tensor] = None
                                         ) -> tuple[torch.Tensor, torch.Tensor]:
    """silu_and_mul + scaled_fp8_quant(use_per_token_if_dynamic=True) in one kernel;
    bit-identical to the two reference ops run back to back.  x [T, 2d] -> (fp8 [T, d], fp32 [T, 1])."""
    require_cuda(x, scale_ub)
    if x.dim() != 2 or not x.is_contiguous():
        raise RuntimeError("silu_and_mul_dynamic_per_token_quant: contiguous 2-D input")
    t, d = x.shape[0], x.shape[1] // 2
    out = torch.empty((t, d), device=x.device, dtype=FP8_DTYPE)
    scales = torch.empty((t, 1), device=x.device, dtype=torch.float32)
    with _device_of(x):
        check(load().vb200_silu_and_mul_dynamic_per_token_quant(ptr(out), ptr(scales), ptr(x), ptr(scale_ub), t, d,
                                                                dtype_enum(x.dtype), stream()))
    return out, scales


# ---------------------------------------------------------------------------------------------------
# callers either side of the path (vllm/_custom_ops.py:142-149, 335-346)
# ---------------------------------------------------------------------------------------------------
def _verify_tensor(name: str, t: torch.Tensor, size_0: int, size_1: int, dtype: torch.dtype) -> None:
    """advance_step.cu:67-89."""
    ok = (size_0 == -1 or t.size(0) == size_0) and (size_1 == -1 or (t.dim() > 1 and t.size(1) == size_1)) and \
        t.is_contiguous() and t.dtype == dtype
    if not ok:
        raise RuntimeError(f"tensor: name = {name}, shape = {list(t.shape)} is_cont = {t.is_contiguous()}, type = "
                           f"{t.dtype} is not as expected: shape = [{size_0}, {size_1}], type = {dtype}")


def advance_step_flashattn(num_seqs: int, num_queries: int, block_size: int, input_tokens: torch.Tensor,
                           sampled_token_ids: torch.Tensor, input_positions: torch.Tensor, seq_lens: torch.Tensor,
                           slot_mapping: torch.Tensor, block_tables: torch.Tensor) -> None:
    """torch.ops._C.advance_step_flashattn (csrc/prepare_inputs/advance_step.cu:176-232): advance the decode inputs
    one step on the device."""
    require_cuda(input_tokens, sampled_token_ids, input_positions, seq_lens, slot_mapping, block_tables)
    _verify_tensor("input_tokens", input_tokens, num_seqs, -1, torch.int64)
    _verify_tensor("sampled_token_ids", sampled_token_ids, num_queries, 1, torch.int64)
    _verify_tensor("input_positions", input_positions, num_seqs, -1, torch.int64)
    _verify_tensor("seq_lens", seq_lens, num_seqs, -1, torch.int32)
    _verify_tensor("slot_mapping", slot_mapping, num_seqs, -1, torch.int64)
    _verify_tensor("block_tables", block_tables, num_seqs, -1, torch.int32)
    with _device_of(sampled_token_ids):
        check(load().vb200_advance_step_flashattn(
            num_seqs, num_queries, block_size, ptr(input_tokens), ptr(sampled_token_ids), ptr(input_positions),
            ptr(seq_lens), ptr(slot_mapping), ptr(block_tables), block_tables.stride(0), stream()))


def greedy_argmax(logits: torch.Tensor, return_max: bool = False):
    """Greedy sampling (`logits.argmax(dim=-1)`, vllm/v1/sample/sampler.py:87-96) on the logits as the lm_head GEMM
    wrote them -- no fp32 copy of the [tokens, vocab] matrix.  logits [T, vocab] f32 / f16 / bf16 with a contiguous
    last dimension.  Returns int64 [T], plus the fp32 row maxima when `return_max` (vocab-parallel merge)."""
    require_cuda(logits)
    if logits.dim() != 2 or logits.stride(1) != 1:
        raise RuntimeError("greedy_argmax: logits must be [num_tokens, vocab] with a contiguous last dimension")
    t, v = logits.shape
    idx = torch.empty(t, dtype=torch.int64, device=logits.device)
    val = torch.empty(t, dtype=torch.float32, device=logits.device) if return_max else None
    with _device_of(logits):
        check(load().vb200_greedy_argmax(ptr(idx), ptr(val), ptr(logits), t, v, logits.stride(0) if t > 1 else v,
                                         dtype_enum(logits.dtype), stream()))
    return (idx, val) if return_max else idx


def merge_attn_states(output: torch.Tensor, prefix_output: torch.Tensor, prefix_lse: torch.Tensor,
                      suffix_output: torch.Tensor, suffix_lse: torch.Tensor,
                      output_lse: Optional[torch.Tensor] = None) -> None:
    """torch.ops._C.merge_attn_states (vllm/_custom_ops.py:142-149; csrc/attention/merge_attn_states.cu:121-178)."""
    require_cuda(output, prefix_output, prefix_lse, suffix_output, suffix_lse, output_lse)
    if output.dtype not in (torch.float32, torch.float16, torch.bfloat16):
        raise RuntimeError(f"Unsupported data type of O: {output.dtype}")
    if output.dim() != 3 or prefix_output.shape != output.shape or suffix_output.shape != output.shape:
        raise RuntimeError("merge_attn_states: outputs must be [num_tokens, num_heads, head_size]")
    t, h, d = output.shape
    for x in (output, prefix_output, suffix_output):
        if x.dtype != output.dtype or not x.is_contiguous():
            raise RuntimeError("merge_attn_states: outputs must be contiguous and of one dtype")
    for x in (prefix_lse, suffix_lse, output_lse):
        if x is not None and (x.dtype != torch.float32 or tuple(x.shape) != (h, t) or not x.is_contiguous()):
            raise RuntimeError("merge_attn_states: lse tensors must be contiguous fp32 [num_heads, num_tokens]")
    with _device_of(output):
        check(load().vb200_merge_attn_states(ptr(output), ptr(output_lse), ptr(prefix_output), ptr(prefix_lse),
                                             ptr(suffix_output), ptr(suffix_lse), t, h, d, dtype_enum(output.dtype),
                                             stream()))
