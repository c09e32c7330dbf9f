"""The `Attention` module and the two opaque ops a compiled / captured model graph calls it through
(reference: /root/reference/vllm/attention/layer.py:56-310 `Attention`, :434-520
`unified_attention` / `unified_attention_with_output`, /root/reference/vllm/forward_context.py `ForwardContext`).

What the model code sees is the reference's contract:
  * `Attention(num_heads, head_size, scale, num_kv_heads, alibi_slopes, cache_config, ..., prefix=<layer name>)`
    owns `_k_scale` / `_v_scale` (fp32 tensors the backend reads), `kv_cache` (bound later with `bind_kv_cache`) and
    `impl` (the backend's AttentionImpl);
  * `forward(query, key, value)` takes no metadata: the per-step AttentionMetadata travels in a forward context that
    the runner sets around the model call (`with set_forward_context(metadata): model(...)`), and the layer is found
    by NAME inside an opaque custom op, so a tracing compiler (or a CUDA-graph capture) sees one node per layer:
        torch.ops.vllm_b200.unified_attention_with_output(query, key, value, output, layer_name)   (output buffer)
        torch.ops.vllm_b200.unified_attention(query, key, value, layer_name) -> Tensor
What is B200-specific: there is one backend (no selector cascade -- `get_attn_backend` of a vLLM that loaded the
platform plugin answers with the same class), and the registry / context are module-level dictionaries instead of a
VllmConfig, because nothing else of that object is on the hot path.
"""
import contextlib
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional

import torch
from torch import nn

from .backend import B200PagedAttentionBackend, B200PagedAttentionMetadata


# ---- forward context ------------------------------------------------------------------------------------------------
@dataclass
class ForwardContext:
    """forward_context.py: what a layer may look up during one model call."""
    attn_metadata: Any = None                      # one AttentionMetadata, or {layer_name: metadata}
    virtual_engine: int = 0
    no_compile_layers: Dict[str, "Attention"] = field(default_factory=dict)


_LAYERS: Dict[str, "Attention"] = {}               # static_forward_context: layer name -> module
_CONTEXT: Optional[ForwardContext] = None


def get_forward_context() -> ForwardContext:
    if _CONTEXT is None:
        raise RuntimeError("Forward context is not set. Use `with set_forward_context(attn_metadata):` around the "
                           "model call.")
    return _CONTEXT


@contextlib.contextmanager
def set_forward_context(attn_metadata: Any, virtual_engine: int = 0):
    global _CONTEXT
    prev = _CONTEXT
    _CONTEXT = ForwardContext(attn_metadata=attn_metadata, virtual_engine=virtual_engine, no_compile_layers=_LAYERS)
    try:
        yield _CONTEXT
    finally:
        _CONTEXT = prev


def reset_attention_layers() -> None:
    """Forget every registered layer (a new model is being built)."""
    _LAYERS.clear()


def _lookup(layer_name: str):
    ctx = get_forward_context()
    meta = ctx.attn_metadata
    if isinstance(meta, dict):
        meta = meta[layer_name]
    layer = ctx.no_compile_layers[layer_name]
    return layer, layer.kv_cache[ctx.virtual_engine], meta


# ---- the opaque ops -------------------------------------------------------------------------------------------------
@torch.library.custom_op("vllm_b200::unified_attention", mutates_args=())
def unified_attention(query: torch.Tensor, key: Optional[torch.Tensor], value: Optional[torch.Tensor],
                      layer_name: str) -> torch.Tensor:
    layer, kv_cache, meta = _lookup(layer_name)
    out = layer.impl.forward(layer, query, key, value, kv_cache, meta)
    return out.view(query.shape[0], -1)


@unified_attention.register_fake
def _(query, key, value, layer_name):
    return torch.empty_like(query).contiguous()


@torch.library.custom_op("vllm_b200::unified_attention_with_output", mutates_args=("output",))
def unified_attention_with_output(query: torch.Tensor, key: Optional[torch.Tensor], value: Optional[torch.Tensor],
                                  output: torch.Tensor, layer_name: str,
                                  output_scale: Optional[torch.Tensor] = None) -> None:
    layer, kv_cache, meta = _lookup(layer_name)
    layer.impl.forward(layer, query, key, value, kv_cache, meta, output=output, output_scale=output_scale)


@unified_attention_with_output.register_fake
def _(query, key, value, output, layer_name, output_scale=None):
    return None


# ---- the module -----------------------------------------------------------------------------------------------------
@dataclass
class CacheConfig:
    """The four fields of vllm.config.CacheConfig the layer reads (layer.py:99-108)."""
    block_size: int = 16
    cache_dtype: str = "auto"
    sliding_window: Optional[int] = None
    calculate_kv_scales: bool = False


class Attention(nn.Module):
    """Stores this step's K/V in the paged cache, attends over it, returns [T, num_heads * head_size]."""

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: Optional[int] = None,
                 alibi_slopes: Optional[List[float]] = None, cache_config: Optional[CacheConfig] = None,
                 quant_config: Any = None, logits_soft_cap: Optional[float] = None,
                 per_layer_sliding_window: Optional[int] = None, use_mla: bool = False, prefix: str = "",
                 attn_type: str = "decoder", kv_sharing_target_layer_name: Optional[str] = None,
                 pipeline_parallel_size: int = 1, **extra_impl_args) -> None:
        super().__init__()
        if use_mla:
            raise NotImplementedError("MLA is outside the B200 paged-attention backend")
        cache_config = cache_config or CacheConfig()
        sliding_window = per_layer_sliding_window if per_layer_sliding_window is not None else cache_config.sliding_window
        num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        if num_heads % num_kv_heads:
            raise ValueError(f"num_heads ({num_heads}) is not divisible by num_kv_heads ({num_kv_heads})")
        self.num_heads, self.head_size, self.num_kv_heads = num_heads, head_size, num_kv_heads
        self.kv_cache_dtype = cache_config.cache_dtype
        self.block_size = cache_config.block_size
        self.sliding_window = sliding_window
        self.calculate_kv_scales = cache_config.calculate_kv_scales
        self.use_mla = False
        # fp32 scales the kernels read from device memory; 1.0 unless a checkpoint (or calc_kv_scales) sets them
        self.register_buffer("_k_scale", torch.tensor(1.0, dtype=torch.float32), persistent=False)
        self.register_buffer("_v_scale", torch.tensor(1.0, dtype=torch.float32), persistent=False)
        self.register_buffer("_q_scale", torch.tensor(1.0, dtype=torch.float32), persistent=False)
        self._k_scale_float = self._v_scale_float = 1.0
        self.k_range, self.v_range, self.q_range = 200.0, 100.0, 200.0     # envs.{K,V,Q}_SCALE_CONSTANT defaults
        backend = B200PagedAttentionBackend
        backend.validate_head_size(head_size)
        self.backend = backend.get_name()
        self.impl = backend.get_impl_cls()(num_heads, head_size, scale, num_kv_heads, alibi_slopes, sliding_window,
                                           self.kv_cache_dtype, logits_soft_cap, attn_type,
                                           kv_sharing_target_layer_name, **extra_impl_args)
        self.use_output = backend.accept_output_buffer
        self.use_direct_call = False        # CUDA: always through the opaque op
        if prefix in _LAYERS:
            raise ValueError(f"Duplicate layer name: {prefix}")
        _LAYERS[prefix] = self
        self.layer_name = prefix
        self.attn_type = attn_type
        # placeholder until the runner binds the real cache (one entry per virtual engine / pipeline stage)
        self.kv_cache: List[torch.Tensor] = [torch.tensor([]) for _ in range(pipeline_parallel_size)]

    def bind_kv_cache(self, kv_cache: torch.Tensor, virtual_engine: int = 0) -> None:
        """worker utils `bind_kv_cache`: [2, num_blocks, block_size * num_kv_heads * head_size]."""
        want = B200PagedAttentionBackend.get_kv_cache_shape(kv_cache.shape[1], self.block_size, self.num_kv_heads,
                                                            self.head_size)
        if tuple(kv_cache.shape) != tuple(want):
            raise ValueError(f"kv cache shape {tuple(kv_cache.shape)} does not match {want}")
        self.kv_cache[virtual_engine] = kv_cache

    def forward(self, query: torch.Tensor, key: Optional[torch.Tensor], value: Optional[torch.Tensor],
                output_shape: Optional[torch.Size] = None) -> torch.Tensor:
        if self.calculate_kv_scales:
            meta = get_forward_context().attn_metadata
            if isinstance(meta, dict):
                meta = meta[self.layer_name]
            if getattr(meta, "enable_kv_scales_calculation", False):
                self.calc_kv_scales(query, key, value)
        if not self.use_output:
            return torch.ops.vllm_b200.unified_attention(query, key, value, self.layer_name)
        output_shape = query.shape if output_shape is None else output_shape
        output = torch.empty(output_shape, dtype=query.dtype, device=query.device)
        hidden = output_shape[-1]
        q3 = query.view(-1, self.num_heads, self.head_size)
        o3 = output.view(-1, self.num_heads, self.head_size)
        k3 = None if key is None else key.view(-1, self.num_kv_heads, self.head_size)
        v3 = None if value is None else value.view(-1, self.num_kv_heads, self.head_size)
        torch.ops.vllm_b200.unified_attention_with_output(q3, k3, v3, o3, self.layer_name)
        return output.view(-1, hidden)

    def calc_kv_scales(self, query, key, value) -> None:
        """layer.py:288-296: one-shot dynamic scales from the first batch."""
        self._q_scale.copy_(query.abs().max() / self.q_range)
        self._k_scale.copy_(key.abs().max() / self.k_range)
        self._v_scale.copy_(value.abs().max() / self.v_range)
        self._k_scale_float, self._v_scale_float = float(self._k_scale), float(self._v_scale)
        self.calculate_kv_scales = False

    def extra_repr(self) -> str:
        return (f"head_size={self.head_size}, num_heads={self.num_heads}, num_kv_heads={self.num_kv_heads}, "
                f"scale={self.impl.scale}, backend={type(self.impl).__name__}")


__all__ = ["Attention", "CacheConfig", "ForwardContext", "get_forward_context", "set_forward_context",
           "reset_attention_layers", "unified_attention", "unified_attention_with_output", "B200PagedAttentionMetadata"]
