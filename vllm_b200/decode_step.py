"""Dummy-weight decode-step driver: one forward pass of a Llama-style decoder for a batch of single-token
(decode) requests over a paged KV cache, built ONLY from the hot-path ops of this package.  It stands where the
reference's model definition + model runner stand (/root/reference/vllm/model_executor/models/llama.py:98-330,
/root/reference/vllm/v1/worker/gpu_model_runner.py:1417) -- those are host orchestration and out of scope; this
is the minimum of them needed to measure the path end to end.

Per layer (FP8 w8a8, dynamic per-token activation scales, per-channel weight scales):
  rms_norm_dynamic_per_token_quant(+residual) -> cutlass_scaled_mm (qkv, column-parallel) ->
  rotary_embedding_reshape_and_cache (one launch) -> AttentionImpl.forward (paged_attention) -> scaled_fp8_quant -> cutlass_scaled_mm (o,
  row-parallel) -> all-reduce -> rms_norm_dynamic_per_token_quant(+residual) -> cutlass_scaled_mm (gate_up) ->
  silu_and_mul_dynamic_per_token_quant -> cutlass_scaled_mm (down, row-parallel) -> all-reduce
or, for Mixtral layers, the MoE block: router GEMM -> fused_topk -> fused_experts -> all-reduce.
Tensor parallelism: heads / columns / expert-intermediate shards, one all-reduce per row-parallel GEMM
(/root/reference/vllm/model_executor/layers/linear.py:1389).  The bf16 variant (BASELINE config 2) and every model's
lm_head use the library GEMM (torch.matmul -> cuBLAS) for their unquantised projections, as the reference does;
VB200_DENSE16=1 runs them on the tcgen05 grouped GEMM with one expert instead (quantization/unquantized_linear.py:
correct, but a dense decode GEMM has too few weight-row tiles for that kernel: 8.66 vs 6.22 ms/step on config 2).
"""
from dataclasses import dataclass
from typing import List, Optional

import os

import torch

from . import _custom_ops as ops
from .attention import B200PagedAttentionBackend, B200PagedAttentionMetadata
from .attention.paged_attn import PagedAttention
from .distributed import (get_peer_allreduce, get_tensor_model_parallel_rank,
                          get_tensor_model_parallel_world_size, tensor_model_parallel_all_reduce)

FP8 = torch.float8_e4m3fn


def _dense(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """Unquantised y = x W^T (UnquantizedLinearMethod.apply, layers/linear.py:221-230)."""
    if os.environ.get("VB200_DENSE16", "0") == "1" and x.dtype in (torch.float16, torch.bfloat16) and x.shape[1] % 8 == 0:
        from .quantization.unquantized_linear import linear_16
        return linear_16(x, w)
    return torch.matmul(x, w.t())


@dataclass
class DecoderConfig:
    name: str
    hidden_size: int
    num_layers: int
    num_heads: int
    num_kv_heads: int
    head_dim: int
    intermediate_size: int
    vocab_size: int
    rope_theta: float = 500000.0
    rms_eps: float = 1e-5
    max_position: int = 131072
    num_experts: int = 0     # > 0: Mixtral-style MoE MLP
    top_k: int = 0


LLAMA_3_1_8B = DecoderConfig("Llama-3.1-8B", 4096, 32, 32, 8, 128, 14336, 128256)
LLAMA_3_1_70B = DecoderConfig("Llama-3.1-70B", 8192, 80, 64, 8, 128, 28672, 128256)
MIXTRAL_8X7B = DecoderConfig("Mixtral-8x7B", 4096, 32, 32, 8, 128, 14336, 32000, rope_theta=1e6, num_experts=8,
                             top_k=2)
TINY = DecoderConfig("tiny-llama", 256, 2, 8, 2, 32, 512, 1024, rope_theta=10000.0, max_position=4096)
TINY_MOE = DecoderConfig("tiny-mixtral", 256, 2, 8, 2, 32, 512, 1024, rope_theta=10000.0, max_position=4096,
                         num_experts=4, top_k=2)
# big enough to shard 8 ways (per-rank K of the row-parallel GEMMs >= 128), small enough for a TP-N == TP-1 test
SMALL = DecoderConfig("small-llama", 1024, 2, 16, 8, 64, 4096, 2048, rope_theta=10000.0, max_position=4096)
SMALL_MOE = DecoderConfig("small-mixtral", 1024, 2, 16, 8, 64, 2048, 2048, rope_theta=10000.0, max_position=4096,
                          num_experts=4, top_k=2)
CONFIGS = {c.name: c for c in (LLAMA_3_1_8B, LLAMA_3_1_70B, MIXTRAL_8X7B, TINY, TINY_MOE, SMALL, SMALL_MOE)}


def _dummy_fp8_weight(n: int, k: int, device, gen: torch.Generator, std: float = 0.02):
    """Random [N, K] weight quantised per output channel: returns (fp8 [N, K], fp32 scale [N, 1])."""
    w = torch.empty(n, k, device=device, dtype=torch.bfloat16).normal_(0.0, std, generator=gen)
    amax = w.abs().amax(dim=1, keepdim=True).float().clamp_(min=1e-8)
    scale = amax / 448.0
    q = (w.float() / scale).clamp_(-448.0, 448.0).to(FP8)
    return q, scale


def _dummy_weight(n: int, k: int, device, dtype, quant: str, gen: torch.Generator, rows=None, cols=None):
    """ONE seeded [N, K] weight of the unsharded model, quantised (fp8: per output channel over the FULL row, as a
    checkpoint would carry it), then cut to this rank's shard: `rows` = list of (start, stop) row ranges that are
    concatenated (column-parallel layers), `cols` = (start, stop) column range (row-parallel layers).  Every rank
    draws the same numbers, so a TP-N step computes the same function as the TP-1 step."""
    if quant == "fp8":
        w, scale = _dummy_fp8_weight(n, k, device, gen)
    else:
        w, scale = torch.empty(n, k, device=device, dtype=dtype).normal_(0.0, 0.02, generator=gen), None
    if rows is not None and not (len(rows) == 1 and rows[0] == (0, n)):
        w = torch.cat([w[a:b] for a, b in rows], dim=0)
        scale = None if scale is None else torch.cat([scale[a:b] for a, b in rows], dim=0)
    if cols is not None and cols != (0, k):
        w = w[:, cols[0]:cols[1]].contiguous()
    return w, scale


def make_cos_sin_cache(rot_dim: int, max_position: int, base: float, dtype, device) -> torch.Tensor:
    """RotaryEmbedding._compute_cos_sin_cache (vllm/model_executor/layers/rotary_embedding.py): [max_pos, rot_dim]."""
    inv_freq = 1.0 / (base ** (torch.arange(0, rot_dim, 2, dtype=torch.float32, device=device) / rot_dim))
    t = torch.arange(max_position, dtype=torch.float32, device=device)
    freqs = torch.outer(t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(dtype)


def tp_shard(cfg: DecoderConfig, tp: int, rank: int):
    """Tensor-parallel partition of one decoder layer (reference: QKVParallelLinear / MergedColumnParallelLinear /
    RowParallelLinear, vllm/model_executor/layers/linear.py): rank r owns q heads [r*H/tp, (r+1)*H/tp), the kv heads
    of those q heads (replicated when tp > num_kv_heads), and intermediate columns [r*I/tp, (r+1)*I/tp).
    Returns dict of index ranges into the UNSHARDED weights."""
    assert cfg.num_heads % tp == 0 and cfg.intermediate_size % tp == 0
    nh = cfg.num_heads // tp
    if tp <= cfg.num_kv_heads:
        assert cfg.num_kv_heads % tp == 0
        nkv = cfg.num_kv_heads // tp
        kv0 = rank * nkv
    else:
        assert tp % cfg.num_kv_heads == 0
        nkv = 1
        kv0 = rank // (tp // cfg.num_kv_heads)
    inter = cfg.intermediate_size // tp
    return {"q_heads": (rank * nh, (rank + 1) * nh), "kv_heads": (kv0, kv0 + nkv),
            "inter": (rank * inter, (rank + 1) * inter), "num_heads": nh, "num_kv_heads": nkv, "inter_size": inter}


def merge_vocab_parallel_argmax(val: torch.Tensor, idx: torch.Tensor, tp: int, group=None) -> torch.Tensor:
    """Global arg-max from per-rank (max value, global index) pairs of a vocab-parallel lm_head: all-gather the
    pairs, take the rank with the largest value (lowest rank wins ties, i.e. the lowest token id)."""
    if tp == 1:
        return idx
    import torch.distributed as dist
    vals = [torch.empty_like(val) for _ in range(tp)]
    idxs = [torch.empty_like(idx) for _ in range(tp)]
    dist.all_gather(vals, val, group=group)
    dist.all_gather(idxs, idx, group=group)
    best = torch.stack(vals).argmax(dim=0, keepdim=True)
    return torch.stack(idxs).gather(0, best).squeeze(0)


class _KVScales:
    """What an `Attention` layer object supplies to AttentionImpl.forward (abstract.py:240-247)."""

    def __init__(self, device, k_scale: float = 1.0, v_scale: float = 1.0):
        self._k_scale = torch.tensor(k_scale, dtype=torch.float32, device=device)
        self._v_scale = torch.tensor(v_scale, dtype=torch.float32, device=device)


class DecoderLayer:
    def __init__(self, cfg: DecoderConfig, tp: int, device, dtype, quant: str, kv_cache_dtype: str,
                 gen: torch.Generator):
        self.cfg, self.tp, self.quant, self.dtype = cfg, tp, quant, dtype
        H, D = cfg.hidden_size, cfg.head_dim
        shard = tp_shard(cfg, tp, get_tensor_model_parallel_rank())
        self.nh, self.nkv = shard["num_heads"], shard["num_kv_heads"]   # kv heads replicated when tp > num_kv_heads
        self.q_size, self.kv_size = self.nh * D, self.nkv * D
        self.inter = shard["inter_size"]
        self.input_ln = torch.ones(H, device=device, dtype=dtype)
        self.post_ln = torch.ones(H, device=device, dtype=dtype)
        mk = lambda n, k, rows=None, cols=None: _dummy_weight(n, k, device, dtype, quant, gen, rows, cols)  # noqa: E731
        (q0, q1), (kv0, kv1), (i0, i1) = shard["q_heads"], shard["kv_heads"], shard["inter"]
        nq, nk, I = cfg.num_heads * D, cfg.num_kv_heads * D, cfg.intermediate_size
        # QKVParallelLinear: this rank's q heads, then its k heads, then its v heads of the full [q | k | v] weight
        self.w_qkv, self.s_qkv = mk(nq + 2 * nk, H, rows=[(q0 * D, q1 * D), (nq + kv0 * D, nq + kv1 * D),
                                                           (nq + nk + kv0 * D, nq + nk + kv1 * D)])
        self.w_o, self.s_o = mk(H, nq, cols=(q0 * D, q1 * D))
        if cfg.num_experts:
            from .fused_moe import make_dummy_experts
            self.w_gate = torch.empty(cfg.num_experts, H, device=device, dtype=dtype).normal_(0.0, 0.02, generator=gen)
            self.experts = make_dummy_experts(cfg.num_experts, H, I, device, dtype, quant, gen, inter_range=(i0, i1))
        else:
            # MergedColumnParallelLinear: [gate | up], each cut to this rank's intermediate columns
            self.w_gate_up, self.s_gate_up = mk(2 * I, H, rows=[(i0, i1), (I + i0, I + i1)])
            self.w_down, self.s_down = mk(H, I, cols=(i0, i1))
        impl_cls = B200PagedAttentionBackend.get_impl_cls()
        self.attn = impl_cls(self.nh, D, D ** -0.5, self.nkv, None, None, kv_cache_dtype)
        self.kv_scales = _KVScales(device, 0.05 if kv_cache_dtype != "auto" else 1.0,
                                   0.05 if kv_cache_dtype != "auto" else 1.0)

    # ---- linear helpers -------------------------------------------------------------------------------
    def _linear_q(self, xq, xs, w, ws, out=None):
        if out is None:
            return ops.cutlass_scaled_mm(xq, w.t(), xs, ws.view(1, -1), self.dtype)
        ops.cutlass_scaled_mm_out(out, xq, w.t(), xs, ws.view(1, -1))
        return out

    def _peer_ar(self, rows):
        """The peer-memory all-reduce when it is set up and the message fits its staging buffer, else None (NCCL)."""
        par = get_peer_allreduce()
        return par if par is not None and par.fits(rows, self.cfg.hidden_size, self.dtype) else None

    def forward(self, positions, hidden, residual, kv_cache, meta: B200PagedAttentionMetadata, cos_sin,
                pending: bool = False):
        """`pending`: `hidden` is this rank's PARTIAL sum of the previous layer's row-parallel GEMM, sitting in the
        peer all-reduce staging buffer; the all-reduce is fused into the first norm of this layer.  Returns
        (hidden, residual, pending) with the same meaning for the caller."""
        cfg, eps = self.cfg, self.cfg.rms_eps
        fp8 = self.quant == "fp8"
        par = self._peer_ar(hidden.shape[0])
        # ---- attention block ----
        if pending:
            assert fp8 and par is not None and residual is not None
            xq, xs = par.all_reduce_add_rms_norm_quant(hidden, residual, self.input_ln, eps)
            qkv = self._linear_q(xq, xs, self.w_qkv, self.s_qkv)
        elif fp8:
            if residual is None:
                residual = hidden.clone()
                xq, xs = ops.rms_norm_dynamic_per_token_quant(hidden, self.input_ln, eps, FP8)
            else:
                xq, xs = ops.rms_norm_dynamic_per_token_quant(hidden, self.input_ln, eps, FP8, None, residual)
            qkv = self._linear_q(xq, xs, self.w_qkv, self.s_qkv)
        else:
            if residual is None:
                residual = hidden.clone()
                x = torch.empty_like(hidden)
                ops.rms_norm(x, hidden, self.input_ln, eps)
            else:
                ops.fused_add_rms_norm(hidden, residual, self.input_ln, eps)
                x = hidden
            qkv = _dense(x, self.w_qkv)
        q = qkv[:, :self.q_size]
        k = qkv[:, self.q_size:self.q_size + self.kv_size]
        v = qkv[:, self.q_size + self.kv_size:]
        # RoPE + this token's K/V into the paged cache: one launch; the backend then only attends (key = value = None)
        key_cache, value_cache = PagedAttention.split_kv_cache(kv_cache, self.nkv, cfg.head_dim)
        ops.rotary_embedding_reshape_and_cache(positions, q, k, v, cfg.head_dim, cos_sin, True, key_cache, value_cache,
                                               meta.slot_mapping, self.attn.kv_cache_dtype, self.kv_scales._k_scale,
                                               self.kv_scales._v_scale)
        T = q.shape[0]
        fused_ar = fp8 and par is not None   # o_proj's all-reduce is fused into the post-attention norm
        if fp8:
            if self.attn.supports_fused_output_quant(value_cache.shape[3]) and os.environ.get("VB200_ATTN_QUANT", "1") != "0":
                # the per-token fp8 quantisation of the attention output happens inside the split-K reduce kernel
                aq, a_s = self.attn.forward_quant(self.kv_scales, q, kv_cache, meta)
            else:
                attn_out = self.attn.forward(self.kv_scales, q, None, None, kv_cache, meta)
                aq, a_s = ops.scaled_fp8_quant(attn_out, use_per_token_if_dynamic=True)
            hidden = self._linear_q(aq, a_s, self.w_o, self.s_o,
                                    par.staging(T, cfg.hidden_size, self.dtype) if fused_ar else None)
        else:
            attn_out = self.attn.forward(self.kv_scales, q, None, None, kv_cache, meta)
            hidden = _dense(attn_out.view(T, -1), self.w_o)
        if not fused_ar:
            hidden = tensor_model_parallel_all_reduce(hidden)
        # ---- MLP block ----
        if cfg.num_experts:
            if fused_ar:
                hidden = par.all_reduce_add_rms_norm(hidden, residual, self.post_ln, eps)
            else:
                ops.fused_add_rms_norm(hidden, residual, self.post_ln, eps)
            from .fused_moe import moe_forward
            hidden = moe_forward(hidden, self.w_gate, self.experts, cfg.top_k, renormalize=True)
            if fused_ar:   # park the partial sum (one 2 MiB copy) so that the next norm pulls and reduces it
                st = par.staging(T, cfg.hidden_size, self.dtype)
                st.copy_(hidden)
                return st, residual, True
        elif fp8:
            if fused_ar:
                xq, xs = par.all_reduce_add_rms_norm_quant(hidden, residual, self.post_ln, eps)
            else:
                xq, xs = ops.rms_norm_dynamic_per_token_quant(hidden, self.post_ln, eps, FP8, None, residual)
            gu = self._linear_q(xq, xs, self.w_gate_up, self.s_gate_up)
            hq, hs = ops.silu_and_mul_dynamic_per_token_quant(gu)
            if fused_ar:   # leave the partial sum in the staging buffer: the next norm pulls and reduces it
                hidden = self._linear_q(hq, hs, self.w_down, self.s_down, par.staging(T, cfg.hidden_size, self.dtype))
                return hidden, residual, True
            hidden = self._linear_q(hq, hs, self.w_down, self.s_down)
        else:
            ops.fused_add_rms_norm(hidden, residual, self.post_ln, eps)
            gu = _dense(hidden, self.w_gate_up)
            act = torch.empty(gu.shape[0], self.inter, device=gu.device, dtype=gu.dtype)
            ops.silu_and_mul(act, gu)
            hidden = _dense(act, self.w_down)
        hidden = tensor_model_parallel_all_reduce(hidden)
        return hidden, residual, False


class Decoder:
    """Dummy-weight decoder + paged KV cache for one tensor-parallel rank."""

    def __init__(self, cfg: DecoderConfig, num_blocks: int, block_size: int = 16, kv_cache_dtype: str = "auto",
                 quant: str = "fp8", dtype=torch.bfloat16, device="cuda", seed: int = 0,
                 num_layers: Optional[int] = None):
        self.cfg, self.block_size, self.kv_cache_dtype = cfg, block_size, kv_cache_dtype
        self.device, self.dtype, self.quant = torch.device(device), dtype, quant
        tp = get_tensor_model_parallel_world_size()
        self.tp, self.tp_rank = tp, get_tensor_model_parallel_rank()
        assert cfg.num_heads % tp == 0 and cfg.intermediate_size % tp == 0
        # ONE weight set for every TP degree: each rank draws the full tensors from the same seed and keeps its shard
        gen = torch.Generator(device=self.device)
        gen.manual_seed(seed)
        self.num_layers = cfg.num_layers if num_layers is None else num_layers
        self.layers: List[DecoderLayer] = [DecoderLayer(cfg, tp, self.device, dtype, quant, kv_cache_dtype, gen)
                                           for _ in range(self.num_layers)]
        gen_rep = torch.Generator(device=self.device)
        gen_rep.manual_seed(seed)  # replicated tensors: identical on every rank
        self.embed = torch.empty(cfg.vocab_size, cfg.hidden_size, device=self.device, dtype=dtype).normal_(
            0.0, 0.5, generator=gen_rep)
        self.final_ln = torch.ones(cfg.hidden_size, device=self.device, dtype=dtype)
        assert cfg.vocab_size % tp == 0
        self.vocab_shard = cfg.vocab_size // tp
        lm_head = torch.empty(cfg.vocab_size, cfg.hidden_size, device=self.device, dtype=dtype).normal_(
            0.0, 0.02, generator=gen)
        self.lm_head = lm_head[self.tp_rank * self.vocab_shard:(self.tp_rank + 1) * self.vocab_shard].clone()
        del lm_head
        self.cos_sin = make_cos_sin_cache(cfg.head_dim, cfg.max_position, cfg.rope_theta, dtype, self.device)
        nkv = self.layers[0].nkv
        shape = B200PagedAttentionBackend.get_kv_cache_shape(num_blocks, block_size, nkv, cfg.head_dim)
        cache_dtype = torch.uint8 if kv_cache_dtype != "auto" else dtype
        self.kv_caches = [torch.zeros(shape, dtype=cache_dtype, device=self.device) for _ in range(self.num_layers)]

    def fill_kv_cache_random(self, seed: int = 1) -> None:
        """Synthetic cache contents (finite values only), so that attention reads realistic bytes."""
        gen = torch.Generator(device=self.device)
        gen.manual_seed(seed + self.tp_rank)   # synthetic cache contents: per rank (not a shard of one cache)
        for kv in self.kv_caches:
            if kv.dtype == torch.uint8:
                flat = kv.view(-1)
                step = 1 << 28
                for o in range(0, flat.numel(), step):
                    part = flat[o:o + step]
                    part.random_(0, 120, generator=gen)                     # finite e4m3 magnitudes
                    sign = torch.empty_like(part).random_(0, 2, generator=gen)
                    part.bitwise_or_(sign.bitwise_left_shift_(7))
            else:
                kv.normal_(0.0, 1.0, generator=gen)

    def hidden_states(self, input_ids, positions, meta: B200PagedAttentionMetadata) -> torch.Tensor:
        hidden = self.embed.index_select(0, input_ids)
        residual, pending = None, False
        for layer, kv in zip(self.layers, self.kv_caches):
            hidden, residual, pending = layer.forward(positions, hidden, residual, kv, meta, self.cos_sin, pending)
        if pending:   # the last down_proj's all-reduce, fused with the final norm
            return get_peer_allreduce().all_reduce_add_rms_norm(hidden, residual, self.final_ln, self.cfg.rms_eps)
        ops.fused_add_rms_norm(hidden, residual, self.final_ln, self.cfg.rms_eps)
        return hidden

    def forward(self, input_ids, positions, meta: B200PagedAttentionMetadata) -> torch.Tensor:
        """Returns the logits of this rank's vocabulary shard, [T, vocab/tp] (lm_head is not quantised: bf16 GEMM)."""
        return _dense(self.hidden_states(input_ids, positions, meta), self.lm_head)

    def greedy_step(self, input_ids, positions, meta: B200PagedAttentionMetadata) -> torch.Tensor:
        """One decode step with greedy sampling: next token ids [T] int64 (vocab-parallel arg-max)."""
        logits = self.forward(input_ids, positions, meta)
        if self.tp == 1:
            return ops.greedy_argmax(logits)
        idx, val = ops.greedy_argmax(logits, return_max=True)
        return merge_vocab_parallel_argmax(val, idx + self.tp_rank * self.vocab_shard, self.tp)
