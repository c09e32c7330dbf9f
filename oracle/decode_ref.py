"""ORACLE (test infrastructure only): one decode step of the dummy-weight decoder restated on the CPU from the
per-op oracles (layernorm, quant, cache, attention, moe), in the operation order of the reference model
(/root/reference/vllm/model_executor/models/llama.py:98-330: input_layernorm(+residual) -> qkv_proj -> rotary ->
Attention -> o_proj -> post_attention_layernorm(+residual) -> gate_up_proj -> SiluAndMul -> down_proj;
Mixtral: block_sparse_moe in place of the MLP, models/mixtral.py).  The weights are read from the GPU Decoder
object (copied to the CPU); caches are CPU tensors in the reference paged layout."""
import torch

from . import attention as oattn
from . import cache as ocache
from . import layernorm as oln
from . import moe as omoe
from . import quant as oq

_f32 = torch.float32


def _linear_fp8(x, w_q, w_s, dtype):
    """Fp8LinearOp.apply with dynamic per-token activation quant: quant -> scaled_mm (kernel epilogue order)."""
    xq, xs = oq.dynamic_per_token_scaled_fp8_quant(x)
    return oq.scaled_mm_epilogue_order(xq, w_q.t(), xs, w_s.reshape(1, -1), dtype)


def _linear_q(xq, xs, w_q, w_s, dtype):
    return oq.scaled_mm_epilogue_order(xq, w_q.t(), xs, w_s.reshape(1, -1), dtype)


def decode_step_logits(dec, kv_caches_cpu, input_ids, positions, block_tables, seq_lens, slot_mapping):
    """dec: vllm_b200.decode_step.Decoder (tp == 1, quant == 'fp8'); kv_caches_cpu: list of CPU caches
    [2, nb, bs*Hkv*D] (uint8 for fp8 KV), updated in place.  Returns logits fp32 [T, vocab]."""
    cfg, dt = dec.cfg, dec.dtype
    cpu = lambda t: t.detach().cpu()
    hidden = cpu(dec.embed)[input_ids]
    residual = None
    cos_sin = cpu(dec.cos_sin)
    D = cfg.head_dim
    for layer, kv in zip(dec.layers, kv_caches_cpu):
        nh, nkv = layer.nh, layer.nkv
        ln1, ln2 = cpu(layer.input_ln), cpu(layer.post_ln)
        if residual is None:
            residual = hidden.clone()
            xq, xs, _ = oln.rms_norm_dynamic_per_token_quant(hidden, ln1, cfg.rms_eps)
        else:
            xq, xs, residual = oln.rms_norm_dynamic_per_token_quant(hidden, ln1, cfg.rms_eps, None, residual)
        qkv = _linear_q(xq, xs, cpu(layer.w_qkv), cpu(layer.s_qkv), dt)
        q, k, v = qkv[:, :nh * D], qkv[:, nh * D:(nh + nkv) * D], qkv[:, (nh + nkv) * D:]
        q, k = oln.rotary_embedding(positions, q, k, D, cos_sin, True)
        T = q.shape[0]
        x = 16 // kv.element_size()
        nb = kv.shape[1]
        kc = kv[0].view(nb, nkv, D // x, -1, x)
        vc = kv[1].view(nb, nkv, D, -1)
        kvd = dec.kv_cache_dtype
        ks, vs = float(layer.kv_scales._k_scale), float(layer.kv_scales._v_scale)
        ocache.reshape_and_cache(k.reshape(T, nkv, D), v.reshape(T, nkv, D), kc, vc, slot_mapping, kvd, ks, vs)
        attn = oattn.paged_attention(q.reshape(T, nh, D).contiguous(), kc, vc, nkv, D ** -0.5, block_tables, seq_lens,
                                     None, kvd, ks, vs).reshape(T, nh * D)
        hidden = _linear_fp8(attn, cpu(layer.w_o), cpu(layer.s_o), dt)
        if cfg.num_experts:
            hidden, residual = oln.fused_add_rms_norm(hidden, residual, ln2, cfg.rms_eps)
            logits = (hidden.to(_f32) @ cpu(layer.w_gate).to(_f32).t()).to(dt)
            tw, ids, _ = omoe.topk_softmax(logits.to(_f32), cfg.top_k)
            tw = tw / tw.sum(-1, keepdim=True)
            ex = layer.experts
            hidden = omoe.fused_experts_fp8(hidden, cpu(ex["w13"]), cpu(ex["w2"]), cpu(ex["w13_scale"]),
                                            cpu(ex["w2_scale"]), tw, ids)
        else:
            xq, xs, residual = oln.rms_norm_dynamic_per_token_quant(hidden, ln2, cfg.rms_eps, None, residual)
            gu = _linear_q(xq, xs, cpu(layer.w_gate_up), cpu(layer.s_gate_up), dt)
            hq, hs = oq.dynamic_per_token_scaled_fp8_quant(oq.silu_and_mul(gu))
            hidden = _linear_q(hq, hs, cpu(layer.w_down), cpu(layer.s_down), dt)
    hidden, _ = oln.fused_add_rms_norm(hidden, residual, cpu(dec.final_ln), cfg.rms_eps)
    return hidden.to(_f32) @ cpu(dec.lm_head).to(_f32).t()
