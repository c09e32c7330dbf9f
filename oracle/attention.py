"""ORACLE (test infrastructure only): paged-attention decode.

Restates /root/reference/csrc/attention/attention_kernels.cuh with its rounding points:
  :289-300  logits  = scale * dot_fp32(q, k) (+ alibi_slope * (token - seq_len + 1))
  :327-340  softmax in fp32, normaliser 1 / (sum + 1e-6)
  :393-395  probabilities are ROUNDED to the query dtype before the PV product
  :415-426  fp32 PV accumulation, tail tokens zeroed
  :343-352  v2 partials: max_logits, exp_sums (un-normalised sum at the partition max), tmp_out
            (normalised partition output in the query dtype, :533)
  :632-663  v2 reduce: out = sum_j tmp_out_j * exp_sums_j*exp(m_j-m) / (sum_j ... + 1e-6)
fp8 KV: k = qdtype(float(half(fp8)) * k_scale) (quant_utils.cuh:348-359), indexed with x = 16.
The same function also stands in for the test reference ref_single_query_cached_kv_attention
(/root/reference/tests/kernels/attention/test_attention.py:68-121).
"""
import torch

from .fp8 import dequant_e4m3

PARTITION_SIZE = 512  # csrc/attention/paged_attention_v2.cu:29


def gather_kv(key_cache, value_cache, block_table, seq_len, kv_cache_dtype, k_scale, v_scale, dtype):
    """-> k, v [seq_len, num_kv_heads, head_size] in `dtype` (dequantised if fp8)."""
    nb, Hkv, Dx, bs, x = key_cache.shape
    D = Dx * x
    idx = torch.arange(seq_len)
    blk = block_table[idx // bs].long()
    off = idx % bs
    k = key_cache[blk, :, :, off, :].reshape(seq_len, Hkv, D)
    v = value_cache[blk, :, :, off]
    if kv_cache_dtype in ("fp8", "fp8_e4m3"):
        k = dequant_e4m3(k.contiguous().view(torch.uint8), float(k_scale), dtype)
        v = dequant_e4m3(v.contiguous().view(torch.uint8), float(v_scale), dtype)
    return k, v


def _attend(q, k, v, scale, alibi_slopes, seq_len, tok0):
    """q [H, D]; k, v [L, H, D] (already repeated to q heads) for tokens tok0..tok0+L-1.
    -> (max_logit [H], exp_sum [H], normalised out fp32 [H, D])."""
    logits = torch.einsum("hd,lhd->hl", q.float(), k.float()) * scale
    if alibi_slopes is not None:
        pos = torch.arange(tok0, tok0 + k.shape[0]).float() - seq_len + 1
        logits = logits + alibi_slopes.float()[:, None] * pos[None, :]
    m = logits.max(dim=-1).values
    e = torch.exp(logits - m[:, None])
    s = e.sum(-1)
    p = (e * (1.0 / (s + 1e-6))[:, None]).to(q.dtype)  # rounded to the query dtype
    out = torch.einsum("hl,lhd->hd", p.float(), v.float())
    return m, s, out


def paged_attention(query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens,
                    alibi_slopes=None, kv_cache_dtype="auto", k_scale=1.0, v_scale=1.0):
    """paged_attention_v1 semantics -> out [num_seqs, num_heads, head_size] (query dtype)."""
    S, H, D = query.shape
    G = H // num_kv_heads
    out = torch.zeros_like(query)
    for i in range(S):
        L = int(seq_lens[i])
        if L <= 0:
            continue
        k, v = gather_kv(key_cache, value_cache, block_tables[i], L, kv_cache_dtype, k_scale, v_scale,
                         query.dtype)
        k = k.repeat_interleave(G, dim=1)
        v = v.repeat_interleave(G, dim=1)
        _, _, o = _attend(query[i], k, v, scale, alibi_slopes, L, 0)
        out[i] = o.to(query.dtype)
    return out


def paged_attention_v2(query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens,
                       max_seq_len, alibi_slopes=None, kv_cache_dtype="auto", k_scale=1.0, v_scale=1.0,
                       partition_size=PARTITION_SIZE):
    """-> (out, exp_sums, max_logits, tmp_out) with the reference's partial formats."""
    S, H, D = query.shape
    G = H // num_kv_heads
    P = max(1, (max_seq_len + partition_size - 1) // partition_size)
    exp_sums = torch.zeros(S, H, P, dtype=torch.float32)
    max_logits = torch.full((S, H, P), float("-inf"), dtype=torch.float32)
    tmp_out = torch.zeros(S, H, P, D, dtype=query.dtype)
    out = torch.zeros_like(query)
    for i in range(S):
        L = int(seq_lens[i])
        if L <= 0:
            continue
        k, v = gather_kv(key_cache, value_cache, block_tables[i], L, kv_cache_dtype, k_scale, v_scale,
                         query.dtype)
        k = k.repeat_interleave(G, dim=1)
        v = v.repeat_interleave(G, dim=1)
        np_ = (L + partition_size - 1) // partition_size
        for j in range(np_):
            a, b = j * partition_size, min((j + 1) * partition_size, L)
            m, s, o = _attend(query[i], k[a:b], v[a:b], scale, alibi_slopes, L, a)
            max_logits[i, :, j] = m
            exp_sums[i, :, j] = s
            tmp_out[i, :, j] = o.to(query.dtype)
        if np_ == 1:  # attention_kernels.cuh:577-589
            out[i] = tmp_out[i, :, 0]
            continue
        m = max_logits[i, :, :np_]
        mg = m.max(dim=-1).values
        r = exp_sums[i, :, :np_] * torch.exp(m - mg[:, None])
        inv = 1.0 / (r.sum(-1) + 1e-6)
        o = (tmp_out[i, :, :np_].float() * (r * inv[:, None])[:, :, None]).sum(1)
        out[i] = o.to(query.dtype)
    return out, exp_sums, max_logits, tmp_out


def flash_attn_varlen_paged(query, key_cache, value_cache, query_lens, kv_lens, block_tables, scale, causal=True,
                            sliding_window=None, soft_cap=None, alibi_slopes=None):
    """Variable-length attention over the FLASH cache layout (prefill / chunked prefill / mixed batches): the golden of
    the reference's own test, /root/reference/tests/kernels/attention/test_flash_attn.py:25-77 (`ref_paged_attn`),
    restated: per sequence gather its kv_len tokens through the block table from key_cache / value_cache
    `[num_blocks, block_size, num_kv_heads, head_size]` (:48-54), repeat kv heads for GQA (:56-58), fp32 logits of the
    scaled query (:46,59), bottom-right aligned causal mask `triu(diagonal = kv_len - query_len + 1)` (:60-61,70),
    softmax rounded to the value dtype (:71), times V (:72).  query [sum(query_lens), H, D] -> same shape.
    sliding_window (:62-67): a query at absolute position p sees keys p - sliding_window + 1 .. p only; soft_cap
    (:68-69): logits become soft_cap * tanh(logits / soft_cap) before the mask.  alibi_slopes [H] (not in the reference
    test; flash-attn's documented bias -slope * |i + kv_len - query_len - j|, which for query_len 1 is the paged
    kernels' slope * (j - kv_len + 1), attention_kernels.cuh:283-285) is added after the cap."""
    nb, bs, Hkv, D = key_cache.shape
    outs, start = [], 0
    for i, (ql, kl) in enumerate(zip(query_lens, kv_lens)):
        q = query[start:start + ql].float() * scale
        nblk = (kl + bs - 1) // bs
        idx = block_tables[i, :nblk].long()
        k = key_cache[idx].reshape(-1, Hkv, D)[:kl]
        v = value_cache[idx].reshape(-1, Hkv, D)[:kl]
        if q.shape[1] != Hkv:
            k = torch.repeat_interleave(k, q.shape[1] // Hkv, dim=1)
            v = torch.repeat_interleave(v, q.shape[1] // Hkv, dim=1)
        attn = torch.einsum("qhd,khd->hqk", q, k.float())
        if soft_cap:
            attn = soft_cap * torch.tanh(attn / soft_cap)
        if alibi_slopes is not None:
            dist = (torch.arange(kl)[None, :] - (torch.arange(ql)[:, None] + kl - ql)).abs().float()
            attn = attn - alibi_slopes.float()[:, None, None] * dist[None]
        if causal:
            mask = torch.triu(torch.ones(ql, kl), diagonal=kl - ql + 1).bool()
            if sliding_window is not None:
                mask |= torch.triu(torch.ones(ql, kl), diagonal=kl - (ql + sliding_window) + 1).bool().logical_not()
            attn.masked_fill_(mask, float("-inf"))
        elif sliding_window is not None:
            raise ValueError("a sliding window needs the causal mask")
        attn = torch.softmax(attn, dim=-1).to(v.dtype)
        outs.append(torch.einsum("hqk,khd->qhd", attn.float(), v.float()).to(query.dtype))
        start += ql
    return torch.cat(outs, dim=0)
