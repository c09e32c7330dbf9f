"""CPU tests of host-side logic that needs no GPU: block-size rules, routing tables, argument checks that run before
any kernel launch, and the oracle's sliding-window / soft-cap / ALiBi restatement against independent formulas."""
import pytest
import torch

from oracle import attention as oattn


def test_get_block_m_rule():
    """M-block of the grouped GEMM (fused_moe/fused_moe.py: get_block_m): an average expert fits one block up to 128
    rows; 256 only between 192 and 256 rows per expert; 128 again above (several blocks per expert anyway)."""
    from vllm_b200.fused_moe.fused_moe import get_block_m
    assert get_block_m(2, 8) == 16
    assert get_block_m(256, 8) == 64          # Mixtral, batch 128, top-2: 32 rows per expert
    assert get_block_m(1024, 8) == 128
    assert get_block_m(1600, 8) == 256
    assert get_block_m(2048, 8) == 256
    assert get_block_m(2056, 8) == 128
    assert get_block_m(4096, 8) == 128
    for pairs in range(1, 5000, 37):
        bm = get_block_m(pairs, 8)
        assert bm % 16 == 0 and 16 <= bm <= 256


def test_dense_linear_routing_table():
    """quantization/unquantized_linear.py: one expert owning rows 0 .. T-1, padding ids == T."""
    from vllm_b200.quantization.unquantized_linear import _routing
    bm, ids, expert_ids, num_post = _routing(5, torch.device("cpu"))
    assert bm == 16 and ids.tolist() == [0, 1, 2, 3, 4] + [5] * 11
    assert expert_ids.tolist() == [0] and num_post.tolist() == [16]
    bm, ids, expert_ids, num_post = _routing(600, torch.device("cpu"))
    assert bm == 256 and ids.numel() == 768 and int(ids.max()) == 600 and expert_ids.numel() == 3
    assert _routing(5, torch.device("cpu"))[1] is _routing(5, torch.device("cpu"))[1]     # cached


def test_ops_refuse_cpu_tensors():
    from vllm_b200.quantization import linear_16
    with pytest.raises(RuntimeError, match="CUDA"):
        linear_16(torch.zeros(2, 8, dtype=torch.bfloat16), torch.zeros(4, 8, dtype=torch.bfloat16))
    from vllm_b200.fused_moe import router_topk
    with pytest.raises(RuntimeError):
        router_topk(torch.zeros(2, 16, dtype=torch.bfloat16), torch.zeros(4, 16, dtype=torch.bfloat16), 2, True)


def test_flash_backend_argument_checks_run_before_any_launch():
    from vllm_b200.attention.flash_backend import (B200FlashAttentionImpl, cascade_attention, flash_attn_varlen_func)
    q = torch.zeros(2, 8, 128, dtype=torch.bfloat16)
    kc = torch.zeros(4, 16, 2, 128, dtype=torch.bfloat16)
    bt = torch.zeros(2, 2, dtype=torch.int32)
    sl = torch.ones(2, dtype=torch.int32)
    cu = torch.arange(3, dtype=torch.int32)
    with pytest.raises(NotImplementedError):
        flash_attn_varlen_func(q, kc, kc, 1, cu, 16, sl)                                   # no paged cache
    with pytest.raises(NotImplementedError):
        flash_attn_varlen_func(q, kc, kc, 1, cu, 16, sl, block_table=bt, window_size=(128, 0), causal=False)
    with pytest.raises(RuntimeError):
        flash_attn_varlen_func(q, kc, kc, 1, cu[:2], 16, sl, block_table=bt)               # cu_seqlens_q too short
    common = dict(output=q, query=q, key_cache=kc, value_cache=kc, cu_query_lens=cu, max_query_len=1,
                  cu_prefix_query_lens=cu[:2], prefix_kv_lens=sl[:1], suffix_kv_lens=sl, max_kv_len=32, softmax_scale=0.1,
                  logits_soft_cap=0.0, block_table=bt)
    with pytest.raises(AssertionError):   # flash_attn.py:698-707
        cascade_attention(alibi_slopes=torch.ones(8), sliding_window=(-1, -1), common_prefix_len=16, **common)
    with pytest.raises(AssertionError):
        cascade_attention(alibi_slopes=None, sliding_window=(255, 0), common_prefix_len=16, **common)
    with pytest.raises(AssertionError):
        cascade_attention(alibi_slopes=None, sliding_window=(-1, -1), common_prefix_len=24, **common)
    impl = B200FlashAttentionImpl(8, 128, 0.1, 2, sliding_window=256, logits_soft_cap=None)
    assert impl.sliding_window == (255, 0) and impl.logits_soft_cap == 0.0                 # flash_attn.py:366-374
    assert B200FlashAttentionImpl(8, 128, 0.1, 2).sliding_window == (-1, -1)


@pytest.mark.parametrize("sliding_window,soft_cap,alibi", [(7, None, False), (None, 5.0, False), (4, 3.0, True),
                                                          (None, None, True)])
def test_oracle_window_softcap_alibi_against_a_per_row_loop(sliding_window, soft_cap, alibi):
    """oracle.attention.flash_attn_varlen_paged's masks / cap / bias restated row by row with explicit key ranges (the
    reference golden builds them from triu masks, test_flash_attn.py:60-70)."""
    torch.manual_seed(3)
    H, Hkv, D, bs, nb = 4, 2, 16, 4, 12
    lens = [(3, 9), (1, 6), (5, 5)]
    ql, kl = [x[0] for x in lens], [x[1] for x in lens]
    q = torch.randn(sum(ql), H, D)
    kc, vc = torch.randn(nb, bs, Hkv, D), torch.randn(nb, bs, Hkv, D)
    bt = torch.randint(0, nb, (3, 3), dtype=torch.int32)
    slopes = torch.tensor([0.1, 0.2, 0.3, 0.05]) if alibi else None
    got = oattn.flash_attn_varlen_paged(q, kc, vc, ql, kl, bt, 0.25, sliding_window=sliding_window, soft_cap=soft_cap,
                                        alibi_slopes=slopes)
    o = 0
    for s, (n_q, n_k) in enumerate(lens):
        k = kc[bt[s].long()].reshape(-1, Hkv, D)[:n_k]
        v = vc[bt[s].long()].reshape(-1, Hkv, D)[:n_k]
        for i in range(n_q):
            p = n_k - n_q + i                                   # absolute position of this query
            lo = 0 if sliding_window is None else max(0, p - sliding_window + 1)
            for h in range(H):
                logit = (k[lo:p + 1, h // 2] @ q[o + i, h]) * 0.25
                if soft_cap:
                    logit = soft_cap * torch.tanh(logit / soft_cap)
                if alibi:
                    logit = logit - slopes[h] * (p - torch.arange(lo, p + 1)).float()
                w = torch.softmax(logit, dim=0)
                want = w @ v[lo:p + 1, h // 2]
                torch.testing.assert_close(got[o + i, h], want, rtol=1e-5, atol=1e-5)
        o += n_q


def test_decode_window_partition_plan_covers_every_visible_key_once():
    """The index arithmetic of the windowed flash-layout decode (csrc/paged_attention_flash.cuh: partition 0 starts at
    the 16-token block of the first visible key; csrc/paged_attention.cu: the reduce kernel counts partitions from the
    same block, the host plans for window_left + 16 tokens) restated in Python: for every (seq_len, window, partition)
    the launched partitions cover exactly the keys seq_len - 1 - window_left .. seq_len - 1, each once, the reduce
    kernel reads exactly the partitions that were written, and none of them lies past the planned grid."""
    for seq_len in list(range(1, 70)) + [255, 256, 257, 1000, 4097]:
        for window_left in (0, 1, 15, 16, 17, 99, 511, 512, 5000):
            for part in (16, 32, 128, 512):
                max_seq_len = 4097
                planned = min(max_seq_len, window_left + 16)                  # pa_flash_impl's clamp
                grid_np = max(1, -(-planned // part))                         # partitions launched per sequence
                win_start = max(0, seq_len - 1 - window_left)
                base = win_start & ~15
                seen, written = [], []
                for p in range(grid_np):
                    tok_begin = base + p * part
                    if tok_begin >= seq_len:
                        continue                                              # the CTA returns
                    tok_end = min(tok_begin + part, seq_len)
                    written.append(p)
                    seen += [t for t in range(tok_begin, tok_end) if t >= win_start]
                assert seen == list(range(win_start, seq_len)), (seq_len, window_left, part)
                reduce_np = min(max(1, -(-(seq_len - base) // part)), grid_np)  # paged_attention_reduce_kernel
                assert written == list(range(reduce_np)), (seq_len, window_left, part)
