"""GPU parity of the kernels either side of the decode path through the C ABI:
advance_step_flashattn (bit-exact vs oracle/decode_loop.py) and merge_attn_states (1e-2 rel / 2e-2 abs vs the
oracle, which follows the reference test's torch golden, tests/kernels/attention/test_merge_attn_states.py:16-56;
shapes from that test's parametrisation :95-100), plus a multi-step decode loop replayed from ONE CUDA graph."""
import pytest
import torch

from oracle import decode_loop as odl
from parity_utils import seed_all

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("num_seqs,num_queries,block_size,max_blocks", [(1, 1, 16, 4), (8, 5, 16, 9), (256, 256, 16, 257),
                                                                      (1000, 777, 32, 70), (64, 0, 16, 8)])
def test_advance_step_bit_exact(num_seqs, num_queries, block_size, max_blocks):
    from vllm_b200 import _custom_ops as ops
    seed_all(num_seqs)
    tok = torch.randint(0, 32000, (num_seqs,), dtype=torch.int64)
    samp = torch.randint(0, 32000, (num_queries, 1), dtype=torch.int64)
    pos = torch.randint(0, 100, (num_seqs,), dtype=torch.int64)
    sl = torch.randint(1, block_size * (max_blocks - 1), (num_seqs,), dtype=torch.int32)
    slots = torch.randint(0, 1000, (num_seqs,), dtype=torch.int64)
    bt = torch.randint(0, 100000, (num_seqs, max_blocks), dtype=torch.int32)
    d = [t.clone().to(DEV) for t in (tok, samp, pos, sl, slots, bt)]
    ops.advance_step_flashattn(num_seqs, num_queries, block_size, *d)
    torch.cuda.synchronize()
    odl.advance_step_flashattn(num_seqs, num_queries, block_size, tok, samp, pos, sl, slots, bt)
    for got, want, name in zip(d, (tok, samp, pos, sl, slots, bt), ("tokens", "sampled", "positions", "seq_lens",
                                                                  "slot_mapping", "block_tables")):
        assert torch.equal(got.cpu(), want), name


def test_advance_step_rejects_bad_tensors():
    from vllm_b200 import _custom_ops as ops
    t64 = torch.zeros(4, dtype=torch.int64, device=DEV)
    t32 = torch.zeros(4, dtype=torch.int32, device=DEV)
    bt = torch.zeros(4, 2, dtype=torch.int32, device=DEV)
    with pytest.raises(RuntimeError, match="seq_lens"):
        ops.advance_step_flashattn(4, 4, 16, t64, t64.view(4, 1), t64, t64, t64, bt)   # seq_lens must be int32
    with pytest.raises(RuntimeError, match="sampled_token_ids"):
        ops.advance_step_flashattn(4, 4, 16, t64, t64, t64, t32, t64, bt)              # must be [num_queries, 1]


@pytest.mark.parametrize("T", [1, 3, 256, 1023])
@pytest.mark.parametrize("H", [4, 9, 32])
@pytest.mark.parametrize("D", [32, 128, 256])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
def test_merge_attn_states(T, H, D, dtype):
    from vllm_b200 import _custom_ops as ops
    seed_all(T + H + D)
    p_lse, s_lse = torch.randn(H, T), torch.randn(H, T)
    # inf / -inf entries as in the reference test (test_merge_attn_states.py:118-124)
    mask = torch.rand(H, T)
    p_lse[mask < 0.1] = float("inf")
    s_lse[(mask >= 0.1) & (mask < 0.2)] = float("-inf")
    p_out, s_out = torch.randn(T, H, D).to(dtype), torch.randn(T, H, D).to(dtype)
    want, want_lse = odl.merge_attn_states(p_out, p_lse, s_out, s_lse)
    out = torch.empty(T, H, D, dtype=dtype, device=DEV)
    out_lse = torch.empty(H, T, device=DEV)
    ops.merge_attn_states(out, p_out.to(DEV), p_lse.to(DEV), s_out.to(DEV), s_lse.to(DEV), out_lse)
    out2 = torch.empty_like(out)
    ops.merge_attn_states(out2, p_out.to(DEV), p_lse.to(DEV), s_out.to(DEV), s_lse.to(DEV))   # output_lse=None
    torch.cuda.synchronize()
    assert torch.equal(out, out2)
    torch.testing.assert_close(out.cpu().float(), want.float(), rtol=1e-2, atol=2e-2)
    ulp = {torch.float32: 2.0 ** -22, torch.float16: 2.0 ** -10, torch.bfloat16: 2.0 ** -7}[dtype]
    assert (out.cpu().float() - want.float()).abs().max() <= ulp * max(1.0, want.float().abs().max().item())
    torch.testing.assert_close(out_lse.cpu(), want_lse, rtol=1e-5, atol=1e-5)


def test_merge_equals_split_kv_attention():
    """Property: attention over [prefix ++ suffix] == merge(attention(prefix), attention(suffix)) with their LSEs
    (what the op is for: cascade / split-KV)."""
    from vllm_b200 import _custom_ops as ops
    seed_all(4)
    T, H, D, L1, L2 = 7, 8, 64, 50, 33
    q = torch.randn(T, H, D)
    k, v = torch.randn(T, H, L1 + L2, D), torch.randn(T, H, L1 + L2, D)
    logits = torch.einsum("thd,thld->thl", q, k) / D ** 0.5
    full = torch.einsum("thl,thld->thd", torch.softmax(logits, -1), v)

    def part(sl):
        lg = logits[..., sl]
        return torch.einsum("thl,thld->thd", torch.softmax(lg, -1), v[:, :, sl]), torch.logsumexp(lg, -1).t().contiguous()
    (po, pl), (so, sl_) = part(slice(0, L1)), part(slice(L1, L1 + L2))
    out = torch.empty(T, H, D, device=DEV)
    ops.merge_attn_states(out, po.contiguous().to(DEV), pl.to(DEV), so.contiguous().to(DEV), sl_.to(DEV))
    torch.testing.assert_close(out.cpu(), full, rtol=1e-4, atol=1e-5)


def test_decode_loop_in_one_graph():
    """8f row 3: greedy decode for several steps replaying ONE CUDA graph (step + advance_step_flashattn, no host
    round trip) gives the same tokens as the host-driven loop that re-uploads ids/positions/slots every step."""
    from vllm_b200.attention import B200PagedAttentionMetadata
    from vllm_b200.decode_step import CONFIGS, Decoder
    from vllm_b200 import _custom_ops as ops
    seed_all(0)
    cfg = CONFIGS["tiny-llama"]
    B, bs, ctx0, steps = 6, 16, 37, 5
    bps = (ctx0 + steps + bs - 1) // bs + 1
    nb = B * bps + 4

    def run(graph_mode):
        dec = Decoder(cfg, nb, bs, "fp8", "fp8", device=DEV, seed=3)
        dec.fill_kv_cache_random(seed=5)
        g = torch.Generator().manual_seed(1)
        bt = torch.randperm(nb, generator=g)[:B * bps].to(torch.int32).reshape(B, bps).to(DEV)
        ids = torch.randint(0, cfg.vocab_size, (B,), generator=g).to(DEV)
        sl = torch.full((B,), ctx0, dtype=torch.int32, device=DEV)
        pos = (sl - 1).to(torch.int64)
        slots = (bt[torch.arange(B), ((ctx0 - 1) // bs)].to(torch.int64) * bs + (ctx0 - 1) % bs)
        meta = B200PagedAttentionMetadata(seq_lens_tensor=sl, max_decode_seq_len=ctx0 + steps, block_tables=bt,
                                          num_decode_tokens=B, slot_mapping=slots)
        nxt = torch.empty(B, 1, dtype=torch.int64, device=DEV)
        outs = []

        def step():
            nxt.copy_(dec.greedy_step(ids, pos, meta).view(B, 1))
            ops.advance_step_flashattn(B, B, bs, ids, nxt, pos, sl, slots, bt)
        if graph_mode:
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                # warm-up on copies of the mutable state so the captured loop starts from step 0
                saved = [t.clone() for t in (ids, pos, sl, slots)]
                caches = [kv.clone() for kv in dec.kv_caches]
                step()
                for t, c in zip((ids, pos, sl, slots), saved):
                    t.copy_(c)
                for kv, c in zip(dec.kv_caches, caches):
                    kv.copy_(c)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=s):
                    step()
            torch.cuda.synchronize()
            for _ in range(steps):
                graph.replay()
                outs.append(nxt.flatten().clone())
        else:
            for i in range(steps):
                tok = dec.greedy_step(ids, pos, meta)
                outs.append(tok.clone())
                ids.copy_(tok)
                sl.add_(1)
                pos.copy_((sl - 1).to(torch.int64))
                p = ctx0 + i
                slots.copy_(bt[torch.arange(B), p // bs].to(torch.int64) * bs + p % bs)
        torch.cuda.synchronize()
        return torch.stack(outs).cpu(), sl.cpu()
    a, sla = run(False)
    b, slb = run(True)
    assert torch.equal(a, b) and torch.equal(sla, slb) and int(sla[0]) == ctx0 + steps


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("rows,vocab", [(1, 7), (3, 1000), (16, 32000), (256, 128256), (5, 50257), (2, 151936)])
def test_greedy_argmax_matches_numpy_first_max(rows, vocab, dtype):
    """Bit-exact index parity with `logits.argmax(dim=-1)` under its documented rule (first index among equal maxima),
    checked against numpy on the host; bf16 logits tie often, so the rule is exercised at every size."""
    from vllm_b200 import _custom_ops as ops
    seed_all(rows * 31 + vocab)
    logits = (torch.randn(rows, vocab) * 4).to(dtype)
    idx, val = ops.greedy_argmax(logits.to(DEV), return_max=True)
    torch.cuda.synchronize()
    ref = logits.float().numpy()
    assert idx.cpu().tolist() == ref.argmax(axis=1).tolist()
    assert torch.equal(val.cpu(), logits.float().max(dim=1).values)
    assert torch.equal(ops.greedy_argmax(logits.to(DEV)), idx)


def test_greedy_argmax_ties_nan_and_strided_rows():
    from vllm_b200 import _custom_ops as ops
    x = torch.zeros(4, 4099, dtype=torch.bfloat16)
    x[0, 17] = x[0, 4000] = 3.0                 # tie: the lower index
    x[1, :] = -float("inf")                     # all equal: index 0
    x[2, 5] = 9.0
    x[2, 77] = float("nan")                     # NaN beats every number
    x[3, 4098] = 1.0                            # last element of an odd-length row (scalar path: stride not 16 B)
    got = ops.greedy_argmax(x.to(DEV)).cpu().tolist()
    assert got == [17, 0, 77, 4098]
    wide = torch.randn(6, 2 * 4096, device=DEV, dtype=torch.float16)
    view = wide[:, :4096]                       # row stride 8192, vector path
    assert torch.equal(ops.greedy_argmax(view).cpu(), torch.from_numpy(view.float().cpu().numpy().argmax(axis=1)))
    with pytest.raises(RuntimeError, match="contiguous last"):
        ops.greedy_argmax(wide.t())
