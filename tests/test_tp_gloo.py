"""N > 1 path on CPU: world_size-2 (and 4) gloo process groups on 127.0.0.1 check the tensor-parallel plumbing of
vllm_b200.distributed / decode_step: the all-reduce call site, the head / column / row sharding algebra of one
decoder layer (sharded partial outputs summed by the all-reduce == the unsharded layer), replicated kv heads when
tp > num_kv_heads, and the vocab-parallel arg-max merge.  The per-rank math is plain fp32 torch (the CUDA ops need
a GPU); what is under test is the partition + collective logic the GPU path uses unchanged."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _layer_fp32(x, w_qkv, w_o, w_gu, w_dn, nh, nkv, D):
    """One decoder layer's contractions with a trivially checkable 'attention' (each q head attends to ONE cached
    token = its kv head's v): enough to exercise the head-sharding of qkv / o."""
    T = x.shape[0]
    qkv = x @ w_qkv.t()
    q = qkv[:, :nh * D].reshape(T, nh, D)
    v = qkv[:, (nh + nkv) * D:].reshape(T, nkv, D)
    g = nh // nkv
    attn = (q.abs().mean(-1, keepdim=True) * v.repeat_interleave(g, dim=1)).reshape(T, nh * D)
    h = attn @ w_o.t()                      # row-parallel: partial sum on each rank
    return h, w_gu, w_dn


def _worker(rank, world, port, cfg_name):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      GLOO_SOCKET_IFNAME="lo", OMP_NUM_THREADS="1")
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from vllm_b200.decode_step import CONFIGS, merge_vocab_parallel_argmax, tp_shard
        from vllm_b200.distributed import (get_tensor_model_parallel_rank, get_tensor_model_parallel_world_size,
                                           init_tensor_parallel, tensor_model_parallel_all_reduce)
        init_tensor_parallel()
        assert get_tensor_model_parallel_world_size() == world and get_tensor_model_parallel_rank() == rank
        t = torch.full((4,), float(rank + 1))
        out = tensor_model_parallel_all_reduce(t)
        assert out is t and torch.equal(t, torch.full((4,), world * (world + 1) / 2.0))

        cfg = CONFIGS[cfg_name]
        H, D, I = cfg.hidden_size, cfg.head_dim, cfg.intermediate_size
        NH, NKV = cfg.num_heads, cfg.num_kv_heads
        g = torch.Generator().manual_seed(0)           # identical full weights on every rank
        w_q = torch.randn(NH * D, H, generator=g, dtype=torch.float64) / 16
        w_k = torch.randn(NKV * D, H, generator=g, dtype=torch.float64) / 16
        w_v = torch.randn(NKV * D, H, generator=g, dtype=torch.float64) / 16
        w_o = torch.randn(H, NH * D, generator=g, dtype=torch.float64) / 16
        w_gu = torch.randn(2 * I, H, generator=g, dtype=torch.float64) / 16
        w_dn = torch.randn(H, I, generator=g, dtype=torch.float64) / 16
        x = torch.randn(5, H, generator=g, dtype=torch.float64)
        # unsharded reference
        full_h, _, _ = _layer_fp32(x, torch.cat([w_q, w_k, w_v]), w_o, w_gu, w_dn, NH, NKV, D)
        gu = full_h @ w_gu.t()
        full_out = (torch.nn.functional.silu(gu[:, :I]) * gu[:, I:]) @ w_dn.t()
        # this rank's shard
        sh = tp_shard(cfg, world, rank)
        (q0, q1), (k0, k1), (i0, i1) = sh["q_heads"], sh["kv_heads"], sh["inter"]
        assert sh["num_heads"] == NH // world and sh["num_kv_heads"] == max(NKV // world, 1)
        wq_s = torch.cat([w_q[q0 * D:q1 * D], w_k[k0 * D:k1 * D], w_v[k0 * D:k1 * D]])
        h_part, _, _ = _layer_fp32(x, wq_s, w_o[:, q0 * D:q1 * D], None, None, sh["num_heads"], sh["num_kv_heads"], D)
        h = tensor_model_parallel_all_reduce(h_part)                 # after o_proj (linear.py:1389)
        torch.testing.assert_close(h, full_h, rtol=1e-10, atol=1e-10)
        gu_s = h @ torch.cat([w_gu[i0:i1], w_gu[I + i0:I + i1]]).t()  # column-parallel gate_up (merged shards)
        part = (torch.nn.functional.silu(gu_s[:, :i1 - i0]) * gu_s[:, i1 - i0:]) @ w_dn[:, i0:i1].t()
        out = tensor_model_parallel_all_reduce(part)                  # after down_proj
        torch.testing.assert_close(out, full_out, rtol=1e-9, atol=1e-9)
        # vocab-parallel greedy sampling
        V = cfg.vocab_size
        logits = torch.randn(5, V, generator=g)
        logits[0, 7] = logits[0, V - 3] = 99.0                        # tie across ranks: lowest token id wins
        vs = V // world
        val, idx = logits[:, rank * vs:(rank + 1) * vs].max(dim=-1)
        tok = merge_vocab_parallel_argmax(val, idx + rank * vs, world)
        assert torch.equal(tok, logits.argmax(-1)) and int(tok[0]) == 7
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,cfg_name", [(2, "tiny-llama"), (4, "tiny-llama")])
def test_tensor_parallel_gloo(world, cfg_name):
    # The rendezvous port is picked by binding port 0 and releasing it: another process can take it in between, and a
    # busy host can miss the store's connect window.  Those are failures of the plumbing, not of what is tested, so a
    # rendezvous error gets a fresh port; an assertion inside a worker is not retried.
    last = None
    for _ in range(3):
        try:
            mp.spawn(_worker, args=(world, _free_port(), cfg_name), nprocs=world, join=True)
            return
        except Exception as e:   # noqa: BLE001  (mp.spawn wraps worker errors in ProcessRaisedException)
            if "AssertionError" in str(e):
                raise
            last = e
    raise last


def test_tp_shard_replicates_kv_heads():
    from vllm_b200.decode_step import LLAMA_3_1_70B, LLAMA_3_1_8B, tp_shard
    s = tp_shard(LLAMA_3_1_70B, 8, 5)
    assert s["q_heads"] == (40, 48) and s["kv_heads"] == (5, 6) and s["inter_size"] == 3584
    s = tp_shard(LLAMA_3_1_8B, 16, 5) if LLAMA_3_1_8B.num_heads % 16 == 0 else None
    assert s is None or (s["num_kv_heads"] == 1 and s["kv_heads"] == (2, 3))   # 2 ranks share each kv head
    assert tp_shard(LLAMA_3_1_8B, 1, 0)["q_heads"] == (0, 32)
