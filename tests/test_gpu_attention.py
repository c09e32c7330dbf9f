"""GPU parity: paged_attention_v1 / v2 through the C ABI vs the oracle and vs the reference's own CUDA
kernels (torch.ops._ref_C from oracle/_ref/libvllm_ref_cuda.so).  Tolerance from BASELINE.json north_star:
1e-2 rel / 2e-2 abs against the reference upcast to fp32 (fp8: against the dequantised reference).
Modelled on /root/reference/tests/kernels/attention/test_attention.py:124-330."""
import os

import numpy as np
import pytest
import torch

from oracle import attention as oattn
from parity_utils import load_ref_cuda, make_block_tables, make_paged_kv, seed_all

pytestmark = pytest.mark.gpu
DEV = "cuda"
RTOL, ATOL = 1e-2, 2e-2
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _run_ours(version, q, kc, vc, hkv, scale, bt, sl, bs, alibi, kvd, ks, vs, partition_size=0):
    from vllm_b200 import _custom_ops as ops
    S, H, D = q.shape
    max_len = int(sl.max()) if sl.numel() else 0
    out = torch.full_like(q, float("nan"))
    if version == "v1":
        ops.paged_attention_v1(out, q, kc, vc, hkv, scale, bt, sl, bs, max_len, alibi, kvd, ks, vs,
                               partition_size=partition_size)
        return out, None
    P = max(1, (max_len + 511) // 512)
    es = torch.zeros(S, H, P, dtype=torch.float32, device=q.device)
    ml = torch.zeros(S, H, P, dtype=torch.float32, device=q.device)
    tmp = torch.zeros(S, H, P, D, dtype=q.dtype, device=q.device)
    ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, hkv, scale, bt, sl, bs, max_len, alibi, kvd, ks, vs)
    return out, (es, ml, tmp)


def _run_ref(q, kc, vc, hkv, scale, bt, sl, bs, alibi, kvd, ks, vs):
    ref = load_ref_cuda()
    if ref is None:
        return None
    out = torch.zeros_like(q)
    ref.paged_attention_v1(out, q, kc, vc, hkv, scale, bt, sl, bs, int(sl.max()), alibi, kvd, ks, vs, 0, 0, 0, 64, 0)
    return out


CASES = [
    # S, H, Hkv, D, bs, seq_lens
    (7, 32, 8, 128, 16, [1, 15, 16, 17, 511, 513, 1300]),
    (3, 8, 8, 64, 16, [33, 700, 2]),
    (4, 16, 2, 128, 32, [100, 1025, 31, 64]),
    (2, 12, 12, 80, 8, [9, 260]),
    (2, 40, 1, 128, 16, [77, 640]),      # 40 q heads on one kv head: 5 head groups
    (2, 6, 2, 96, 16, [50, 530]),        # GQA 3 (partial head group)
    (2, 4, 2, 256, 16, [40, 300]),
    (2, 4, 4, 32, 16, [40, 300]),
    (2, 4, 2, 112, 16, [40, 300]),
    (2, 4, 2, 192, 16, [40, 300]),
]


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("kvd", ["auto", "fp8"])
@pytest.mark.parametrize("case", CASES, ids=lambda c: f"S{c[0]}H{c[1]}kv{c[2]}D{c[3]}b{c[4]}")
def test_paged_attention(version, dtype, kvd, case):
    S, H, Hkv, D, bs, seq_lens = case
    seed_all(S * 1000 + D)
    nb = S * ((max(seq_lens) + bs - 1) // bs) + 4
    ksc, vsc = (0.043, 0.029) if kvd == "fp8" else (1.0, 1.0)
    kc, vc = make_paged_kv(nb, bs, Hkv, D, dtype, kvd, dist="normal", k_scale=ksc, v_scale=vsc)
    # q is a slice of a fused qkv projection: row stride != H*D (attention_kernels.cuh:172-174)
    qkv = (torch.randn(S, H + 2 * Hkv, D) * 0.5).to(dtype)
    q = qkv[:, :H]
    bt = make_block_tables(S, seq_lens, bs, nb)
    sl = torch.tensor(seq_lens, dtype=torch.int32)
    scale = float(D ** -0.5)
    ks, vs = torch.tensor(ksc), torch.tensor(vsc)
    want = oattn.paged_attention(q, kc, vc, Hkv, scale, bt, sl, None, kvd, ksc, vsc)
    qd = qkv.to(DEV)[:, :H]
    args = (qd, kc.to(DEV), vc.to(DEV), Hkv, scale, bt.to(DEV), sl.to(DEV), bs, None, kvd, ks.to(DEV), vs.to(DEV))
    got, partials = _run_ours(version, *args)
    torch.cuda.synchronize()
    torch.testing.assert_close(got.cpu().float(), want.float(), rtol=RTOL, atol=ATOL)
    ref_out = _run_ref(*args)
    if ref_out is not None:
        torch.testing.assert_close(got.float(), ref_out.float(), rtol=RTOL, atol=ATOL)
    if version == "v2" and max(seq_lens) > 512:
        # caller-visible partials keep the reference's formats (attention_kernels.cuh:343-352)
        _, es_o, ml_o, tmp_o = oattn.paged_attention_v2(q, kc, vc, Hkv, scale, bt, sl, max(seq_lens), None, kvd, ksc, vsc)
        es, ml, tmp = [t.cpu() for t in partials]
        for i, L in enumerate(seq_lens):
            n = (L + 511) // 512
            if n < 2:
                continue  # single-partition rows are written straight to `out`
            torch.testing.assert_close(ml[i, :, :n], ml_o[i, :, :n], rtol=1e-3, atol=1e-3)
            torch.testing.assert_close(es[i, :, :n], es_o[i, :, :n], rtol=2e-3, atol=1e-3)
            torch.testing.assert_close(tmp[i, :, :n].float(), tmp_o[i, :, :n].float(), rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("version", ["v1", "v2"])
def test_paged_attention_alibi_and_uniform(version):
    seed_all(11)
    S, H, Hkv, D, bs = 3, 8, 4, 128, 16
    seq_lens = [19, 555, 1024]
    nb = S * 64 + 2
    kc, vc = make_paged_kv(nb, bs, Hkv, D, torch.bfloat16, dist="uniform")
    q = (torch.rand(S, H, D) * 2 - 1).bfloat16() * D ** -0.5
    bt = make_block_tables(S, seq_lens, bs, nb)
    sl = torch.tensor(seq_lens, dtype=torch.int32)
    alibi = torch.randn(H) * 0.05
    want = oattn.paged_attention(q, kc, vc, Hkv, 0.09, bt, sl, alibi)
    one = torch.ones(1, device=DEV)
    got, _ = _run_ours(version, q.to(DEV), kc.to(DEV), vc.to(DEV), Hkv, 0.09, bt.to(DEV), sl.to(DEV), bs, alibi.to(DEV),
                       "auto", one, one)
    torch.testing.assert_close(got.cpu().float(), want.float(), rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("name", ["pa_gqa_uniform", "pa_mha_normal_alibi"])
def test_paged_attention_vs_reference_cpu_golden(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    bf = lambda a: torch.from_numpy(a.view(np.int16).copy()).view(torch.bfloat16).to(DEV)
    q, kc, vc = bf(g["q"]), bf(g["key_cache"]), bf(g["value_cache"])
    bt, sl = torch.from_numpy(g["block_tables"]).to(DEV), torch.from_numpy(g["seq_lens"]).to(DEV)
    alibi = torch.from_numpy(g["alibi"]).to(DEV) if g["alibi"].size else None
    one = torch.ones(1, device=DEV)
    for version, key in (("v1", "out_v1"), ("v2", "out_v2")):
        got, _ = _run_ours(version, q, kc, vc, int(g["num_kv_heads"]), float(g["scale"]), bt, sl, 16, alibi, "auto", one, one)
        torch.testing.assert_close(got.float(), bf(g[key]).float(), rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("partition_size", [-1, 0, 64, 256, 1024])
def test_paged_attention_v1_split_k_invariance(partition_size):
    """Split-K is an implementation detail of v1: every partition size gives the same answer (within the
    bf16 rounding of the partials) as the unsplit kernel."""
    seed_all(5)
    S, H, Hkv, D, bs = 4, 8, 2, 128, 16
    seq_lens = [2048, 1, 1500, 333]
    nb = S * 128
    kc, vc = make_paged_kv(nb, bs, Hkv, D, torch.bfloat16, dist="normal")
    q = (torch.randn(S, H, D) * 0.3).bfloat16()
    bt = make_block_tables(S, seq_lens, bs, nb)
    sl = torch.tensor(seq_lens, dtype=torch.int32)
    one = torch.ones(1, device=DEV)
    want = oattn.paged_attention(q, kc, vc, Hkv, D ** -0.5, bt, sl)
    got, _ = _run_ours("v1", q.to(DEV), kc.to(DEV), vc.to(DEV), Hkv, D ** -0.5, bt.to(DEV), sl.to(DEV), bs, None, "auto",
                       one, one, partition_size=partition_size)
    torch.testing.assert_close(got.cpu().float(), want.float(), rtol=RTOL, atol=ATOL)


def test_paged_attention_shared_blocks_and_errors():
    from vllm_b200 import _custom_ops as ops
    seed_all(6)
    S, H, Hkv, D, bs, nb = 6, 8, 8, 128, 16, 40
    seq_lens = [300] * S
    kc, vc = make_paged_kv(nb, bs, Hkv, D, torch.float16, dist="normal")
    q = (torch.randn(S, H, D) * 0.3).half()
    bt = make_block_tables(S, seq_lens, bs, nb, shared=True)  # prefix sharing: blocks referenced many times
    sl = torch.tensor(seq_lens, dtype=torch.int32)
    one = torch.ones(1, device=DEV)
    want = oattn.paged_attention(q, kc, vc, Hkv, 0.1, bt, sl)
    got, _ = _run_ours("v1", q.to(DEV), kc.to(DEV), vc.to(DEV), Hkv, 0.1, bt.to(DEV), sl.to(DEV), bs, None, "auto", one, one)
    torch.testing.assert_close(got.cpu().float(), want.float(), rtol=RTOL, atol=ATOL)
    out = torch.empty_like(q, device=DEV)
    with pytest.raises(RuntimeError, match="Unsupported head size"):
        bad_q = torch.zeros(1, 2, 48, dtype=torch.float16, device=DEV)
        ops.paged_attention_v1(torch.empty_like(bad_q), bad_q, torch.zeros(2, 2, 6, 16, 8, dtype=torch.float16, device=DEV),
                               torch.zeros(2, 2, 48, 16, dtype=torch.float16, device=DEV), 2, 1.0,
                               torch.zeros(1, 1, dtype=torch.int32, device=DEV), torch.ones(1, dtype=torch.int32, device=DEV),
                               16, 1, None, "auto", one, one)
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        ops.paged_attention_v1(out, q.to(DEV), kc.to(DEV), vc.to(DEV), Hkv, 0.1, bt.to(DEV), sl.to(DEV), bs, 300, None,
                               "fp8_e5m2", one, one)


def test_paged_attention_cuda_graph_capture():
    """The op must be capturable (no host sync, scratch through the caching allocator)."""
    seed_all(8)
    S, H, Hkv, D, bs, nb = 8, 32, 8, 128, 16, 8 * 128
    seq_lens = [2048] * S
    kc, vc = make_paged_kv(nb, bs, Hkv, D, torch.bfloat16, "fp8", dist="normal", k_scale=0.05, v_scale=0.05)
    q = (torch.randn(S, H, D) * 0.3).bfloat16().to(DEV)
    bt = make_block_tables(S, seq_lens, bs, nb).to(DEV)
    sl = torch.tensor(seq_lens, dtype=torch.int32, device=DEV)
    sc = torch.full((1,), 0.05, device=DEV)
    kc, vc = kc.to(DEV), vc.to(DEV)
    eager, _ = _run_ours("v1", q, kc, vc, Hkv, 0.088, bt, sl, bs, None, "fp8", sc, sc)
    torch.cuda.synchronize()
    from vllm_b200 import _custom_ops as ops
    out = torch.zeros_like(q)
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        ops.paged_attention_v1(out, q, kc, vc, Hkv, 0.088, bt, sl, bs, 2048, None, "fp8", sc, sc)  # warm-up
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            ops.paged_attention_v1(out, q, kc, vc, Hkv, 0.088, bt, sl, bs, 2048, None, "fp8", sc, sc)
    out.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, eager)


@pytest.mark.parametrize("kvd", ["auto", "fp8"])
@pytest.mark.parametrize("case", [(7, 32, 8, 128, [1, 15, 16, 17, 511, 513, 1300]), (3, 8, 8, 64, [33, 700, 2]),
                                  (2, 40, 1, 128, [77, 640]), (2, 4, 2, 256, [40, 300]), (2, 64, 8, 128, [5000, 300]),
                                  (4, 16, 2, 128, [100, 90, 80, 70]), (2, 128, 8, 128, [600, 40]),
                                  (2, 72, 8, 128, [700, 30]), (3, 24, 8, 64, [1500, 2, 513])],
                         ids=lambda c: f"S{c[0]}H{c[1]}kv{c[2]}D{c[3]}")
def test_paged_attention_v1_quant_matches_unfused(kvd, case):
    """paged_attention_v1_quant (per-token fp8 quantisation inside the split-K reduce kernel) == paged_attention_v1
    followed by scaled_fp8_quant(use_per_token_if_dynamic=True), bit for bit (codes and scales)."""
    from vllm_b200 import _custom_ops as ops
    S, H, Hkv, D, seq_lens = case
    bs, dtype = 16, torch.bfloat16
    seed_all(S + H + D)
    nb = S * ((max(seq_lens) + bs - 1) // bs) + 4
    ksc, vsc = (0.043, 0.029) if kvd == "fp8" else (1.0, 1.0)
    kc, vc = make_paged_kv(nb, bs, Hkv, D, dtype, kvd, dist="normal", k_scale=ksc, v_scale=vsc)
    q = (torch.randn(S, H, D) * 0.5).to(dtype).to(DEV)
    bt = make_block_tables(S, seq_lens, bs, nb).to(DEV)
    sl = torch.tensor(seq_lens, dtype=torch.int32, device=DEV)
    ks, vs = torch.tensor([ksc], device=DEV), torch.tensor([vsc], device=DEV)
    kc, vc = kc.to(DEV), vc.to(DEV)
    scale = float(D ** -0.5)
    out = torch.empty_like(q)
    ops.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, bs, max(seq_lens), None, kvd, ks, vs)
    want_q, want_s = ops.scaled_fp8_quant(out.view(S, H * D), None, use_per_token_if_dynamic=True)
    got_q, got_s = ops.paged_attention_v1_quant(q, kc, vc, Hkv, scale, bt, sl, bs, max(seq_lens), None, kvd, ks, vs)
    torch.cuda.synchronize()
    assert got_q.shape == (S, H * D) and got_s.shape == (S, 1)
    assert torch.equal(got_s, want_s.view(S, 1))
    assert torch.equal(got_q.view(torch.uint8), want_q.view(torch.uint8))


# ---- BASELINE.json-sized problems: the per-layer attention call of configs[1], [2], [4] (per-GPU shard at TP 8) and a
# ragged batch.  Too big for the CPU oracle, so the checks are (1) the reference's own CUDA kernel on the same inputs
# when oracle/_ref is present, (2) size-independent properties: the split-K plan does not change the result, v2 == v1,
# a sequence's output does not depend on its batch neighbours (a sub-batch at oracle size reproduces its rows), and
# that sub-batch against the CPU oracle. ------------------------------------------------------------------------------
BASELINE_SHAPES = [
    # name, S, H, Hkv, D, kv dtype, sequence lengths
    ("cfg2_bf16_64x2k", 64, 32, 8, 128, "auto", [2048] * 64),
    ("cfg3_fp8_256x4k", 256, 32, 8, 128, "fp8", [4096] * 256),
    ("cfg5_tp8_shard_fp8_128x8k", 128, 8, 1, 128, "fp8", [8192] * 128),
    ("ragged_bf16_64", 64, 32, 8, 128, "auto", [1 + (i * 977) % 2048 for i in range(64)]),
]


def _device_paged_kv(nb, bs, Hkv, D, dtype, kvd, ksc, vsc):
    """Caches in the reference paged layout generated on the GPU (the CPU helper would take minutes at these sizes)."""
    x = 16 if kvd == "fp8" else 8
    k = torch.randn(nb, Hkv, D // x, bs, x, device=DEV)
    v = torch.randn(nb, Hkv, D, bs, device=DEV)
    if kvd == "fp8":
        return ((k / ksc).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8),
                (v / vsc).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8))
    return k.to(dtype), v.to(dtype)


@pytest.mark.parametrize("shape", BASELINE_SHAPES, ids=lambda s: s[0])
def test_paged_attention_baseline_sized(shape):
    name, S, H, Hkv, D, kvd, seq_lens = shape
    bs, dtype = 16, torch.bfloat16
    seed_all(len(name))
    blocks_per_seq = (max(seq_lens) + bs - 1) // bs
    nb = S * blocks_per_seq
    ksc, vsc = (0.043, 0.029) if kvd == "fp8" else (1.0, 1.0)
    kc, vc = _device_paged_kv(nb, bs, Hkv, D, dtype, kvd, ksc, vsc)
    q = (torch.randn(S, H, D, device=DEV) * 0.5).to(dtype)
    bt = torch.randperm(nb, device=DEV).to(torch.int32).view(S, blocks_per_seq)
    sl = torch.tensor(seq_lens, dtype=torch.int32, device=DEV)
    scale = float(D ** -0.5)
    ks, vs = torch.tensor(ksc, device=DEV), torch.tensor(vsc, device=DEV)
    args = (q, kc, vc, Hkv, scale, bt, sl, bs, None, kvd, ks, vs)
    got, _ = _run_ours("v1", *args)                       # the library's own split-K plan
    torch.cuda.synchronize()
    assert torch.isfinite(got.float()).all()
    ref_out = _run_ref(*args)
    if ref_out is not None:
        torch.testing.assert_close(got.float(), ref_out.float(), rtol=RTOL, atol=ATOL)
    unsplit, _ = _run_ours("v1", *args, partition_size=-1)
    torch.testing.assert_close(got.float(), unsplit.float(), rtol=RTOL, atol=ATOL)
    v2, _ = _run_ours("v2", *args)
    torch.testing.assert_close(got.float(), v2.float(), rtol=RTOL, atol=ATOL)
    # three sequences on their own (the longest, the shortest, the last): same rows, and equal to the CPU oracle
    pick = sorted({max(range(S), key=lambda i: seq_lens[i]), min(range(S), key=lambda i: seq_lens[i]), S - 1})
    idx = torch.tensor(pick, device=DEV)
    sub_args = (q[idx].contiguous(), kc, vc, Hkv, scale, bt[idx].contiguous(), sl[idx].contiguous(), bs, None, kvd, ks, vs)
    sub, _ = _run_ours("v1", *sub_args)
    torch.testing.assert_close(sub.float(), got[idx].float(), rtol=RTOL, atol=ATOL)
    # the oracle gathers through the block table: hand it only the blocks these sequences own, renumbered
    own = bt[idx].reshape(-1).long()
    kc_s, vc_s = kc[own].cpu(), vc[own].cpu()
    bt_s = torch.arange(own.numel(), dtype=torch.int32).view(len(pick), blocks_per_seq)
    want = oattn.paged_attention(q[idx].cpu(), kc_s, vc_s, Hkv, scale, bt_s, sl[idx].cpu(), None, kvd, ksc, vsc)
    torch.testing.assert_close(sub.cpu().float(), want.float(), rtol=RTOL, atol=ATOL)
