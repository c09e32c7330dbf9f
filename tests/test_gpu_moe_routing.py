"""GPU parity: topk_softmax, moe_align_block_size, moe_sum through the C ABI vs the oracle and the reference's own
CUDA kernels.  Modelled on /root/reference/tests/kernels/moe/test_moe_align_block_size.py and test_moe.py."""
import pytest
import torch

from oracle import moe as omoe
from parity_utils import load_ref_cuda, seed_all

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("T,E,k", [(1, 8, 2), (128, 8, 2), (33, 64, 6), (257, 256, 8), (64, 60, 4), (16, 128, 1), (5, 3, 3)])
@pytest.mark.parametrize("idx_dtype", [torch.int32, torch.int64, torch.uint32])
def test_topk_softmax(T, E, k, idx_dtype):
    from vllm_b200 import _custom_ops as ops
    seed_all(T + E)
    g = torch.randn(T, E)
    g[0, : min(E, 4)] = 1.5  # exact ties -> lowest index first
    w = torch.empty(T, k, device=DEV)
    ids = torch.empty(T, k, dtype=idx_dtype, device=DEV)
    tei = torch.empty(T, k, dtype=torch.int32, device=DEV)
    ops.topk_softmax(w, ids, tei, g.to(DEV))
    wo, io, to = omoe.topk_softmax(g, k)
    assert torch.equal(ids.cpu().to(torch.int64), io.to(torch.int64))   # bit-exact ids
    assert torch.equal(tei.cpu(), to)
    torch.testing.assert_close(w.cpu(), wo, rtol=1e-5, atol=1e-7)
    ref = load_ref_cuda()
    if ref is not None and idx_dtype != torch.uint32:
        wr, ir, tr = torch.empty_like(w), torch.empty_like(ids), torch.empty_like(tei)
        ref.topk_softmax(wr, ir, tr, g.to(DEV))
        assert torch.equal(ir, ids) and torch.equal(tr, tei)
        torch.testing.assert_close(w, wr, rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("T,k,E,block", [(1, 2, 8, 16), (128, 2, 8, 128), (128, 2, 8, 64), (1000, 6, 64, 16), (4096, 8, 256, 128),
                                         (77, 3, 5, 4), (3, 1, 512, 32), (2048, 2, 1024, 16)])
def test_moe_align_block_size(T, k, E, block):
    from vllm_b200 import _custom_ops as ops
    seed_all(T)
    ids = torch.stack([torch.randperm(E)[:k] for _ in range(T)]).to(torch.int32)
    max_padded = T * k + E * (block - 1)
    s = torch.empty(max_padded, dtype=torch.int32, device=DEV)
    e = torch.empty((max_padded + block - 1) // block, dtype=torch.int32, device=DEV)
    n = torch.empty(1, dtype=torch.int32, device=DEV)
    ops.moe_align_block_size(ids.to(DEV), E, block, s, e, n)
    so, eo, no = omoe.moe_align_block_size(ids, block, E)
    assert int(n) == int(no)
    assert torch.equal(s.cpu(), so)          # stable order == the reference test's golden
    nblk = int(no) // block
    assert torch.equal(e.cpu()[:nblk], eo[:nblk]) and (e.cpu()[nblk:] == 0).all()
    ref = load_ref_cuda()
    if ref is not None and E < 1024:
        sr = torch.full_like(s, T * k)
        er, nr = torch.zeros_like(e), torch.zeros_like(n)
        ref.moe_align_block_size(ids.to(DEV), E, block, sr, er, nr)
        assert int(nr) == int(n) and torch.equal(er[:nblk], e[:nblk])
        # the reference's order inside an expert is atomic-order: compare as per-block sets
        a = s[: int(n)].reshape(-1, block).sort(-1).values
        b = sr[: int(n)].reshape(-1, block).sort(-1).values
        # blocks of one expert may be permuted among themselves only if the expert spans several blocks; compare per expert
        for ex in range(E):
            m = e[:nblk] == ex
            if m.any():
                assert torch.equal(a[m].flatten().sort().values, b[m].flatten().sort().values)


def test_moe_align_docstring_known_answer():
    from vllm_b200 import _custom_ops as ops
    ids = torch.tensor([[2, 3, 4], [1, 2, 4], [1, 3, 4], [1, 2, 3]], dtype=torch.int32, device=DEV)
    s = torch.empty(12 + 5 * 3, dtype=torch.int32, device=DEV)
    e = torch.empty(7, dtype=torch.int32, device=DEV)
    n = torch.empty(1, dtype=torch.int32, device=DEV)
    ops.moe_align_block_size(ids, 5, 4, s, e, n)
    assert s[:16].tolist() == [3, 6, 9, 12, 0, 4, 10, 12, 1, 7, 11, 12, 2, 5, 8, 12] and int(n) == 16


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("T,k,H", [(1, 2, 64), (128, 2, 4096), (33, 3, 1000), (17, 4, 128), (9, 6, 256)])
def test_moe_sum(dtype, T, k, H):
    from vllm_b200 import _custom_ops as ops
    seed_all(9)
    x = torch.randn(T, k, H).to(dtype)
    out = torch.empty(T, H, dtype=dtype, device=DEV)
    ops.moe_sum(x.to(DEV), out)
    want = omoe.moe_sum(x)
    if k in (2, 3, 4):
        assert torch.equal(out.cpu(), want)  # same adds in the same order and dtype
    else:
        torch.testing.assert_close(out.cpu().float(), want.float(), rtol=1e-2, atol=1e-2)
    ref = load_ref_cuda()
    if ref is not None and k in (2, 3, 4):
        outr = torch.empty_like(out)
        ref.moe_sum(x.to(DEV), outr)
        assert torch.equal(out, outr)


@pytest.mark.parametrize("T,E,k", [(1, 8, 2), (128, 8, 2), (33, 64, 6), (257, 16, 4), (64, 160, 8)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("renorm", [False, True])
def test_fused_topk_one_launch(T, E, k, dtype, renorm):
    """fused_topk (fused_moe.py:882-915): the single-launch path (logits read in their own dtype, renormalisation in the
    kernel) gives the ids of the explicit arg-max oracle bit for bit and the weights of `softmax -> topk -> w / sum(w)`
    to fp32 rounding."""
    from vllm_b200.fused_moe import fused_topk
    seed_all(T + E + k)
    gate = torch.randn(T, E).to(dtype)
    hidden = torch.zeros(T, 16, dtype=torch.bfloat16, device=DEV)
    w_o, ids_o, _ = omoe.topk_softmax(gate.float(), k)
    if renorm:
        w_o = w_o / w_o.sum(-1, keepdim=True)
    w, ids, tei = fused_topk(hidden, gate.to(DEV), k, renorm)
    torch.cuda.synchronize()
    assert ids.dtype == torch.int32 and torch.equal(ids.cpu(), ids_o.to(torch.int32))
    torch.testing.assert_close(w.cpu(), w_o, rtol=2e-6, atol=1e-7)
    assert torch.equal(tei.cpu(), (torch.arange(k)[None, :] * T + torch.arange(T)[:, None]).to(torch.int32))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("renorm", [True, False])
@pytest.mark.parametrize("T,H,E,k", [(1, 4096, 8, 2), (128, 4096, 8, 2), (33, 1024, 32, 6), (256, 8192, 16, 1), (7, 136, 3, 3)])
def test_router_topk_one_launch(T, H, E, k, renorm, dtype):
    """vb200_moe_router_topk = gate GEMM + fused_topk (+ abs-max of the hidden states).  The logits are a floating-point
    GEMM: against a plain PyTorch fp32 reference of hidden @ w_gate^T within one unit in the last place of the
    activation dtype; the routing on THOSE logits is integer / exact work: bit-identical to fused_topk
    (fused_moe.py:882-915) run on them; the abs-max scale equals scaled_fp8_quant's dynamic per-tensor scale."""
    from vllm_b200 import _custom_ops as ops
    from vllm_b200.fused_moe import fused_topk, router_topk
    seed_all(T + H + E)
    x = torch.randn(T, H, device=DEV).to(dtype)
    w = (torch.randn(E, H, device=DEV) / H ** 0.5 * 3).to(dtype)
    scale = torch.zeros(1, dtype=torch.float32, device=DEV)
    tw, ids, logits = router_topk(x, w, k, renorm, absmax_scale=scale, return_logits=True)
    torch.cuda.synchronize()
    want_logits = x.float() @ w.float().t()
    ulp = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    assert (logits.float() - want_logits).abs().max() <= ulp * want_logits.abs().clamp_min(1.0).max() * 1.01
    tw2, ids2, _ = fused_topk(x, logits, k, renorm)
    assert torch.equal(ids, ids2)
    assert torch.equal(tw, tw2)
    _, want_scale = ops.scaled_fp8_quant(x, None)
    assert torch.equal(scale.view(-1), want_scale.view(-1))
