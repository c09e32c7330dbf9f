"""GPU parity of the CUTLASS-style fused-MoE API through the C ABI: get_cutlass_moe_mm_data,
get_cutlass_pplx_moe_mm_data, shuffle_rows, cutlass_moe_mm, run_cutlass_moe_fp8 / cutlass_moe_fp8, against
oracle/moe.py and -- when oracle/_ref/libvllm_ref_cutlass.so was built -- the reference's own sm100 kernels.
Modelled on /root/reference/tests/kernels/moe/test_cutlass_moe.py:226-330 and
tests/kernels/quantization/test_cutlass_scaled_mm.py:514-645 (grouped GEMM, rtol 1e-2 / atol 5e-4 there).
Bars: metadata bit-exact (offsets, problem sizes, per-expert sets; the permutation equal to the oracle's stable
order), GEMM outputs 1e-2 rel / 2e-2 abs."""
import pytest
import torch

from oracle import moe as omoe
from oracle import quant as oq
from parity_utils import load_ref_cutlass, seed_all

pytestmark = pytest.mark.gpu
DEV = "cuda"
FP8 = torch.float8_e4m3fn
RTOL, ATOL = 1e-2, 2e-2


def _fp8_weights(E, N, K, per_channel):
    w = torch.randn(E, N, K) / 10
    if per_channel:
        s = w.abs().amax(dim=2).clamp(min=1e-6) / 448.0
        q = (w / s[:, :, None]).clamp(-448, 448).to(FP8)
    else:
        s = w.abs().amax(dim=(1, 2)).clamp(min=1e-6) / 448.0
        q = (w / s[:, None, None]).clamp(-448, 448).to(FP8)
    return q, s.float()


def _ids(T, k, E, invalid_frac=0.0):
    ids = torch.stack([torch.randperm(E)[:k] for _ in range(T)]).to(torch.int32)
    if invalid_frac:
        ids[torch.rand(T, k) < invalid_frac] = -1
    return ids


def _run_data(ids, E, n, k, blockscale=False):
    from vllm_b200 import _custom_ops as ops
    numel = ids.numel()
    eo = torch.full((E + 1,), -7, dtype=torch.int32, device=DEV)
    ps1 = torch.full((E, 3), -7, dtype=torch.int32, device=DEV)
    ps2 = torch.full((E, 3), -7, dtype=torch.int32, device=DEV)
    a_map = torch.zeros(numel, dtype=torch.int32, device=DEV)
    c_map = torch.full((numel,), -7, dtype=torch.int32, device=DEV)
    bo = torch.full((E + 1,), -7, dtype=torch.int32, device=DEV) if blockscale else None
    ops.get_cutlass_moe_mm_data(ids.to(DEV), eo, ps1, ps2, a_map, c_map, E, n, k, bo)
    torch.cuda.synchronize()
    return eo.cpu(), ps1.cpu(), ps2.cpu(), a_map.cpu(), c_map.cpu(), (bo.cpu() if blockscale else None)


@pytest.mark.parametrize("T,k,E", [(1, 1, 1), (1, 2, 8), (4, 3, 5), (32, 2, 8), (33, 2, 8), (128, 2, 8), (257, 6, 64),
                                   (1500, 8, 256), (64, 1, 1024), (7, 2, 40)])
@pytest.mark.parametrize("invalid", [0.0, 0.3])
def test_moe_mm_data_matches_oracle(T, k, E, invalid):
    seed_all(T * 31 + E)
    ids = _ids(T, k, E, invalid)
    n, kk = 96, 160
    eo, ps1, ps2, a_map, c_map, _ = _run_data(ids, E, n, kk)
    w_eo, w_ps1, w_ps2, w_a, w_c = omoe.get_cutlass_moe_mm_data(ids, E, n, kk)
    assert torch.equal(eo, w_eo)
    assert torch.equal(ps1, w_ps1) and torch.equal(ps2, w_ps2)
    total = int(w_eo[-1])
    assert torch.equal(a_map[:total], w_a[:total])
    assert torch.equal(c_map, w_c)
    # size-independent properties: c_map inverts a_map on the routed rows; every row lands inside its expert's range
    flat = ids.flatten()
    valid = flat >= 0
    pos = c_map[valid].long()
    assert torch.equal(torch.sort(pos).values, torch.arange(total))
    assert torch.equal(a_map[pos].long(), torch.nonzero(valid).flatten() // k)
    e_of = flat[valid].long()
    assert bool(((pos >= eo[e_of]) & (pos < eo[e_of + 1])).all())
    assert bool((c_map[~valid] == total).all())


def test_moe_mm_data_docstring_example_and_blockscale():
    """The worked routing example of moe_align_block_size.py:52-66 (experts 1..4 of 5) through this op, plus the
    fp4 block-scale offsets (rows rounded up to 128 per expert, never swapped: moe_data.cu:61-79,121-122)."""
    ids = torch.tensor([[2, 3, 4], [1, 2, 4], [1, 3, 4], [1, 2, 3]], dtype=torch.int32)
    eo, ps1, ps2, a_map, c_map, _ = _run_data(ids, 5, 8, 16)
    assert eo.tolist() == [0, 0, 3, 6, 9, 12]
    assert ps1.tolist() == [[16, 0, 16]] + [[16, 3, 16]] * 4          # swapped: 12 routed rows <= 64
    assert ps2.tolist() == [[16, 0, 8]] + [[16, 3, 8]] * 4
    assert a_map.tolist() == [1, 2, 3, 0, 1, 3, 0, 2, 3, 0, 1, 2]
    assert c_map.tolist() == [3, 6, 9, 0, 4, 10, 1, 7, 11, 2, 5, 8]
    eo, ps1, ps2, _, _, bo = _run_data(ids, 5, 8, 16, blockscale=True)
    assert ps1.tolist() == [[0, 16, 16]] + [[3, 16, 16]] * 4
    assert bo.tolist() == [0, 0, 128, 256, 384, 512]


def test_moe_mm_data_vs_reference_kernel():
    """Against the reference's own moe_data.cu kernels: offsets and problem sizes bit-exact; its permutation is
    atomics-ordered, so compare the per-expert SETS of source tokens and the c_map/a_map consistency."""
    ref = load_ref_cutlass()
    if ref is None:
        pytest.skip("oracle/_ref/libvllm_ref_cutlass.so not built")
    for T, k, E in [(16, 2, 8), (200, 2, 8), (513, 4, 32)]:
        seed_all(T)
        ids = _ids(T, k, E)
        eo, ps1, ps2, a_map, c_map, _ = _run_data(ids, E, 128, 256)
        r = [torch.zeros_like(x, device=DEV) for x in (eo, ps1, ps2, a_map, c_map)]
        ref.get_cutlass_moe_mm_data(ids.to(DEV), *r, E, 128, 256, None)
        torch.cuda.synchronize()
        reo, rps1, rps2, ra, rc = [x.cpu() for x in r]
        assert torch.equal(eo, reo) and torch.equal(ps1, rps1) and torch.equal(ps2, rps2)
        for e in range(E):
            lo, hi = int(eo[e]), int(eo[e + 1])
            assert torch.equal(torch.sort(a_map[lo:hi]).values, torch.sort(ra[lo:hi]).values)
        assert torch.equal(ra[rc.long()].long(), torch.arange(T * k) // k)


def test_pplx_data():
    from vllm_b200 import _custom_ops as ops
    E, padded_m, n, k = 6, 48, 64, 128
    cnt = torch.tensor([0, 5, 48, 1, 17, 33], dtype=torch.int32, device=DEV)
    eo = torch.empty(E, dtype=torch.int32, device=DEV)
    ps1 = torch.empty(E, 3, dtype=torch.int32, device=DEV)
    ps2 = torch.empty(E, 3, dtype=torch.int32, device=DEV)
    ops.get_cutlass_pplx_moe_mm_data(eo, ps1, ps2, cnt, E, padded_m, n, k)
    assert eo.tolist() == [e * padded_m for e in range(E)]
    assert ps1.tolist() == [[int(c), 2 * n, k] for c in cnt.tolist()]
    assert ps2.tolist() == [[int(c), k, n] for c in cnt.tolist()]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float8_e4m3fn, torch.float32, torch.int32])
@pytest.mark.parametrize("rows,cols,dst", [(1, 16, 1), (33, 4096, 66), (128, 24, 256), (7, 1, 20), (64, 14336, 128)])
def test_shuffle_rows(dtype, rows, cols, dst):
    """shuffle_rows == input[dst2src_map] bit for bit (vllm/_custom_ops.py:912-922), incl. expansion (dst > rows)
    and rows whose byte length is not a multiple of 16."""
    from vllm_b200 import _custom_ops as ops
    seed_all(rows + cols)
    x = torch.randn(rows, cols) * 4 if dtype != torch.int32 else torch.randint(-99, 99, (rows, cols)).float()
    x = x.to(dtype).to(DEV)
    m = torch.randint(0, rows, (dst,), dtype=torch.int32, device=DEV)
    got = ops.shuffle_rows(x, m)
    want = x.view(torch.uint8).view(rows, -1)[m.long()]
    assert got.shape == (dst, cols) and got.dtype == dtype
    assert torch.equal(got.view(torch.uint8).view(dst, -1), want)


def _grouped_problem(T, k, E, N, K, per_act_token, per_out_ch, skew=False):
    """Permuted fp8 activations + per-expert metadata, built with the oracle (CPU)."""
    wq, ws = _fp8_weights(E, N, K, per_out_ch)
    a = torch.randn(T, K).bfloat16()
    if per_act_token:
        aq, a_s = oq.dynamic_per_token_scaled_fp8_quant(a)
    else:
        aq, a_s = oq.dynamic_scaled_fp8_quant(a)
        a_s = a_s.reshape(1)
    ids = _ids(T, k, E)
    if skew:   # most rows on expert 0: several M-blocks for one expert, none for others
        ids[: T * 3 // 4, 0] = 0
        ids[: T * 3 // 4, 1:] = torch.arange(1, k).to(torch.int32) if k > 1 else ids[: T * 3 // 4, 1:]
    eo, ps1, _, a_map, _ = omoe.get_cutlass_moe_mm_data(ids, E, N // 2, K)
    ap = aq[a_map.long()]
    asp = a_s.reshape(-1)[a_map.long()].contiguous() if per_act_token else a_s
    return wq, ws, ap, asp, eo, ps1


@pytest.mark.parametrize("T,k,E,N,K", [
    (1, 2, 8, 256, 128), (16, 2, 8, 512, 256), (33, 2, 8, 512, 256), (128, 2, 8, 1024, 512), (128, 2, 8, 384, 4096),
    (64, 6, 64, 256, 512), (300, 2, 4, 256, 256), (700, 2, 2, 128, 128), (17, 1, 3, 144, 272), (2048, 2, 8, 256, 384),
    (2048, 2, 8, 128, 14336)])   # long K (> 8192) with > 128 rows per expert: the block_m clamp of the ragged launcher
@pytest.mark.parametrize("per_act_token,per_out_ch", [(False, False), (True, True), (True, False)])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float16])
def test_cutlass_moe_mm(T, k, E, N, K, per_act_token, per_out_ch, out_dtype):
    from vllm_b200 import _custom_ops as ops
    seed_all(T + N + K)
    wq, ws, ap, asp, eo, ps1 = _grouped_problem(T, k, E, N, K, per_act_token, per_out_ch, skew=(T == 300))
    counts = eo[1:] - eo[:-1]
    ps = ps1.clone()
    ps[:, 1 if T * k > 64 else 0] = N   # the oracle's n argument is N/2 (gate_up convention): store N itself
    want = omoe.cutlass_moe_mm(ap, wq, asp, ws, eo[:E], counts, out_dtype)
    out = torch.full((T * k, N), float("nan"), dtype=out_dtype, device=DEV)
    strides = torch.full((E,), K, dtype=torch.int64, device=DEV)
    c_strides = torch.full((E,), N, dtype=torch.int64, device=DEV)
    ops.cutlass_moe_mm(out, ap.to(DEV), wq.to(DEV), asp.to(DEV), ws.to(DEV), eo[:E].contiguous().to(DEV), ps.to(DEV),
                       strides, strides, c_strides, per_act_token, per_out_ch)
    torch.cuda.synchronize()
    got = out.cpu().float()
    assert torch.isfinite(got).all(), "every routed row must be written"
    torch.testing.assert_close(got, want.float(), rtol=RTOL, atol=ATOL)
    assert (got - want.float()).abs().max() <= 2.0 ** -7 * want.float().abs().max() + 1e-6


def test_cutlass_moe_mm_leaves_unowned_rows_and_graph():
    """Rows past the last expert's range (expert-parallel ranks) are not touched; the op is CUDA-graph capturable
    (tests/kernels/moe/test_cutlass_moe.py:93-107 runs it under a graph)."""
    from vllm_b200 import _custom_ops as ops
    seed_all(5)
    T, k, E, N, K = 40, 2, 4, 256, 256
    wq, ws, ap, asp, eo, ps1 = _grouped_problem(T, k, E, N, K, True, True)
    ps = ps1.clone()
    ps[:, 1] = N
    # drop the last expert: its rows become unowned
    ps[E - 1, 0] = 0
    counts = (eo[1:] - eo[:-1]).clone()
    counts[E - 1] = 0
    sentinel = torch.full((T * k, N), 3.0, dtype=torch.bfloat16)
    want = omoe.cutlass_moe_mm(ap, wq, asp, ws, eo[:E], counts, torch.bfloat16, out=sentinel)
    out = sentinel.to(DEV)
    args = (out, ap.to(DEV), wq.to(DEV), asp.to(DEV), ws.to(DEV), eo[:E].contiguous().to(DEV), ps.to(DEV),
            torch.full((E,), K, dtype=torch.int64, device=DEV), torch.full((E,), K, dtype=torch.int64, device=DEV),
            torch.full((E,), N, dtype=torch.int64, device=DEV), True, True)
    ops.cutlass_moe_mm(*args)
    torch.cuda.synchronize()
    torch.testing.assert_close(out.cpu().float(), want.float(), rtol=RTOL, atol=ATOL)
    assert bool((out.cpu()[int(eo[E - 1]):] == 3.0).all())
    g = torch.cuda.CUDAGraph()
    out.fill_(3.0)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        ops.cutlass_moe_mm(*args)
        with torch.cuda.graph(g, stream=s):
            ops.cutlass_moe_mm(*args)
    out.fill_(3.0)
    g.replay()
    torch.cuda.synchronize()
    torch.testing.assert_close(out.cpu().float(), want.float(), rtol=RTOL, atol=ATOL)


def test_cutlass_moe_mm_errors():
    from vllm_b200 import _custom_ops as ops
    E, N, K = 2, 128, 128
    a = torch.zeros(4, K, device=DEV).to(FP8)
    b = torch.zeros(E, N, K, device=DEV).to(FP8)
    out = torch.zeros(4, N, dtype=torch.bfloat16, device=DEV)
    sc = torch.ones(1, device=DEV)
    bs = torch.ones(E, device=DEV)
    eo = torch.zeros(E, dtype=torch.int32, device=DEV)
    ps = torch.zeros(E, 3, dtype=torch.int32, device=DEV)
    st = torch.full((E,), K, dtype=torch.int64, device=DEV)
    with pytest.raises(RuntimeError, match="float8_e4m3fn"):
        ops.cutlass_moe_mm(out, a.to(torch.bfloat16), b, sc, bs, eo, ps, st, st, st, False, False)
    with pytest.raises(RuntimeError, match="Invalid output type"):
        ops.cutlass_moe_mm(out.float(), a, b, sc, bs, eo, ps, st, st, st, False, False)
    with pytest.raises(RuntimeError, match="No input A"):
        ops.cutlass_moe_mm(out[:0], a[:0], b, sc, bs, eo, ps, st, st, st, False, False)
    assert ops.cutlass_group_gemm_supported(100) and not ops.cutlass_group_gemm_supported(90)


@pytest.mark.parametrize("T,k,E,N,K", [(64, 2, 8, 512, 256), (200, 2, 8, 1024, 512), (16, 2, 8, 256, 128)])
@pytest.mark.parametrize("per_act_token,per_out_ch", [(False, False), (True, True)])
def test_cutlass_moe_mm_vs_reference_kernel(T, k, E, N, K, per_act_token, per_out_ch):
    """Against the reference's own sm100 CUTLASS grouped GEMM (moe/grouped_mm_c3x_sm100.cu) on the same inputs.
    Both kernels get the problem sizes in the order get_cutlass_moe_mm_data produces for that row count."""
    ref = load_ref_cutlass()
    if ref is None:
        pytest.skip("oracle/_ref/libvllm_ref_cutlass.so not built")
    from vllm_b200 import _custom_ops as ops
    seed_all(T + N)
    wq, ws, ap, asp, eo, ps1 = _grouped_problem(T, k, E, N, K, per_act_token, per_out_ch)
    counts = eo[1:] - eo[:-1]
    nn, kk = torch.full((E,), N, dtype=torch.int32), torch.full((E,), K, dtype=torch.int32)
    # <= 64 rows: the reference dispatches its swap-AB config (grouped_mm_c3x_sm100.cu:30-44,91-95), which reads the
    # problem sizes as (N, m_e, K) -- the order get_cutlass_moe_mm_data writes for such inputs (moe_data.cu:36-43)
    ps = torch.stack([nn, counts.to(torch.int32), kk] if T * k <= 64 else [counts.to(torch.int32), nn, kk],
                     1).contiguous()
    st = torch.full((E,), K, dtype=torch.int64, device=DEV)
    cst = torch.full((E,), N, dtype=torch.int64, device=DEV)
    ws_ref = ws.reshape(E, -1).contiguous()
    asp_ref = asp.reshape(-1, 1).contiguous() if per_act_token else asp
    outs = []
    for fn in (ops.cutlass_moe_mm, ref.cutlass_moe_mm):
        out = torch.zeros(T * k, N, dtype=torch.bfloat16, device=DEV)
        fn(out, ap.to(DEV), wq.to(DEV), asp_ref.to(DEV), ws_ref.to(DEV), eo[:E].contiguous().to(DEV), ps.to(DEV),
           st, st, cst, per_act_token, per_out_ch)
        torch.cuda.synchronize()
        outs.append(out.cpu().float())
    torch.testing.assert_close(outs[0], outs[1], rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("T,k,E,inter,hidden", [(1, 2, 8, 256, 128), (33, 2, 8, 256, 256), (128, 2, 8, 512, 512),
                                                (64, 6, 64, 128, 256), (222, 2, 8, 1024, 1024)])
@pytest.mark.parametrize("per_act_token,per_out_ch", [(False, False), (True, True)])
@pytest.mark.parametrize("use_expert_map", [False, True])
def test_cutlass_moe_fp8_end_to_end(T, k, E, inter, hidden, per_act_token, per_out_ch, use_expert_map):
    """cutlass_moe_fp8 (cutlass_moe.py:324-409) vs the oracle pipeline; with an expert_map this rank owns the first
    half of the experts and the other ids contribute zeros (test_cutlass_moe.py `test_run_cutlass_moe_fp8` /
    `ep` cases)."""
    from vllm_b200.fused_moe import cutlass_moe_fp8
    seed_all(T + inter)
    w1q, w1s = _fp8_weights(E, 2 * inter, hidden, per_out_ch)
    w2q, w2s = _fp8_weights(E, hidden, inter, per_out_ch)
    a = (torch.randn(T, hidden) / 2).bfloat16()
    ids = _ids(T, k, E)
    tw = torch.softmax(torch.randn(T, k), -1)
    emap = None
    if use_expert_map:
        if E < 2:
            pytest.skip("needs >= 2 experts")
        local_E = E // 2
        emap = torch.full((E,), -1, dtype=torch.int32)
        emap[:local_E] = torch.arange(local_E, dtype=torch.int32)
        w1q, w1s, w2q, w2s = w1q[:local_E], w1s[:local_E], w2q[:local_E], w2s[:local_E]
    want = omoe.cutlass_moe_fp8(a, w1q, w2q, tw, ids, w1s, w2s, per_act_token, expert_map=emap)
    got = cutlass_moe_fp8(a.to(DEV), w1q.contiguous().to(DEV), w2q.contiguous().to(DEV), tw.to(DEV), ids.to(DEV),
                          w1s.contiguous().to(DEV), w2s.contiguous().to(DEV), per_act_token=per_act_token,
                          expert_map=None if emap is None else emap.to(DEV),
                          global_num_experts=E)
    torch.cuda.synchronize()
    assert got.shape == (T, hidden) and got.dtype == torch.bfloat16
    # two chained fp8 GEMMs: an intermediate that rounds to the neighbouring e4m3 value moves single outputs by one
    # fp8 step, so -- as tests/test_gpu_fused_moe.py and the reference's test_moe.py do -- the bound is 2e-2 scaled
    # by the output magnitude, plus a tight bound on the mean error
    g, w = got.cpu().float(), want.float()
    scale = max(w.abs().max().item(), 1.0)
    torch.testing.assert_close(g, w, rtol=2e-2, atol=2e-2 * scale)
    assert (g - w).abs().mean().item() <= 3e-3 * scale


def test_cutlass_moe_fp8_matches_fused_experts():
    """The two MoE entry points of the package (moe_align-driven fused_experts and the CUTLASS-style pipeline) agree
    on the same inputs (test_cutlass_moe.py:226-270 compares cutlass_moe_fp8 with fused_experts the same way)."""
    from vllm_b200.fused_moe import cutlass_moe_fp8, fused_experts
    seed_all(11)
    T, k, E, inter, hidden = 96, 2, 8, 512, 512
    w1q, w1s = _fp8_weights(E, 2 * inter, hidden, True)
    w2q, w2s = _fp8_weights(E, hidden, inter, True)
    a = (torch.randn(T, hidden) / 2).bfloat16().to(DEV)
    ids = _ids(T, k, E).to(DEV)
    tw = torch.softmax(torch.randn(T, k), -1).to(DEV)
    w1q, w2q, w1s, w2s = w1q.to(DEV), w2q.to(DEV), w1s.to(DEV), w2s.to(DEV)
    x = cutlass_moe_fp8(a, w1q, w2q, tw, ids, w1s, w2s, per_act_token=True)
    y = fused_experts(a, w1q, w2q, tw, ids, use_fp8_w8a8=True, per_channel_quant=True, w1_scale=w1s, w2_scale=w2s)
    scale = max(y.float().abs().max().item(), 1.0)
    torch.testing.assert_close(x.float(), y.float(), rtol=2e-2, atol=2e-2 * scale)
    assert (x.float() - y.float()).abs().mean().item() <= 3e-3 * scale


# ---- the fused ends of the permuted pipeline (one launch each) vs the stand-alone ops they replace --------------
@pytest.mark.parametrize("T,k,E,hidden", [(1, 2, 8, 128), (33, 2, 8, 1024), (128, 2, 8, 4096), (64, 6, 64, 256),
                                          (200, 4, 16, 2048)])
@pytest.mark.parametrize("mode", ["per_token", "static"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("invalid", [0.0, 0.3])
def test_gather_quant_matches_quant_then_shuffle(T, k, E, hidden, mode, dtype, invalid):
    """moe_gather_quant_fp8 == scaled_fp8_quant followed by shuffle_rows of codes and scales (cutlass_moe.py:142-150),
    bit for bit, on every routed row; rows past the routed count are not written."""
    from vllm_b200 import _custom_ops as ops
    seed_all(T * 7 + hidden)
    a = (torch.randn(T, hidden) * 2).to(dtype).to(DEV)
    ids = _ids(T, k, E, invalid).to(DEV)
    rows = T * k
    eo = torch.empty(E + 1, dtype=torch.int32, device=DEV)
    ps = torch.empty(E, 3, dtype=torch.int32, device=DEV)
    a_map = torch.zeros(rows, dtype=torch.int32, device=DEV)
    c_map = torch.empty(rows, dtype=torch.int32, device=DEV)
    ops.get_cutlass_moe_mm_data(ids, eo, ps, ps.clone(), a_map, c_map, E, 64, hidden)
    static = torch.tensor([0.03], device=DEV) if mode == "static" else None
    q, s = ops.scaled_fp8_quant(a, static, use_per_token_if_dynamic=mode == "per_token")
    want_q = ops.shuffle_rows(q, a_map)
    got_q, got_s = ops.moe_gather_quant_fp8(a, a_map, eo[-1:], scale=static)
    torch.cuda.synchronize()
    n = int(eo[-1])
    assert n == int((ids >= 0).sum())
    assert torch.equal(got_q[:n].view(torch.uint8), want_q[:n].view(torch.uint8))
    if mode == "per_token":
        want_s = ops.shuffle_rows(s.view(T, 1), a_map)
        assert torch.equal(got_s[:n], want_s[:n])
    else:
        assert got_s.data_ptr() == static.data_ptr()
    # without the bound every row is written (a_map is zero-filled past the routed rows)
    got_q2, _ = ops.moe_gather_quant_fp8(a, a_map, None, scale=static)
    torch.cuda.synchronize()
    assert torch.equal(got_q2.view(torch.uint8), want_q.view(torch.uint8))


@pytest.mark.parametrize("T,k,hidden", [(1, 1, 128), (33, 2, 1024), (128, 2, 4096), (64, 3, 256), (50, 4, 512),
                                        (40, 6, 256)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("weighted", [True, False])
@pytest.mark.parametrize("invalid", [0.0, 0.3])
def test_unpermute_reduce_matches_shuffle_mul_sum(T, k, hidden, dtype, weighted, invalid):
    """moe_unpermute_reduce == shuffle_rows(c3, c_map) -> mul_(topk_weights) -> moe_sum (cutlass_moe.py:205-214,
    topk_weight_and_reduce.py:60-100), bit for bit; (token, j) pairs of foreign experts contribute zero."""
    from vllm_b200 import _custom_ops as ops
    seed_all(T + k + hidden)
    E = 8
    ids = _ids(T, k, E, invalid).to(DEV)
    rows = T * k
    eo = torch.empty(E + 1, dtype=torch.int32, device=DEV)
    ps = torch.empty(E, 3, dtype=torch.int32, device=DEV)
    a_map = torch.zeros(rows, dtype=torch.int32, device=DEV)
    c_map = torch.empty(rows, dtype=torch.int32, device=DEV)
    ops.get_cutlass_moe_mm_data(ids, eo, ps, ps.clone(), a_map, c_map, E, 64, hidden)
    torch.cuda.synchronize()
    n = int(eo[-1])
    c3 = torch.randn(rows, hidden, device=DEV).to(dtype)
    c3_ref = c3.clone()
    c3_ref[n:] = 0            # the reference zero-fills c3 so that row `n` reads as zero (cutlass_moe.py:196-197)
    c3[n:] = float("nan")     # the fused kernel must never read those rows
    w = torch.softmax(torch.randn(T, k, device=DEV), -1)
    fused = ops.shuffle_rows(c3_ref, c_map.clamp(max=rows - 1)).view(T, k, hidden)
    if weighted:
        fused.mul_(w.view(T, k, 1))
    want = torch.empty(T, hidden, device=DEV, dtype=dtype)
    ops.moe_sum(fused, want)
    got = ops.moe_unpermute_reduce(c3, c_map, w if weighted else None, eo[-1:], T, k)
    torch.cuda.synchronize()
    assert torch.equal(got.view(torch.int16), want.view(torch.int16))


@pytest.mark.parametrize("per_act_token,per_out_ch", [(False, False), (True, True)])
@pytest.mark.parametrize("use_expert_map", [False, True])
def test_run_cutlass_moe_fp8_reference_entry(per_act_token, per_out_ch, use_expert_map):
    """run_cutlass_moe_fp8 called the way the reference's CutlassExpertsFp8.apply does (cutlass_moe.py:287-321):
    pre-quantised activations, caller-owned workspaces, un-weighted [M*topk, K] output in (token, j) order; its
    weighted top-k sum must equal cutlass_moe_fp8 on the same inputs."""
    from vllm_b200 import _custom_ops as ops
    from vllm_b200.fused_moe import cutlass_moe_fp8, run_cutlass_moe_fp8
    seed_all(5)
    T, k, E, inter, hidden = 96, 2, 8, 512, 256
    w1q, w1s = _fp8_weights(E, 2 * inter, hidden, per_out_ch)
    w2q, w2s = _fp8_weights(E, hidden, inter, per_out_ch)
    a = (torch.randn(T, hidden) / 2).bfloat16().to(DEV)
    ids = _ids(T, k, E).to(DEV)
    tw = torch.softmax(torch.randn(T, k), -1).to(DEV)
    emap = None
    if use_expert_map:
        emap = torch.full((E,), -1, dtype=torch.int32)
        emap[:E // 2] = torch.arange(E // 2, dtype=torch.int32)
        emap = emap.to(DEV)
        w1q, w1s, w2q, w2s = w1q[:E // 2], w1s[:E // 2], w2q[:E // 2], w2s[:E // 2]
    w1q, w2q, w1s, w2s = (t.contiguous().to(DEV) for t in (w1q, w2q, w1s, w2s))
    a1q, a1s = ops.scaled_fp8_quant(a, None, use_per_token_if_dynamic=per_act_token)
    ws13 = torch.empty(T * k * max(2 * inter, hidden), dtype=a.dtype, device=DEV)
    ws2 = torch.empty(T * k * inter, dtype=a.dtype, device=DEV)
    out = torch.full((T * k, hidden), float("nan"), dtype=a.dtype, device=DEV)

    def act(o, i):
        ops.silu_and_mul(o, i)
    run_cutlass_moe_fp8(out, a1q, w1q, w2q, ids, act, E, emap, w1s, w2s, a1s, None, ws13, ws2, None, a.dtype,
                        per_act_token, per_out_ch, False)
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all()
    if use_expert_map:   # (token, j) pairs of foreign experts read as zero rows
        foreign = (emap[ids.long()] < 0).view(-1)
        assert (out[foreign] == 0).all() and foreign.any()
    x = (out.view(T, k, hidden).float() * tw.view(T, k, 1)).to(a.dtype)
    got = torch.empty(T, hidden, dtype=a.dtype, device=DEV)
    ops.moe_sum(x, got)
    want = cutlass_moe_fp8(a, w1q, w2q, tw, ids, w1s, w2s, per_act_token=per_act_token, expert_map=emap,
                           global_num_experts=E)
    torch.cuda.synchronize()
    scale = max(want.float().abs().max().item(), 1.0)
    # the same ops in the same order on both routes
    torch.testing.assert_close(got.float(), want.float(), rtol=0, atol=1e-3 * scale)


def test_run_cutlass_moe_fp8_batched_format():
    """Batched-experts format (cutlass_moe.py:105-126): hidden_states [E, padded_M, K], expert_num_tokens rows valid
    per expert; valid rows equal the per-expert dense computation of the oracle."""
    from vllm_b200 import _custom_ops as ops
    from vllm_b200.fused_moe import run_cutlass_moe_fp8
    seed_all(9)
    E, padded_M, inter, hidden = 4, 32, 256, 128
    counts = torch.tensor([32, 0, 7, 19], dtype=torch.int32)
    w1q, w1s = _fp8_weights(E, 2 * inter, hidden, True)
    w2q, w2s = _fp8_weights(E, hidden, inter, True)
    a = (torch.randn(E, padded_M, hidden) / 2).bfloat16()
    a1q, a1s = oq.dynamic_per_token_scaled_fp8_quant(a.view(-1, hidden))
    out = torch.zeros(E, padded_M, hidden, dtype=torch.bfloat16, device=DEV)
    ws13 = torch.empty(E * padded_M * max(2 * inter, hidden), dtype=torch.bfloat16, device=DEV)
    ws2 = torch.empty(E * padded_M * inter, dtype=torch.bfloat16, device=DEV)
    run_cutlass_moe_fp8(out, a1q.view(E, padded_M, hidden).to(DEV), w1q.to(DEV), w2q.to(DEV), None,
                        lambda o, i: ops.silu_and_mul(o, i), E, None, w1s.view(E, -1, 1).to(DEV),
                        w2s.view(E, -1, 1).to(DEV), a1s.view(E, padded_M, 1).to(DEV), None, ws13, ws2, counts.to(DEV),
                        torch.bfloat16, True, True, True)
    torch.cuda.synchronize()
    offs = torch.arange(E, dtype=torch.int32) * padded_M
    c1 = omoe.cutlass_moe_mm(a1q, w1q, a1s.view(-1), w1s, offs, counts, torch.bfloat16)
    c2 = omoe.silu_and_mul(c1)
    a2q, a2s = oq.dynamic_per_token_scaled_fp8_quant(c2)
    c3 = omoe.cutlass_moe_mm(a2q, w2q, a2s.view(-1), w2s, offs, counts, torch.bfloat16).view(E, padded_M, hidden)
    for e in range(E):
        m = int(counts[e])
        if m:
            g, w = out[e, :m].cpu().float(), c3[e, :m].float()
            torch.testing.assert_close(g, w, rtol=2e-2, atol=2e-2 * max(w.abs().max().item(), 1.0))
