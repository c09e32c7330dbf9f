"""GPU parity: cutlass_scaled_mm (fp8 w8a8, tcgen05 kernel) through the C ABI vs the oracle.
Tolerance: north_star 1e-2 rel / 2e-2 abs against the reference result computed in fp32 from the exactly
dequantised fp8 operands (baseline_scaled_mm, /root/reference/tests/kernels/utils.py:1212-1251); the reference's own
test uses rtol 5e-1 / atol 1.5e-1 (tests/kernels/quantization/test_cutlass_scaled_mm.py:67-102)."""
import pytest
import torch

from oracle import quant as oq
from parity_utils import seed_all

pytestmark = pytest.mark.gpu
DEV = "cuda"
FP8 = torch.float8_e4m3fn


def to_fp8(t):
    return t.clamp(-448, 448).to(FP8)


def _check(M, N, K, per_token, per_channel, out_dtype, use_bias, seed=0, scale=1.0):
    from vllm_b200 import _custom_ops as ops
    seed_all(seed)
    a = to_fp8(torch.randn(M, K) * scale)
    b = to_fp8(torch.randn(N, K) * scale).t()        # [K, N] column-major
    sa = (torch.rand(M if per_token else 1, 1) * 0.1 + 0.01)
    sb = (torch.rand(1, N if per_channel else 1) * 0.1 + 0.01)
    bias = (torch.rand(N) * 10).to(out_dtype) if use_bias else None
    got = ops.cutlass_scaled_mm(a.to(DEV), b.to(DEV), sa.to(DEV), sb.to(DEV), out_dtype,
                                None if bias is None else bias.to(DEV))
    torch.cuda.synchronize()
    want = oq.baseline_scaled_mm(a, b, sa, sb, torch.float32, None if bias is None else bias.float())
    # values are O(sqrt(K)) * scales: relative check with an absolute floor scaled to the output magnitude
    torch.testing.assert_close(got.cpu().float(), want, rtol=1e-2, atol=2e-2 * max(1.0, float(want.abs().mean())))
    # the kernel's own rounding order must agree with the oracle restatement to 1 ulp of the output type
    exact = oq.scaled_mm_epilogue_order(a, b, sa, sb, out_dtype, bias)
    ulp = 2.0 ** -7 if out_dtype == torch.bfloat16 else 2.0 ** -10
    rel = ((got.cpu().float() - exact.float()).abs() / exact.float().abs().clamp(min=1e-3)).max()
    assert rel <= 2 * ulp, f"max rel err vs exact-order oracle {rel}"


@pytest.mark.parametrize("M", [1, 16, 33, 64, 128, 129, 200, 256, 257, 512, 1000])
@pytest.mark.parametrize("N,K", [(128, 128), (256, 512), (4096, 4096), (6144, 4096), (1280, 8192), (4096, 14336)])
def test_scaled_mm_shapes(M, N, K):
    if M > 300 and N * K > 4096 * 4096:
        pytest.skip("oracle time")
    _check(M, N, K, True, True, torch.bfloat16, False, seed=M + N)


@pytest.mark.parametrize("per_token", [False, True])
@pytest.mark.parametrize("per_channel", [False, True])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("use_bias", [False, True])
def test_scaled_mm_scales_bias(per_token, per_channel, out_dtype, use_bias):
    _check(48, 272, 496, per_token, per_channel, out_dtype, use_bias, seed=5)  # N, K multiples of 16 only


def test_scaled_mm_full_size_linearity():
    """BASELINE cfg 3 gate_up shape (M=256, N=28672, K=4096): size-independent properties instead of the CPU
    oracle -- doubling a_scale doubles the output exactly (power of two), zero rows give zero, and a sampled set
    of outputs matches an fp64 dot product."""
    from vllm_b200 import _custom_ops as ops
    seed_all(1)
    M, N, K = 256, 28672, 4096
    a = to_fp8(torch.randn(M, K, device=DEV))
    a[7] = 0
    b = to_fp8(torch.randn(N, K, device=DEV)).t()
    sa = torch.rand(M, 1, device=DEV) * 0.1 + 0.01
    sb = torch.rand(1, N, device=DEV) * 0.1 + 0.01
    o1 = ops.cutlass_scaled_mm(a, b, sa, sb, torch.bfloat16)
    o2 = ops.cutlass_scaled_mm(a, b, sa * 2, sb, torch.bfloat16)
    assert torch.equal(o2, o1 * 2)
    assert (o1[7] == 0).all()
    rows = torch.randint(0, M, (64,), device=DEV)
    cols = torch.randint(0, N, (64,), device=DEV)
    ref = (a[rows].double() * b.t()[cols].double()).sum(-1) * sa[rows, 0].double() * sb[0, cols].double()
    torch.testing.assert_close(o1[rows, cols].double(), ref, rtol=1e-2, atol=2e-2)
    # run-to-run determinism (split-K sums in a fixed order)
    assert torch.equal(o1, ops.cutlass_scaled_mm(a, b, sa, sb, torch.bfloat16))


def _gpu_reference(a, b, sa, sb):
    """fp32 torch matmul of the exactly dequantised operands on the GPU (sizes the CPU oracle cannot finish)."""
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        return sa * (sb * (a.float() @ b.float()))
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


@pytest.mark.parametrize("M,N,K", [(2048, 16384, 512), (1000, 6144 + 16, 1024 + 16), (300, 28672, 256)])
def test_scaled_mm_persistent_many_tiles(M, N, K):
    """More tiles than CTA pairs: every pair walks several tiles, both TMEM accumulator stages and the
    shared-memory ring wrap their mbarrier phases several times; ragged M / N / K tails included."""
    from vllm_b200 import _custom_ops as ops
    seed_all(M + K)
    a = to_fp8(torch.randn(M, K, device=DEV))
    b = to_fp8(torch.randn(N, K, device=DEV)).t()
    sa = torch.rand(M, 1, device=DEV) * 0.1 + 0.01
    sb = torch.rand(1, N, device=DEV) * 0.1 + 0.01
    got = ops.cutlass_scaled_mm(a, b, sa, sb, torch.bfloat16)
    want = _gpu_reference(a, b, sa, sb)
    torch.testing.assert_close(got.float(), want, rtol=1e-2, atol=2e-2 * max(1.0, float(want.abs().mean())))


@pytest.mark.parametrize("M", [16, 100, 256])
@pytest.mark.parametrize("N,K", [(4096, 4096), (6144, 4096), (4096, 14336), (512, 8192)])
def test_scaled_mm_split_k_repeated(M, N, K):
    """Decode shapes that take the split-K path: repeated launches re-arm the arrival counters, results are
    bit-identical run to run and match the fp32 reference."""
    from vllm_b200 import _custom_ops as ops
    seed_all(N + M)
    a = to_fp8(torch.randn(M, K, device=DEV))
    b = to_fp8(torch.randn(N, K, device=DEV)).t()
    sa = torch.rand(M, 1, device=DEV) * 0.1 + 0.01
    sb = torch.rand(1, N, device=DEV) * 0.1 + 0.01
    want = _gpu_reference(a, b, sa, sb)
    first = None
    for _ in range(5):
        got = ops.cutlass_scaled_mm(a, b, sa, sb, torch.bfloat16)
        if first is None:
            first = got
            torch.testing.assert_close(got.float(), want, rtol=1e-2, atol=2e-2 * max(1.0, float(want.abs().mean())))
        else:
            assert torch.equal(first, got)


def test_scaled_mm_rejects_bad_inputs():
    from vllm_b200 import _custom_ops as ops
    a = torch.zeros(4, 64, device=DEV).to(FP8)
    b = torch.zeros(40, 64, device=DEV).to(FP8).t()
    s = torch.ones(1, device=DEV)
    with pytest.raises(RuntimeError, match="multiples of 16"):
        ops.cutlass_scaled_mm(a, b, s, s, torch.bfloat16)
    b2 = torch.zeros(64, 48, device=DEV).to(FP8)  # row-major b
    with pytest.raises(RuntimeError, match="column-major"):
        ops.cutlass_scaled_mm(a, b2, s, s, torch.bfloat16)
    assert ops.cutlass_scaled_mm_supports_fp8(100) and not ops.cutlass_scaled_mm_supports_fp8(89)


@pytest.mark.parametrize("M", [16, 64, 128, 256])
@pytest.mark.parametrize("N,K", [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336),      # Llama-3.1-8B
                                 (1280, 8192), (8192, 1024), (7168, 8192), (8192, 3584)])       # Llama-3.1-70B / TP 8
def test_scaled_mm_vs_reference_sm100_cutlass(M, N, K):
    """Against the kernel SURVEY.md calls "the Blackwell kernel to beat": the reference's own sm100 CUTLASS
    cutlass_scaled_mm (csrc/quantization/cutlass_w8a8/c3x/scaled_mm_sm100_fp8_dispatch.cuh:17-135) compiled from
    /root/reference into oracle/_ref/libvllm_ref_cutlass.so, same inputs, at the decode batch sizes.  Both kernels
    accumulate the same fp8 products in fp32 (in different orders) and round once to bf16: north_star bound
    1e-2 rel / 2e-2 abs, and at least 99 % of the outputs identical to one bf16 ulp."""
    from parity_utils import load_ref_cutlass
    from vllm_b200 import _custom_ops as ops
    ref = load_ref_cutlass()
    if ref is None:
        pytest.skip("oracle/_ref/libvllm_ref_cutlass.so not built / not loadable")
    seed_all(M + N + K)
    a = to_fp8(torch.randn(M, K, device=DEV))
    b = to_fp8(torch.randn(N, K, device=DEV)).t()
    sa = torch.rand(M, 1, device=DEV) * 0.1 + 0.01
    sb = torch.rand(1, N, device=DEV) * 0.1 + 0.01
    bias = (torch.rand(N, device=DEV) * 4).to(torch.bfloat16)
    for use_bias in (False, True):
        got = ops.cutlass_scaled_mm(a, b, sa, sb, torch.bfloat16, bias if use_bias else None)
        want = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
        ref.cutlass_scaled_mm(want, a, b, sa, sb, bias if use_bias else None)
        torch.cuda.synchronize()
        torch.testing.assert_close(got.float(), want.float(), rtol=1e-2, atol=2e-2 * max(1.0, float(want.float().abs().mean())))
        rel = (got.float() - want.float()).abs() / want.float().abs().clamp(min=1e-2)
        assert (rel <= 2.0 ** -7).float().mean().item() >= 0.99
