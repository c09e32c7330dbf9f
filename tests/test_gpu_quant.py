"""GPU parity: fp8 activation quantisation and silu_and_mul through the C ABI vs the oracle (bit-exact)
and the reference's own CUDA kernels.  Modelled on /root/reference/tests/kernels/quantization/test_fp8_quant.py
and tests/kernels/core/test_activation.py."""
import pytest
import torch

from oracle import quant as oq
from parity_utils import load_ref_cuda, seed_all

pytestmark = pytest.mark.gpu
DEV = "cuda"
SHAPES = [(1, 16), (7, 17), (83, 1024), (256, 4096), (33, 5137), (4, 14336)]


def _u8(t):
    return t.view(torch.uint8)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("shape", SHAPES)
def test_static_and_dynamic_per_tensor(dtype, shape):
    from vllm_b200 import _custom_ops as ops
    seed_all(shape[0])
    x = (torch.randn(shape) * 5).to(dtype)
    x[0, 0] = 1000.0  # saturates
    xd = x.to(DEV)
    scale = torch.tensor([0.73])
    q, s = ops.scaled_fp8_quant(xd, scale.to(DEV))
    assert torch.equal(_u8(q.cpu()), _u8(oq.static_scaled_fp8_quant(x, 0.73)))
    qd, sd = ops.scaled_fp8_quant(xd)
    qo, so = oq.dynamic_scaled_fp8_quant(x)
    assert torch.equal(sd.cpu(), so) and torch.equal(_u8(qd.cpu()), _u8(qo))
    ref = load_ref_cuda()
    if ref is not None:
        qr = torch.empty_like(qd)
        sr = torch.zeros(1, device=DEV)
        ref.dynamic_scaled_fp8_quant(qr, xd, sr)
        assert torch.equal(sr, sd) and torch.equal(_u8(qr), _u8(qd))
        ref.static_scaled_fp8_quant(qr, xd, scale.to(DEV))
        assert torch.equal(_u8(qr), _u8(q))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("ub", [None, 3.0])
def test_dynamic_per_token(dtype, shape, ub):
    from vllm_b200 import _custom_ops as ops
    seed_all(shape[1])
    x = (torch.randn(shape) * 2).to(dtype)
    if shape[0] > 2:
        x[1] = 0  # all-zero row -> min scaling factor
    xd = x.to(DEV)
    ubt = None if ub is None else torch.tensor([ub], device=DEV)
    q, s = ops.scaled_fp8_quant(xd, None, scale_ub=ubt, use_per_token_if_dynamic=True)
    qo, so = oq.dynamic_per_token_scaled_fp8_quant(x, ub)
    assert s.shape == (shape[0], 1)
    assert torch.equal(s.cpu(), so) and torch.equal(_u8(q.cpu()), _u8(qo))
    ref = load_ref_cuda()
    if ref is not None:
        qr, sr = torch.empty_like(q), torch.empty_like(s)
        ref.dynamic_per_token_scaled_fp8_quant(qr, xd, sr, ubt)
        assert torch.equal(sr, s) and torch.equal(_u8(qr), _u8(q))


def test_quant_padding_and_output_arg():
    from vllm_b200 import _custom_ops as ops
    x = torch.randn(5, 64, device=DEV).bfloat16()
    q, s = ops.scaled_fp8_quant(x, None, num_token_padding=17)
    assert q.shape == (17, 64) and s.numel() == 1
    out = torch.empty(5, 64, dtype=torch.float8_e4m3fn, device=DEV)
    q2, _ = ops.scaled_fp8_quant(x, s, output=out)
    assert q2 is out and torch.equal(_u8(out), _u8(q[:5]))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("shape", [(1, 2 * 8), (7, 2 * 13), (83, 2 * 512), (256, 2 * 14336)])
def test_silu_and_mul(dtype, shape):
    from vllm_b200 import _custom_ops as ops
    seed_all(4)
    x = (torch.randn(shape) * 3).to(dtype)
    xd = x.to(DEV)
    out = torch.empty(shape[0], shape[1] // 2, dtype=dtype, device=DEV)
    ops.silu_and_mul(out, xd)
    want = oq.silu_and_mul(x)
    ref = load_ref_cuda()
    if ref is not None:
        outr = torch.empty_like(out)
        ref.silu_and_mul(outr, xd)
        # same operation order as the reference kernel: equal up to expf's 1-2 ulp device/host difference
        torch.testing.assert_close(out.float(), outr.float(), rtol=2e-3 if dtype != torch.float32 else 1e-5, atol=1e-5)
        assert (out != outr).float().mean() < 1e-3
    torch.testing.assert_close(out.cpu().float(), want.float(), rtol=1e-2 if dtype != torch.float32 else 1e-5, atol=1e-5)
