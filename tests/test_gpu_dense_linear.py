"""GPU parity: unquantised dense linear on the tcgen05 grouped GEMM (vllm_b200/quantization/unquantized_linear.py; the
reference's UnquantizedLinearMethod.apply = F.linear, vllm/model_executor/layers/linear.py:221-230) against a plain
PyTorch fp32 reference of the same op (floating-point kernel: fp32 accumulation, one rounding to the activation dtype;
tolerance = two units in the last place of the output dtype relative to the row's magnitude)."""
import pytest
import torch

from parity_utils import seed_all

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("T,N,K", [(1, 128, 64), (7, 300, 136), (64, 6144, 4096), (256, 4096, 4096), (256, 1000, 14336),
                                   (300, 515, 256), (700, 256, 512), (256, 32064, 1024)])
def test_linear_16(T, N, K, dtype):
    from vllm_b200.quantization import linear_16
    seed_all(T + N + K)
    x = torch.randn(T, K, device=DEV).to(dtype)
    w = (torch.randn(N, K, device=DEV) / K ** 0.5).to(dtype)
    out = linear_16(x, w)
    torch.cuda.synchronize()
    want = x.float() @ w.float().t()          # fp32 reference of the same op on the same (rounded) operands
    ulp = 2.0 ** -7 if dtype == torch.bfloat16 else 2.0 ** -10
    tol = 2 * ulp * want.abs().amax() + 1e-6
    assert out.shape == (T, N) and out.dtype == dtype
    assert (out.float() - want).abs().max() <= tol
    # a second call reuses the cached routing table; with a bias
    b = torch.randn(N, device=DEV).to(dtype)
    out2 = linear_16(x, w, bias=b)
    assert (out2.float() - (want + b.float())).abs().max() <= tol + 2 * ulp * b.float().abs().max()


def test_linear_16_in_a_cuda_graph_and_errors():
    from vllm_b200.quantization import linear_16
    seed_all(1)
    x = torch.randn(32, 512, device=DEV).bfloat16()
    w = torch.randn(640, 512, device=DEV).bfloat16() / 22
    out = torch.empty(32, 640, dtype=torch.bfloat16, device=DEV)
    linear_16(x, w, out=out)                  # warm-up: builds the routing table
    want = out.clone()
    g = torch.cuda.CUDAGraph()
    out.zero_()
    with torch.cuda.graph(g):
        linear_16(x, w, out=out)
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    with pytest.raises(RuntimeError):
        linear_16(x, w.float())
    with pytest.raises(RuntimeError):
        linear_16(x.cpu(), w.cpu())
