"""GPU parity: reshape_and_cache / reshape_and_cache_flash / convert_fp8 through the C ABI vs the oracle
(bit-exact) and vs the reference's own CUDA kernels (oracle/_ref/libvllm_ref_cuda.so) when loadable.
Modelled on /root/reference/tests/kernels/attention/test_cache.py:122-330."""
import pytest
import torch

from oracle import cache as ocache
from oracle import fp8 as ofp8
from parity_utils import load_ref_cuda, seed_all

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _bits(t):
    return t.view(torch.uint8) if t.element_size() == 1 else t.view(torch.int16) if t.element_size() == 2 else t.view(torch.int32)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("num_tokens,num_heads,head_size,block_size", [(42, 8, 128, 16), (1, 1, 64, 8), (257, 5, 80, 32),
                                                                       (64, 8, 256, 16), (7, 3, 120, 16)])
def test_reshape_and_cache(dtype, kv_cache_dtype, num_tokens, num_heads, head_size, block_size):
    from vllm_b200 import _custom_ops as ops
    fp8 = kv_cache_dtype == "fp8"
    if fp8 and head_size % 16:
        pytest.skip("fp8 x = 16 needs head_size % 16 == 0")
    seed_all(0)
    x = 16 if fp8 else 16 // torch.tensor([], dtype=dtype).element_size()
    nb = (num_tokens + block_size - 1) // block_size + 3
    cdt = torch.uint8 if fp8 else dtype
    # key/value are slices of a fused qkv tensor: arbitrary row stride (cache_kernels.cu:385-386)
    qkv = torch.randn(num_tokens, 3, num_heads, head_size, dtype=torch.float32).to(dtype)
    key, value = qkv[:, 1], qkv[:, 2]
    kc = torch.randint(0, 255, (nb, num_heads, head_size // x, block_size, x), dtype=torch.uint8).to(cdt) if fp8 else \
        torch.randn(nb, num_heads, head_size // x, block_size, x).to(dtype)
    vc = torch.randint(0, 255, (nb, num_heads, head_size, block_size), dtype=torch.uint8).to(cdt) if fp8 else \
        torch.randn(nb, num_heads, head_size, block_size).to(dtype)
    slots = torch.randperm(nb * block_size)[:num_tokens].to(torch.int64)
    if num_tokens > 4:
        slots[3] = -1  # padding token must be skipped
    ks, vs = torch.tensor(0.037 if fp8 else 1.0), torch.tensor(0.051 if fp8 else 1.0)
    kc_o, vc_o = kc.clone(), vc.clone()
    ocache.reshape_and_cache(key, value, kc_o, vc_o, slots, kv_cache_dtype, ks, vs)
    kc_d, vc_d = kc.to(DEV), vc.to(DEV)
    ops.reshape_and_cache(key.to(DEV), value.to(DEV), kc_d, vc_d, slots.to(DEV), kv_cache_dtype, ks.to(DEV), vs.to(DEV))
    torch.cuda.synchronize()
    assert torch.equal(_bits(kc_d.cpu()), _bits(kc_o))
    assert torch.equal(_bits(vc_d.cpu()), _bits(vc_o))
    ref = load_ref_cuda()
    if ref is not None:
        kc_r, vc_r = kc.to(DEV), vc.to(DEV)
        ref.reshape_and_cache(key.to(DEV), value.to(DEV), kc_r, vc_r, slots.to(DEV), kv_cache_dtype, ks.to(DEV), vs.to(DEV))
        assert torch.equal(_bits(kc_d), _bits(kc_r)) and torch.equal(_bits(vc_d), _bits(vc_r))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("num_tokens,num_heads,head_size,block_size", [(42, 8, 128, 16), (3, 2, 80, 32), (300, 4, 64, 16)])
def test_reshape_and_cache_flash(dtype, kv_cache_dtype, layout, num_tokens, num_heads, head_size, block_size):
    from vllm_b200 import _custom_ops as ops
    fp8 = kv_cache_dtype == "fp8"
    seed_all(1)
    nb = (num_tokens + block_size - 1) // block_size + 2
    cdt = torch.uint8 if fp8 else dtype
    shape = (2, nb, block_size, num_heads, head_size) if layout == "NHD" else (2, nb, num_heads, block_size, head_size)
    base = torch.randint(0, 255, shape, dtype=torch.uint8).to(cdt) if fp8 else torch.randn(shape).to(dtype)

    def views(t):
        k, v = t[0], t[1]
        return (k, v) if layout == "NHD" else (k.permute(0, 2, 1, 3), v.permute(0, 2, 1, 3))

    # key longer than slot_mapping (CUDA-graph padding, cache_kernels.cu:421-431)
    key = torch.randn(num_tokens + 5, num_heads, head_size).to(dtype)
    value = torch.randn(num_tokens + 5, num_heads, head_size).to(dtype)
    slots = torch.randperm(nb * block_size)[:num_tokens].to(torch.int64)
    slots[0] = -1
    ks, vs = torch.tensor(0.02 if fp8 else 1.0), torch.tensor(0.03 if fp8 else 1.0)
    base_o = base.clone()
    ko, vo = views(base_o)
    ocache.reshape_and_cache_flash(key, value, ko, vo, slots, kv_cache_dtype, ks, vs)
    base_d = base.to(DEV)
    kd, vd = views(base_d)
    ops.reshape_and_cache_flash(key.to(DEV), value.to(DEV), kd, vd, slots.to(DEV), kv_cache_dtype, ks.to(DEV), vs.to(DEV))
    torch.cuda.synchronize()
    assert torch.equal(_bits(base_d.cpu()), _bits(base_o))
    ref = load_ref_cuda()
    if ref is not None:
        base_r = base.to(DEV)
        kr, vr = views(base_r)
        ref.reshape_and_cache_flash(key.to(DEV), value.to(DEV), kr, vr, slots.to(DEV), kv_cache_dtype, ks.to(DEV), vs.to(DEV))
        assert torch.equal(_bits(base_d), _bits(base_r))


def test_reshape_and_cache_empty_and_all_padding():
    from vllm_b200 import _custom_ops as ops
    kc = torch.zeros(2, 2, 16, 16, 8, dtype=torch.bfloat16, device=DEV)
    vc = torch.zeros(2, 2, 128, 16, dtype=torch.bfloat16, device=DEV)
    one = torch.ones(1, device=DEV)
    k = torch.randn(0, 2, 128, dtype=torch.bfloat16, device=DEV)
    ops.reshape_and_cache(k, k, kc, vc, torch.zeros(0, dtype=torch.int64, device=DEV), "auto", one, one)
    k = torch.randn(4, 2, 128, dtype=torch.bfloat16, device=DEV)
    ops.reshape_and_cache(k, k, kc, vc, torch.full((4,), -1, dtype=torch.int64, device=DEV), "auto", one, one)
    torch.cuda.synchronize()
    assert kc.abs().sum() == 0 and vc.abs().sum() == 0
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        ops.reshape_and_cache(k, k, kc, vc, torch.zeros(4, dtype=torch.int64, device=DEV), "fp8_e5m2", one, one)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_convert_fp8_roundtrip(dtype):
    from vllm_b200 import _custom_ops as ops
    seed_all(2)
    x = (torch.randn(3, 1000) * 20).to(dtype)
    q = torch.empty(3, 1000, dtype=torch.uint8, device=DEV)
    ops.convert_fp8(q, x.to(DEV), 0.25, "fp8")
    assert torch.equal(q.cpu(), ofp8.quant_e4m3(x, 0.25))
    back = torch.empty(3, 1000, dtype=dtype, device=DEV)
    ops.convert_fp8(back, q, 0.25, "fp8")
    assert torch.equal(_bits(back.cpu()), _bits(ofp8.dequant_e4m3(q.cpu(), 0.25, dtype)))
    ref = load_ref_cuda()
    if ref is not None:
        q2 = torch.empty_like(q)
        ref.convert_fp8(q2, x.to(DEV), 0.25, "fp8")
        assert torch.equal(q, q2)


def test_full_size_scatter_gather_roundtrip():
    """BASELINE cfg 3 decode-step size (256 tokens x 8 kv heads x 128) -- property: what was scattered is what
    the block-table gather reads back (bit-exact), every other slot untouched."""
    from vllm_b200 import _custom_ops as ops
    seed_all(3)
    T, H, D, bs, nb = 256, 8, 128, 16, 1024
    key = torch.randn(T, H, D, device=DEV).bfloat16()
    value = torch.randn(T, H, D, device=DEV).bfloat16()
    kc = torch.zeros(nb, H, D // 8, bs, 8, dtype=torch.bfloat16, device=DEV)
    vc = torch.zeros(nb, H, D, bs, dtype=torch.bfloat16, device=DEV)
    slots = torch.randperm(nb * bs, device=DEV)[:T].to(torch.int64)
    one = torch.ones(1, device=DEV)
    ops.reshape_and_cache(key, value, kc, vc, slots, "auto", one, one)
    b, o = slots // bs, slots % bs
    kg = kc[b, :, :, o, :].reshape(T, H, D)
    vg = vc[b, :, :, o]
    assert torch.equal(kg, key) and torch.equal(vg, value)
    assert int((kc != 0).sum()) == int((key != 0).sum()) and int((vc != 0).sum()) == int((value != 0).sum())
