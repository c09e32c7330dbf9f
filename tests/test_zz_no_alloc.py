"""GPU test of vb200_paged_attention_v1_static (include/vllm_b200.h): the paged_attention_v1 entry point that
allocates nothing -- its split-K partials live in a library-owned static device array.  No counterpart in the
reference; the check is against the existing entry point
vb200_paged_attention_v1 with a caller workspace, through the C ABI, on two calls in a row that share the array,
within the attention tests' tolerance (1e-2 rel / 2e-2 abs), and against the oracle.

The file name sorts last on purpose: a failure here must not hide other tests."""
import pytest
import torch

from oracle import attention as oattn
from parity_utils import make_block_tables, make_paged_kv, seed_all

pytestmark = pytest.mark.gpu
DEV = "cuda"
RTOL, ATOL = 1e-2, 2e-2


def _call(entry, q, kc, vc, hkv, scale, bt, sl, bs, kvd, ks, vs, part):
    """One call through ctypes.  entry = "static" (no workspace) or "workspace" (the existing entry point)."""
    from vllm_b200._lib import check, dtype_enum, kv_dtype_enum, load, ptr, stream
    lib = load()
    S, H, D = q.shape
    max_len = int(sl.max())
    qd, kd = dtype_enum(q.dtype), kv_dtype_enum(kvd)
    out = torch.full((S, H, D), float("nan"), dtype=q.dtype, device=q.device)
    head = (ptr(out), ptr(q), ptr(kc), ptr(vc), S, H, D, hkv, float(scale), ptr(bt), ptr(sl), bs, max_len,
            bt.stride(0), None, qd, kd, ptr(ks), ptr(vs), q.stride(0), kc.stride(0), kc.stride(1), part)
    if entry == "static":
        rc = lib.vb200_paged_attention_v1_static(*head, stream())
    else:
        nbytes = lib.vb200_paged_attention_v1_workspace_size(S, H, D, max_len, qd, part)
        assert nbytes > 0, "the case must split K, otherwise no scratch is exercised"
        ws = torch.empty(nbytes, dtype=torch.uint8, device=q.device)
        rc = lib.vb200_paged_attention_v1(*head, ptr(ws), nbytes, stream())
    check(rc)
    torch.cuda.synchronize()
    return out


def _case(seed, S, H, Hkv, D, bs, seq_lens, dtype, kvd):
    seed_all(seed)
    nb = S * ((max(seq_lens) + bs - 1) // bs) + 4
    ksc, vsc = (0.043, 0.029) if kvd == "fp8" else (1.0, 1.0)
    kc, vc = make_paged_kv(nb, bs, Hkv, D, dtype, kvd, dist="normal", k_scale=ksc, v_scale=vsc)
    q = (torch.randn(S, H, D) * 0.5).to(dtype)
    bt = make_block_tables(S, seq_lens, bs, nb)
    sl = torch.tensor(seq_lens, dtype=torch.int32)
    scale = float(D ** -0.5)
    want = oattn.paged_attention(q, kc, vc, Hkv, scale, bt, sl, None, kvd, ksc, vsc)
    dev = (q.to(DEV), kc.to(DEV), vc.to(DEV), Hkv, scale, bt.to(DEV), sl.to(DEV), bs, kvd,
           torch.tensor(ksc, device=DEV), torch.tensor(vsc, device=DEV))
    return dev, want


def test_static_scratch_matches_workspace_entry_two_calls_in_a_row():
    # first call: 6 sequences, 4 partitions of 256; second call: fewer / shorter sequences and another head count, so
    # its partials land on top of the first call's leftovers with a different row layout
    calls = [
        (101, 6, 32, 8, 128, 16, [1, 255, 257, 700, 1000, 1024], torch.bfloat16, "auto", 256),
        (202, 3, 16, 4, 128, 16, [300, 513, 40], torch.float16, "fp8", 128),
    ]
    for seed, S, H, Hkv, D, bs, seq_lens, dtype, kvd, part in calls:
        args, want = _case(seed, S, H, Hkv, D, bs, seq_lens, dtype, kvd)
        got = _call("static", *args, part)
        ref = _call("workspace", *args, part)
        assert not torch.isnan(got).any()
        torch.testing.assert_close(got.float(), ref.float(), rtol=RTOL, atol=ATOL)
        torch.testing.assert_close(got.cpu().float(), want.float(), rtol=RTOL, atol=ATOL)
    # and the first case once more: the array now holds the second call's partials
    seed, S, H, Hkv, D, bs, seq_lens, dtype, kvd, part = calls[0]
    args, want = _case(seed, S, H, Hkv, D, bs, seq_lens, dtype, kvd)
    got = _call("static", *args, 0)   # library's own partition plan
    torch.testing.assert_close(got.cpu().float(), want.float(), rtol=RTOL, atol=ATOL)


def test_static_scratch_rejects_oversized_requests():
    from vllm_b200._lib import load
    lib = load()
    limit = lib.vb200_paged_attention_v1_static_max_bytes()
    assert limit == 16 << 20
    S, H, Hkv, D, bs, part = 512, 64, 8, 128, 16, 128
    max_len = 4096
    assert lib.vb200_paged_attention_v1_workspace_size(S, H, D, max_len, 1, part) > limit
    seed_all(5)
    q = torch.zeros(S, H, D, dtype=torch.bfloat16, device=DEV)
    kc = torch.zeros(4, Hkv, D // 8, bs, 8, dtype=torch.bfloat16, device=DEV)
    vc = torch.zeros(4, Hkv, D, bs, dtype=torch.bfloat16, device=DEV)
    bt = torch.zeros(S, max_len // bs, dtype=torch.int32, device=DEV)
    sl = torch.full((S,), max_len, dtype=torch.int32, device=DEV)
    one = torch.ones((), device=DEV)
    with pytest.raises(RuntimeError, match="library-owned scratch"):
        _call("static", q, kc, vc, Hkv, D ** -0.5, bt, sl, bs, "auto", one, one, part)
