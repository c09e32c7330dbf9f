"""ORACLE (test infrastructure only): KV-cache scatter.

reshape_and_cache        /root/reference/csrc/cache_kernels.cu:210-262
reshape_and_cache_flash  /root/reference/csrc/cache_kernels.cu:264-306
convert_fp8              /root/reference/csrc/cache_kernels.cu:505-571
Python loops mirror the reference test's own golden loop (tests/kernels/attention/test_cache.py:196-200).
"""
import torch

from .fp8 import dequant_e4m3, quant_e4m3


def _scale(s) -> float:
    return float(s.item()) if isinstance(s, torch.Tensor) else float(s)


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype="auto",
                      k_scale=1.0, v_scale=1.0):
    """In place on CPU tensors.  key_cache [nb, H, D/x, bs, x], value_cache [nb, H, D, bs]."""
    nb, H, Dx, bs, x = key_cache.shape
    D = Dx * x
    fp8 = kv_cache_dtype in ("fp8", "fp8_e4m3")
    if fp8:
        kq = quant_e4m3(key, _scale(k_scale))
        vq = quant_e4m3(value, _scale(v_scale))
        kc, vc = key_cache.view(torch.uint8), value_cache.view(torch.uint8)
    else:
        kq, vq, kc, vc = key, value, key_cache, value_cache
    for t in range(slot_mapping.shape[0]):
        slot = int(slot_mapping[t])
        if slot < 0:  # padding token (cache_kernels.cu:224-227)
            continue
        b, off = slot // bs, slot % bs
        kc[b, :, :, off, :] = kq[t].reshape(H, D // x, x)
        vc[b, :, :, off] = vq[t]


def reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype="auto",
                            k_scale=1.0, v_scale=1.0):
    """In place.  caches are [nb, bs, H, D] tensors or HND-permuted views of [nb, H, bs, D]."""
    bs = key_cache.shape[1]
    fp8 = kv_cache_dtype in ("fp8", "fp8_e4m3")
    if fp8:
        kq = quant_e4m3(key, _scale(k_scale))
        vq = quant_e4m3(value, _scale(v_scale))
        kc, vc = key_cache.view(torch.uint8), value_cache.view(torch.uint8)
    else:
        kq, vq, kc, vc = key, value, key_cache, value_cache
    for t in range(slot_mapping.shape[0]):  # slot_mapping may be shorter than key (:421-431)
        slot = int(slot_mapping[t])
        if slot < 0:
            continue
        b, off = slot // bs, slot % bs
        kc[b, off] = kq[t]
        vc[b, off] = vq[t]


def convert_fp8(dst, src, scale=1.0):
    """Flat conversion (cache_kernels.cu:505-516): fp8->T dequant or T->fp8 quant, in place on dst."""
    if dst.element_size() == 1:
        dst.view(torch.uint8).copy_(quant_e4m3(src, scale).reshape(dst.shape))
    else:
        dst.copy_(dequant_e4m3(src.view(torch.uint8), scale, dst.dtype).reshape(dst.shape))
