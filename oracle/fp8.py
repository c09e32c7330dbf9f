"""ORACLE (test infrastructure only): fp8 e4m3fn conversions.

Follows /root/reference/csrc/quantization/fp8/nvidia/quant_utils.cuh:
  dequant  :348-359   bf16( float( half(fp8) ) * scale )
  quant    :466-479   fp8_e4m3_satfinite( float(x) / scale )
An independent numpy encoder/decoder (no torch float8 support needed) is provided to pin torch's
casts bit-for-bit.
"""
import numpy as np
import torch

FP8_MAX = 448.0


def e4m3_decode_table() -> np.ndarray:
    """All 256 e4m3fn byte values -> float32 (NaN for 0x7f / 0xff)."""
    out = np.zeros(256, dtype=np.float32)
    for b in range(256):
        s = -1.0 if b & 0x80 else 1.0
        e = (b >> 3) & 0xF
        m = b & 0x7
        if e == 0xF and m == 0x7:
            v = np.nan
        elif e == 0:
            v = s * (m / 8.0) * 2.0 ** -6
        else:
            v = s * (1.0 + m / 8.0) * 2.0 ** (e - 7)
        out[b] = v
    return out


_DEC = e4m3_decode_table()


def e4m3_decode_np(b: np.ndarray) -> np.ndarray:
    return _DEC[b.astype(np.uint8)]


def e4m3_encode_satfinite_np(x: np.ndarray) -> np.ndarray:
    """float32 -> e4m3fn byte, round-to-nearest-even, saturating to +-448 (NaN -> 0x7f).
    Pure-numpy restatement of __nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3)."""
    x = np.asarray(x, dtype=np.float32)
    sign = (np.signbit(x)).astype(np.uint8) << 7
    a = np.abs(x).astype(np.float64)
    out = np.zeros(x.shape, dtype=np.uint8)
    nan = np.isnan(a)
    a = np.where(nan, 0.0, a)
    a = np.minimum(a, FP8_MAX)
    # exponent of the value; subnormal quantum 2^-9, normal quantum 2^(e-3)
    with np.errstate(divide="ignore"):
        e = np.floor(np.log2(np.where(a > 0, a, 1.0)))
    e = np.clip(e, -6, 8)
    q = 2.0 ** (e - 3)
    n = np.rint(a / q)  # numpy rint = round half to even; a/q exact in fp64
    v = n * q
    # re-normalise if rounding bumped to the next binade
    v = np.minimum(v, FP8_MAX)
    with np.errstate(divide="ignore"):
        e2 = np.floor(np.log2(np.where(v > 0, v, 1.0)))
    e2 = np.clip(e2, -6, 8)
    is_sub = v < 2.0 ** -6
    mant = np.where(is_sub, np.rint(v / 2.0 ** -9), np.rint((v / 2.0 ** e2 - 1.0) * 8.0))
    expf = np.where(is_sub, 0, e2 + 7)
    out = (expf.astype(np.uint8) << 3) | mant.astype(np.uint8)
    out = np.where(v == 0, 0, out).astype(np.uint8)
    out = out | sign
    out = np.where(nan, np.uint8(0x7F) | sign, out)
    return out.astype(np.uint8)


def quant_e4m3(x: torch.Tensor, scale: float) -> torch.Tensor:
    """fp8 bytes (uint8) of fp8_satfinite(float(x) / scale)  (quant_utils.cuh:466-479)."""
    y = (x.float() / torch.tensor(scale, dtype=torch.float32)).clamp(-FP8_MAX, FP8_MAX)
    return y.to(torch.float8_e4m3fn).view(torch.uint8)


def dequant_e4m3(b: torch.Tensor, scale: float, dtype: torch.dtype) -> torch.Tensor:
    """dtype( float( half(fp8) ) * scale )  (quant_utils.cuh:348-359)."""
    f = b.view(torch.float8_e4m3fn).to(torch.float16).float()
    return (f * torch.tensor(scale, dtype=torch.float32)).to(dtype)
