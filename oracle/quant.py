"""ORACLE (test infrastructure only): FP8 activation quantisation, silu_and_mul, scaled_mm.

static_scaled_fp8_quant             /root/reference/csrc/quantization/fp8/common.cu:14-26, 87-111
                                    q = fp8(clamp(x * (1/scale), +-448))      (common.cuh:38-56)
dynamic_scaled_fp8_quant            common.cu:113-141; scale = amax/448 (common.cuh:64-99) then static
dynamic_per_token_scaled_fp8_quant  common.cu:28-83: s = max(min(amax, ub)/448, 1/(448*512)); q = fp8(clamp(x / s))
   (mirrors /root/reference/tests/kernels/quant_utils.py:22-98, the reference tests' own golden)
silu_and_mul                        /root/reference/csrc/activation_kernels.cu:26-63
baseline_scaled_mm                  /root/reference/tests/kernels/utils.py:1212-1251
scaled_mm_epilogue_order            csrc/cutlass_extensions/epilogue/scaled_mm_epilogues_c3x.hpp:137-216
"""
import torch

FP8_MAX = 448.0
FP8 = torch.float8_e4m3fn
_f32 = torch.float32


def static_scaled_fp8_quant(x: torch.Tensor, scale: float) -> torch.Tensor:
    inv = torch.tensor(1.0, dtype=_f32) / torch.tensor(float(scale), dtype=_f32)
    return (x.to(_f32) * inv).clamp(-FP8_MAX, FP8_MAX).to(FP8)


def dynamic_scaled_fp8_quant(x: torch.Tensor):
    scale = x.abs().max().to(_f32) / torch.tensor(FP8_MAX, dtype=_f32)
    return static_scaled_fp8_quant(x, float(scale)), scale.reshape(1)


def dynamic_per_token_scaled_fp8_quant(x: torch.Tensor, scale_ub=None):
    amax = x.abs().max(dim=-1).values.to(_f32)
    if scale_ub is not None:
        amax = amax.clamp(max=float(scale_ub))
    s = (amax / torch.tensor(FP8_MAX, dtype=_f32)).clamp(min=1.0 / (FP8_MAX * 512.0))[:, None]
    q = (x.to(_f32) / s).clamp(-FP8_MAX, FP8_MAX).to(FP8)
    return q, s


def silu_and_mul(x: torch.Tensor) -> torch.Tensor:
    """out = T(silu_f32(x[:d])) *_T x[d:]: silu rounded to T first, then a T*T product."""
    d = x.shape[-1] // 2
    a, b = x[..., :d], x[..., d:]
    af = a.to(_f32)
    s = (af / (1.0 + torch.exp(-af))).to(x.dtype)
    return (s.to(_f32) * b.to(_f32)).to(x.dtype)


def baseline_scaled_mm(a, b, scale_a, scale_b, out_dtype, bias=None):
    """The reference test's golden: mm((scale_a * a), (scale_b * b)) in fp32, cast, + bias."""
    out = torch.mm(scale_a.to(_f32) * a.to(_f32), scale_b.to(_f32) * b.to(_f32)).to(out_dtype)
    if bias is not None:
        out = out + bias
    return out


def scaled_mm_epilogue_order(a, b, scale_a, scale_b, out_dtype, bias=None):
    """The kernel's own operation order: OutT(a_scale * (b_scale * acc) [+ bias]) with an exact
    fp32 accumulator (fp8 x fp8 products are exact in fp32; fp64 sum stands in for the wide acc)."""
    acc = torch.mm(a.to(torch.float64), b.to(torch.float64)).to(_f32)
    sa = scale_a.to(_f32).reshape(-1, 1) if scale_a.numel() > 1 else scale_a.to(_f32).reshape(1, 1)
    sb = scale_b.to(_f32).reshape(1, -1) if scale_b.numel() > 1 else scale_b.to(_f32).reshape(1, 1)
    t = sb * acc
    if bias is not None:
        return torch.addcmul(bias.to(_f32).reshape(1, -1), sa, t).to(out_dtype)
    return (sa * t).to(out_dtype)
