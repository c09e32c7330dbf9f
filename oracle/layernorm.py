"""ORACLE (test infrastructure only): RMSNorm family and rotary embedding.

rms_norm                          /root/reference/csrc/layernorm_kernels.cu:16-100
                                  out = T(T(x * rsqrt(mean(x^2) + eps)) * w)
fused_add_rms_norm                layernorm_kernels.cu:102-200: z = T(x + r); r <- z; x <- rms_norm(z)
rms_norm_dynamic_per_token_quant  /root/reference/csrc/quantization/fused_kernels/layernorm_utils.cuh:19-125:
                                  z = float(x) + float(r) (unrounded; r <- T(z)), y = T(T(z*inv)*w),
                                  s = max(min(amax|y|, ub)/448, 1/(448*512)), q = fp8(clamp(y / s))
   (mirrors the reference test's own golden, /root/reference/tests/kernels/core/test_fused_quant_layernorm.py:30-64)
rotary_embedding                  /root/reference/csrc/pos_encoding_kernels.cu:10-98: T arithmetic, one rounding
                                  per operation:  x' = T(T(x*c) - T(y*s)),  y' = T(T(y*c) + T(x*s))
   (fp32 variant of the same math: vllm/model_executor/layers/rotary_embedding.py forward_native)
"""
import torch

_f32 = torch.float32
FP8_MAX = 448.0
FP8 = torch.float8_e4m3fn


def _inv_rms(zf: torch.Tensor, eps: float) -> torch.Tensor:
    return torch.rsqrt(zf.pow(2).mean(dim=-1, keepdim=True) + eps)


def rms_norm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    xf = x.to(_f32)
    n = (xf * _inv_rms(xf, eps)).to(x.dtype)
    return (n.to(_f32) * w.to(_f32)).to(x.dtype)


def fused_add_rms_norm(x: torch.Tensor, residual: torch.Tensor, w: torch.Tensor, eps: float):
    """Returns (normed, new_residual)."""
    z = (x.to(_f32) + residual.to(_f32)).to(x.dtype)
    return rms_norm(z, w, eps), z


def rms_norm_dynamic_per_token_quant(x, w, eps, scale_ub=None, residual=None):
    """Returns (q fp8, scales [T,1], new_residual or None)."""
    zf = x.to(_f32)
    new_res = None
    if residual is not None:
        zf = zf + residual.to(_f32)
        new_res = zf.to(x.dtype)
    n = (zf * _inv_rms(zf, eps)).to(x.dtype)
    y = (n.to(_f32) * w.to(_f32)).to(x.dtype).to(_f32)
    amax = y.abs().max(dim=-1, keepdim=True).values
    if scale_ub is not None:
        amax = amax.clamp(max=float(scale_ub))
    s = (amax / torch.tensor(FP8_MAX, dtype=_f32)).clamp(min=1.0 / (FP8_MAX * 512.0))
    q = (y / s).clamp(-FP8_MAX, FP8_MAX).to(FP8)
    return q, s, new_res


def make_cos_sin_cache(head_size: int, rot_dim: int, max_position: int, base: float, dtype) -> torch.Tensor:
    """vllm/model_executor/layers/rotary_embedding.py RotaryEmbedding._compute_cos_sin_cache: [max_pos, rot_dim]."""
    inv_freq = 1.0 / (base ** (torch.arange(0, rot_dim, 2, dtype=_f32) / rot_dim))
    t = torch.arange(max_position, dtype=_f32)
    freqs = torch.einsum("i,j->ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(dtype)


def rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox):
    """Returns (query', key') as new tensors; query/key are [T, H, D] (or [T, H*D])."""
    T = positions.numel()
    rot_dim = cos_sin_cache.shape[1]
    embed = rot_dim // 2
    dt = query.dtype
    cs = cos_sin_cache[positions.long()]  # [T, rot_dim]
    cos = cs[:, :embed].to(_f32)[:, None, :]
    sin = cs[:, embed:].to(_f32)[:, None, :]

    def rnd(v):
        return v.to(dt).to(_f32)

    def apply(t):
        if t is None:
            return None
        shp = t.shape
        t3 = t.reshape(T, -1, head_size).clone()
        rot = t3[..., :rot_dim].to(_f32)
        if is_neox:
            x, y = rot[..., :embed], rot[..., embed:]
        else:
            x, y = rot[..., 0::2], rot[..., 1::2]
        xo = (rnd(x * cos) - rnd(y * sin)).to(dt)
        yo = (rnd(y * cos) + rnd(x * sin)).to(dt)
        if is_neox:
            t3[..., :embed], t3[..., embed:rot_dim] = xo, yo
        else:
            t3[..., 0:rot_dim:2], t3[..., 1:rot_dim:2] = xo, yo
        return t3.reshape(shp)

    return apply(query), apply(key)
