"""Mirror of /root/reference/vllm/model_executor/layers/quantization/utils/w8a8_utils.py:140-152
(cutlass_w8a8_scaled_mm) and :297-378 (Fp8LinearOp): view 2-D, quantise the activation if it is not fp8 already
(static per-tensor / dynamic per-tensor / dynamic per-token), run cutlass_scaled_mm, reshape to [..., N].
There is one GEMM path (the tcgen05 kernel): no torch._scaled_mm / Triton dispatch."""
from typing import Optional

import torch

from .. import _custom_ops as ops

FP8_DTYPE = torch.float8_e4m3fn


def cutlass_w8a8_scaled_mm(*, qinput: torch.Tensor, weight: torch.Tensor, out_dtype: torch.dtype,
                           scale_a: torch.Tensor, scale_b: torch.Tensor, bias: Optional[torch.Tensor],
                           output_shape, **kwargs) -> torch.Tensor:
    """w8a8_utils.py:140-152.  `weight` is [K, N] column-major (the [N, K] checkpoint tensor transposed)."""
    output = ops.cutlass_scaled_mm(qinput, weight, out_dtype=out_dtype, scale_a=scale_a, scale_b=scale_b, bias=bias)
    return output.view(*output_shape)


class Fp8LinearOp:
    """w8a8_utils.py:297-378.  act_quant_static: `input_scale` is a calibrated scalar; otherwise the scale is
    computed from the activation -- per token when `use_per_token_if_dynamic` (the reference's
    act_quant_group_shape == PER_TOKEN), else per tensor."""

    def __init__(self, act_quant_static: bool = False, cutlass_fp8_supported: bool = True,
                 use_per_token_if_dynamic: bool = True, pad_output: Optional[bool] = None):
        if not cutlass_fp8_supported:
            raise RuntimeError("vllm_b200 has only the cutlass_scaled_mm-equivalent path")
        self.cutlass_fp8_supported = True
        self.act_quant_static = act_quant_static
        self.use_per_token_if_dynamic = use_per_token_if_dynamic
        self.output_padding = None  # padding only helps torch._scaled_mm (w8a8_utils.py:311-321)

    def apply(self, input: torch.Tensor, weight: torch.Tensor, weight_scale: torch.Tensor,
              out_dtype: Optional[torch.dtype] = None, input_scale: Optional[torch.Tensor] = None,
              input_scale_ub: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        input_2d = input.view(-1, input.shape[-1])
        output_shape = [*input.shape[:-1], weight.shape[1]]
        if out_dtype is None:
            out_dtype = input.dtype
        if input.dtype != FP8_DTYPE:
            qinput, x_scale = ops.scaled_fp8_quant(input_2d, input_scale if self.act_quant_static else None,
                                                   scale_ub=input_scale_ub,
                                                   use_per_token_if_dynamic=self.use_per_token_if_dynamic)
        else:
            if input_scale is None:
                raise RuntimeError("Fp8LinearOp.apply: an fp8 input needs its input_scale")
            qinput, x_scale = input_2d, input_scale
        return cutlass_w8a8_scaled_mm(qinput=qinput, weight=weight, out_dtype=out_dtype, scale_a=x_scale,
                                      scale_b=weight_scale, bias=bias, output_shape=output_shape)
