"""FP8 w8a8 linear glue (Fp8LinearOp) of the B200 decode path."""
from .fp8_linear import Fp8LinearOp, cutlass_w8a8_scaled_mm
from .unquantized_linear import linear_16

__all__ = ["Fp8LinearOp", "cutlass_w8a8_scaled_mm", "linear_16"]
