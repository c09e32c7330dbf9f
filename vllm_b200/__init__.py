"""vllm_b200: B200-native (sm_100a) implementation of vLLM's decode hot path.

Layout (hot path only -- see DESIGN.md):
  csrc/            hand-written CUDA kernels + the C ABI (include/vllm_b200.h)
  _lib.py          ctypes binding of lib/libvllm_b200.so (fails loudly when it is missing)
  _custom_ops.py   host-side mirror of /root/reference/vllm/_custom_ops.py (same names / args)
  torch_bindings.py  torch.library registration under the reference's schemas
  attention/       PagedAttention helper + AttentionBackend (vllm/attention interface)
  fused_moe/       fused_topk / moe_align_block_size / fused_experts glue
  quantization/    Fp8LinearOp glue
  distributed/     tensor-parallel all-reduce call site (NCCL)
"""
__version__ = "0.1.0"
