"""Shared fixtures for the parity tests (modelled on the reference's tests/kernels helpers:
kv_cache_factory = create_kv_caches_with_random, /root/reference/vllm/utils/__init__.py:1076-1127)."""
import os
import random

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_CPU_SO = os.path.join(ROOT, "oracle", "_ref", "libvllm_ref_cpu.so")
REF_CUDA_SO = os.path.join(ROOT, "oracle", "_ref", "libvllm_ref_cuda.so")

_loaded = {}


def load_ref_cpu():
    """torch.ops._ref_cpu (the reference's own CPU kernels) or None when not built."""
    if "cpu" not in _loaded:
        try:
            torch.ops.load_library(REF_CPU_SO)
            _loaded["cpu"] = torch.ops._ref_cpu
        except Exception:
            _loaded["cpu"] = None
    return _loaded["cpu"]


def load_ref_cuda():
    """torch.ops._ref_C (the reference's own CUDA kernels built for sm_100a) or None."""
    if "cuda" not in _loaded:
        try:
            torch.ops.load_library(REF_CUDA_SO)
            _loaded["cuda"] = torch.ops._ref_C
        except Exception:
            _loaded["cuda"] = None
    return _loaded["cuda"]


def seed_all(seed=0):
    random.seed(seed)
    torch.manual_seed(seed)


def make_paged_kv(num_blocks, block_size, num_kv_heads, head_size, dtype, kv_cache_dtype="auto",
                  device="cpu", dist="uniform", k_scale=1.0, v_scale=1.0):
    """Random caches in the reference paged layout.  Returns (key_cache, value_cache); fp8 caches
    are uint8 tensors quantised from N(0,1)/U(-s,s) values with the given scales (x = 16)."""
    from oracle.fp8 import quant_e4m3
    fp8 = kv_cache_dtype in ("fp8", "fp8_e4m3")
    x = 16 if fp8 else 16 // torch.tensor([], dtype=dtype).element_size()
    s = head_size ** -0.5
    kshape = (num_blocks, num_kv_heads, head_size // x, block_size, x)
    vshape = (num_blocks, num_kv_heads, head_size, block_size)
    if dist == "uniform":
        k = torch.empty(kshape, dtype=torch.float32).uniform_(-s, s)
        v = torch.empty(vshape, dtype=torch.float32).uniform_(-s, s)
    else:  # "normal": peaked softmax, catches K-layout errors (SURVEY.md section 4 caveat)
        k = torch.randn(kshape, dtype=torch.float32)
        v = torch.randn(vshape, dtype=torch.float32)
    if fp8:
        k, v = quant_e4m3(k, k_scale), quant_e4m3(v, v_scale)
    else:
        k, v = k.to(dtype), v.to(dtype)
    return k.to(device), v.to(device)


def make_block_tables(num_seqs, seq_lens, block_size, num_blocks, shared=False):
    """int32 [num_seqs, max_blocks]; distinct random blocks per sequence unless `shared`."""
    max_blocks = (max(seq_lens) + block_size - 1) // block_size
    max_blocks = max(max_blocks, 1)
    if shared:
        bt = torch.randint(0, num_blocks, (num_seqs, max_blocks), dtype=torch.int32)
    else:
        need = num_seqs * max_blocks
        assert need <= num_blocks, (need, num_blocks)
        perm = torch.randperm(num_blocks)[:need].to(torch.int32)
        bt = perm.reshape(num_seqs, max_blocks)
    return bt


REF_CUTLASS_SO = os.path.join(ROOT, "oracle", "_ref", "libvllm_ref_cutlass.so")


def load_ref_cutlass():
    """torch.ops._ref_cutlass (the reference's own sm100 CUTLASS kernels: cutlass_scaled_mm, cutlass_moe_mm,
    get_cutlass_moe_mm_data) or None when not built / not loadable (needs libcuda.so.1)."""
    if "cutlass" not in _loaded:
        try:
            torch.ops.load_library(REF_CUTLASS_SO)
            _loaded["cutlass"] = torch.ops._ref_cutlass
        except Exception:
            _loaded["cutlass"] = None
    return _loaded["cutlass"]
