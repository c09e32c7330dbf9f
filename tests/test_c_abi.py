"""CPU-side checks of the drop-in boundary: libvllm_b200.so loads and exports every symbol that
include/vllm_b200.h declares, and the Python host mirrors the reference wrapper names."""
import ctypes
import inspect
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    txt = open(os.path.join(ROOT, "include", "vllm_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(vb200_[a-z0-9_]+)\s*\(", txt)))


def test_header_declares_symbols():
    syms = _declared_symbols()
    assert "vb200_paged_attention_v1" in syms and "vb200_reshape_and_cache" in syms
    assert len(syms) >= 10


def test_library_exports_every_declared_symbol():
    from vllm_b200 import _lib
    assert os.path.exists(_lib.LIB_PATH), "build with python vllm_b200/build.py"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [s for s in _declared_symbols() if not hasattr(lib, s)]
    assert not missing, f"symbols declared in include/vllm_b200.h but not exported: {missing}"


def test_ctypes_table_matches_header():
    from vllm_b200 import _lib
    assert sorted(_lib.SIGNATURES) == _declared_symbols()
    lib = _lib.load()
    assert lib.vb200_version().decode().startswith("vllm_b200")


def test_missing_library_fails_loudly(monkeypatch):
    from vllm_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libvllm_b200.so")
    with pytest.raises(RuntimeError, match="no fallback"):
        _lib.load()


def test_custom_ops_mirror_reference_signatures():
    """Argument names/order of the reference wrappers (vllm/_custom_ops.py:37-97, 1617-1660)."""
    from vllm_b200 import _custom_ops as ops
    v1 = list(inspect.signature(ops.paged_attention_v1).parameters)
    assert v1[:19] == ["out", "query", "key_cache", "value_cache", "num_kv_heads", "scale", "block_tables",
                       "seq_lens", "block_size", "max_seq_len", "alibi_slopes", "kv_cache_dtype", "k_scale",
                       "v_scale", "tp_rank", "blocksparse_local_blocks", "blocksparse_vert_stride",
                       "blocksparse_block_size", "blocksparse_head_sliding_step"]
    v2 = list(inspect.signature(ops.paged_attention_v2).parameters)
    assert v2[:4] == ["out", "exp_sum", "max_logits", "tmp_out"] and v2[4:] == v1[1:19]
    rc = list(inspect.signature(ops.reshape_and_cache).parameters)
    assert rc == ["key", "value", "key_cache", "value_cache", "slot_mapping", "kv_cache_dtype", "k_scale",
                  "v_scale"]
    assert list(inspect.signature(ops.reshape_and_cache_flash).parameters) == rc


def test_cpu_tensors_are_rejected():
    import torch
    from vllm_b200 import _custom_ops as ops
    k = torch.zeros(2, 1, 16, dtype=torch.bfloat16)
    kc = torch.zeros(4, 1, 2, 16, 8, dtype=torch.bfloat16)
    vc = torch.zeros(4, 1, 16, 16, dtype=torch.bfloat16)
    one = torch.ones(1)
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.reshape_and_cache(k, k, kc, vc, torch.zeros(2, dtype=torch.int64), "auto", one, one)
