"""`B200Platform`: the CUDA platform of the installed vLLM with two overrides --
  * `import_kernels()` loads this repo's shim built under the reference's library names (`_C`, `_C_cache_ops`,
    `_moe_C`: vllm_b200/lib/libvllm_b200_dropin.so) INSTEAD of the wheel's own extension, so every
    `torch.ops._C.*` / `vllm._custom_ops.*` call of the hot path lands in the sm_100a kernels
    (reference: Platform.import_kernels, /root/reference/vllm/platforms/interface.py:164-173);
  * `get_attn_backend_cls()` answers with the paged-attention backend of this repo, which the selector imports by
    qualified name (/root/reference/vllm/attention/selector.py:195-201, /root/reference/vllm/platforms/cuda.py:226-330).
Everything else (device queries, memory info, communicator, config checks) is inherited."""
import os
import sys
import types

BACKEND_CLS = "vllm_b200.attention.backend.B200PagedAttentionBackend"
_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN_SO = os.environ.get("VLLM_B200_DROPIN_LIB", os.path.join(_ROOT, "vllm_b200", "lib", "libvllm_b200_dropin.so"))
# extension modules of the wheel that must NOT be imported next to the shim (both would register TORCH_LIBRARY(_C))
_WHEEL_EXTENSIONS = ("vllm._C", "vllm._moe_C", "vllm._C_stable_libtorch", "vllm._rocm_C", "vllm._flashmla_C",
                     "vllm._flashmla_extension_C")


def keep_wheel_extension_out() -> None:
    """Must run before anything imports `vllm._custom_ops` (whose module body calls import_kernels())."""
    for name in _WHEEL_EXTENSIONS:
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)


def _base_platform():
    keep_wheel_extension_out()
    try:
        from vllm.platforms.cuda import CudaPlatform
        return CudaPlatform
    except Exception:  # noqa: BLE001  (no NVML / trimmed install): the abstract interface still gives the hooks
        from vllm.platforms.interface import Platform
        return Platform


_Base = _base_platform()
_kernels_loaded = False


class B200Platform(_Base):  # type: ignore[misc, valid-type]
    device_name = "cuda"
    device_type = "cuda"

    @classmethod
    def import_kernels(cls) -> None:
        global _kernels_loaded
        if _kernels_loaded:
            return
        import torch
        if not os.path.exists(DROPIN_SO):
            raise RuntimeError(f"vllm_b200: {DROPIN_SO} not found -- build it with __graft_entry__.build() "
                               "(there is no fallback to the wheel's kernels)")
        keep_wheel_extension_out()
        torch.ops.load_library(DROPIN_SO)
        _kernels_loaded = True

    @classmethod
    def get_attn_backend_cls(cls, *args, **kwargs) -> str:
        """Signature differs between vLLM versions (reference: selected_backend, head_size, dtype, kv_cache_dtype,
        block_size, use_v1, use_mla; newer wheels: selected_backend, attn_selector_config, ...): the answer does not
        depend on it -- MLA is the one case this backend does not serve."""
        use_mla = kwargs.get("use_mla", args[6] if len(args) > 6 else False)
        if use_mla:
            raise NotImplementedError("the B200 paged-attention backend does not implement MLA")
        return BACKEND_CLS


def resolve_platform():
    """What `vllm.platforms.resolve_obj_by_qualname(register())` yields."""
    return B200Platform
