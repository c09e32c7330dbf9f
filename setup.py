"""Packaging of the two Python packages of this repo:
  vllm_b200         the B200 hot-path ops (ctypes over lib/libvllm_b200.so) + the host-side mirrors of the reference
                    interfaces (attention backend, fused_moe, Fp8LinearOp, peer all-reduce)
  vllm_b200_plugin  the vLLM platform plugin (entry point group `vllm.platform_plugins`) that makes an unmodified vLLM
                    load these kernels instead of its own extension
Build the native libraries first (`python vllm_b200/build.py --shim`, or __graft_entry__.build()); they ship as
package data."""
from setuptools import setup

setup(
    name="vllm_b200",
    version="0.2",
    packages=["vllm_b200", "vllm_b200.attention", "vllm_b200.distributed", "vllm_b200.fused_moe",
              "vllm_b200.quantization", "vllm_b200_plugin"],
    package_data={"vllm_b200": ["lib/*.so", "csrc/*", "torch_shim.cpp"]},
    entry_points={"vllm.platform_plugins": ["b200 = vllm_b200_plugin:register"]},
)
