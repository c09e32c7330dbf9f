#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY (oracle).  Recipe that compiles the reference's OWN sources, where
they lie under /root/reference, into oracle/_ref/ (git-ignored; travels to the GPU box):

  oracle/_ref/libvllm_ref_cpu.so   csrc/cpu/attention.cpp + csrc/cpu/cache.cpp (AVX512, OpenMP)
                                   -> torch.ops._ref_cpu.{paged_attention_v1,v2,reshape_and_cache}
  oracle/_ref/libvllm_ref_cuda.so  csrc/attention/paged_attention_v{1,2}.cu, cache_kernels.cu,
                                   quantization/fp8/common.cu, activation_kernels.cu,
                                   moe/{topk_softmax,moe_align_sum}_kernels.cu, layernorm_kernels.cu,
                                   pos_encoding_kernels.cu, fused_layernorm_dynamic_per_token_quant.cu (sm_100a)
                                   -> torch.ops._ref_C.*

The reference's build system (CMake + FetchContent) is NOT run; this is direct g++/nvcc on the
few files of the path, driven by a generated ninja file.  No reference source is copied.
  oracle/_ref/libvllm_ref_cutlass.so  quantization/cutlass_w8a8/c3x/scaled_mm_sm100_fp8.cu, moe/grouped_mm_c3x_sm100.cu,
                                   moe/moe_data.cu (CUTLASS headers from flashinfer's vendored tree)
                                   -> torch.ops._ref_cutlass.{cutlass_scaled_mm, cutlass_moe_mm, get_cutlass_moe_mm_data}
Usage: python oracle/build_ref.py [cpu] [cuda] [cutlass] [-j N]
"""
import os, subprocess, sys, sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("VLLM_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(HERE, "_ref")
BLD = os.path.join(OUT, "build")


def main():
    import torch
    targets = [a for a in sys.argv[1:] if a in ("cpu", "cuda", "cutlass")] or ["cpu", "cuda"]
    jobs = "4"
    if "-j" in sys.argv:
        jobs = sys.argv[sys.argv.index("-j") + 1]
    if not os.path.isdir(os.path.join(REF, "csrc")):
        print(f"[build_ref] {REF}/csrc not present; keeping prebuilt files in {OUT}")
        return 0
    os.makedirs(BLD, exist_ok=True)
    tdir = os.path.dirname(torch.__file__)
    pyinc = sysconfig.get_paths()["include"]
    inc = (f"-I{REF}/csrc -isystem {tdir}/include -isystem {tdir}/include/torch/csrc/api/include "
           f"-isystem {pyinc} -isystem /usr/local/cuda/include")
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    cpu_flags = ("-std=c++17 -O3 -fPIC -fopenmp -DVLLM_CPU_EXTENSION -mavx512f -mavx512vl -mavx512bw "
                 f"-mavx512dq -mavx512bf16 -mavx512vnni -mf16c -D_GLIBCXX_USE_CXX11_ABI={abi} "
                 f"-I{REF}/csrc/cpu {inc}")
    cu_flags = ("-std=c++17 -O2 -gencode arch=compute_100a,code=sm_100a -DENABLE_FP8 "
                "--expt-relaxed-constexpr -Xcompiler -fPIC -DTORCH_EXTENSION_NAME=_ref_C "
                f"-D_GLIBCXX_USE_CXX11_ABI={abi} -DUSE_CUDA {inc}")
    libs = f"-L{tdir}/lib -lc10 -ltorch_cpu -ltorch -Wl,-rpath,{tdir}/lib"
    gomp = "/usr/lib/gcc/x86_64-linux-gnu/13/libgomp.so"
    n = ["rule cxx", "  command = /usr/bin/g++ $flags -c $in -o $out", "  description = CXX $out",
         "rule nvcc", "  command = /usr/local/cuda/bin/nvcc $flags -c $in -o $out",
         "  description = NVCC $out",
         "rule link", "  command = /usr/bin/g++ -shared $in -o $out $libs", "  description = LINK $out", ""]
    defaults = []
    if "cpu" in targets:
        srcs = {"ref_bind_cpu": f"{HERE}/ref_bind_cpu.cpp", "cpu_attention": f"{REF}/csrc/cpu/attention.cpp",
                "cpu_cache": f"{REF}/csrc/cpu/cache.cpp"}
        objs = []
        for k, s in srcs.items():
            o = f"{BLD}/{k}.o"
            objs.append(o)
            n += [f"build {o}: cxx {s}", f"  flags = {cpu_flags}"]
        so = f"{OUT}/libvllm_ref_cpu.so"
        n += [f"build {so}: link {' '.join(objs)}", f"  libs = {libs} {gomp}"]
        defaults.append(so)
    if "cuda" in targets:
        cus = {"pa_v1": "attention/paged_attention_v1.cu", "pa_v2": "attention/paged_attention_v2.cu",
               "cache_kernels": "cache_kernels.cu", "fp8_common": "quantization/fp8/common.cu",
               "activation": "activation_kernels.cu", "topk_softmax": "@ref_topk_softmax_compat.cu",
               "moe_align_sum": "moe/moe_align_sum_kernels.cu", "layernorm": "layernorm_kernels.cu",
               "pos_encoding": "pos_encoding_kernels.cu",
               "ln_quant": "quantization/fused_kernels/fused_layernorm_dynamic_per_token_quant.cu"}
        objs = []
        for k, s in cus.items():
            o = f"{BLD}/{k}.o"
            objs.append(o)
            path = f"{HERE}/{s[1:]}" if s.startswith("@") else f"{REF}/csrc/{s}"
            n += [f"build {o}: nvcc {path}", f"  flags = {cu_flags}"]
        o = f"{BLD}/ref_bind_cuda.o"
        objs.append(o)
        n += [f"build {o}: cxx {HERE}/ref_bind_cuda.cpp",
              f"  flags = -std=c++17 -O1 -fPIC -D_GLIBCXX_USE_CXX11_ABI={abi} -DUSE_CUDA {inc}"]
        so = f"{OUT}/libvllm_ref_cuda.so"
        n += [f"build {so}: link {' '.join(objs)}",
              f"  libs = {libs} -ltorch_cuda -lc10_cuda -L/usr/local/cuda/lib64 -lcudart"]
        defaults.append(so)
    if "cutlass" in targets:
        # the reference's Blackwell CUTLASS kernels (cutlass_scaled_mm sm100 fp8, cutlass_moe_mm sm100) against the
        # CUTLASS 4.x header tree vendored inside flashinfer (the pinned v4.0.0 is fetched by CMake, not vendored)
        import flashinfer
        fi = os.path.dirname(flashinfer.__file__)
        cl_flags = (cu_flags.replace("-DTORCH_EXTENSION_NAME=_ref_C", "-DTORCH_EXTENSION_NAME=_ref_cutlass") +
                    " --expt-extended-lambda -DENABLE_SCALED_MM_SM100=1 -DENABLE_CUTLASS_MOE_SM100=1"
                    " -DCUTLASS_ENABLE_DIRECT_CUDA_DRIVER_CALL=1"
                    f" -I{fi}/data/cutlass/include -I{fi}/data/cutlass/tools/util/include")
        cus = {"cl_scaled_mm_sm100_fp8": "quantization/cutlass_w8a8/c3x/scaled_mm_sm100_fp8.cu",
               "cl_grouped_mm_sm100": "quantization/cutlass_w8a8/moe/grouped_mm_c3x_sm100.cu",
               "cl_moe_data": "quantization/cutlass_w8a8/moe/moe_data.cu"}
        objs = []
        for k, s_ in cus.items():
            o = f"{BLD}/{k}.o"
            objs.append(o)
            n += [f"build {o}: nvcc {REF}/csrc/{s_}", f"  flags = {cl_flags}"]
        for k, path in {"cl_common": f"{REF}/csrc/cutlass_extensions/common.cpp",
                        "ref_bind_cutlass": f"{HERE}/ref_bind_cutlass.cpp"}.items():
            o = f"{BLD}/{k}.o"
            objs.append(o)
            n += [f"build {o}: cxx {path}",
                  f"  flags = -std=c++17 -O1 -fPIC -D_GLIBCXX_USE_CXX11_ABI={abi} -DUSE_CUDA {inc} "
                  f"-I{fi}/data/cutlass/include"]
        so = f"{OUT}/libvllm_ref_cutlass.so"
        # CUTLASS calls cuTensorMapEncodeTiled directly: record libcuda.so.1 as needed (stub here, real on the box)
        n += [f"build {so}: link {' '.join(objs)}",
              f"  libs = {libs} -ltorch_cuda -lc10_cuda -L/usr/local/cuda/lib64 -lcudart "
              "-L/usr/local/cuda/lib64/stubs -lcuda"]
        defaults.append(so)
    n.append("default " + " ".join(defaults))
    with open(f"{BLD}/build.ninja", "w") as f:
        f.write("\n".join(n) + "\n")
    return subprocess.call(["ninja", "-C", BLD, "-j", jobs])


if __name__ == "__main__":
    sys.exit(main())
