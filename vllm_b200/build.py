#!/usr/bin/env python3
"""Builds vllm_b200/lib/libvllm_b200.so (C ABI, include/vllm_b200.h) with nvcc for sm_100a.

One nvcc invocation per .cu (parallel), objects cached under vllm_b200/lib/obj keyed by source
mtime.  No torch dependency: the library links only the (static) CUDA runtime; the driver entry
point needed for TMA descriptors is resolved at run time through cudaGetDriverEntryPoint.
"""
import concurrent.futures as cf
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "lib")
OBJ = os.path.join(LIB, "obj")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
         "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v"]


def _stale(out, deps):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(d) > t for d in deps)


def _compile(src, obj, hdrs, verbose):
    if not _stale(obj, [src] + hdrs):
        return obj, ""
    cmd = [NVCC] + FLAGS + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    with open(obj + ".ptxas.log", "w") as f:
        f.write(r.stderr)
    return obj, r.stderr if verbose else ""


def build(verbose=False, jobs=None):
    os.makedirs(OBJ, exist_ok=True)
    srcs = sorted(f for f in os.listdir(SRC) if f.endswith(".cu"))
    hdrs = [os.path.join(SRC, f) for f in os.listdir(SRC) if f.endswith((".cuh", ".h"))]
    hdrs.append(os.path.join(HERE, "..", "include", "vllm_b200.h"))
    objs = []
    with cf.ThreadPoolExecutor(max_workers=jobs or min(8, os.cpu_count() or 4)) as ex:
        futs = [ex.submit(_compile, os.path.join(SRC, s), os.path.join(OBJ, s[:-3] + ".o"), hdrs, verbose)
                for s in srcs]
        for f in futs:
            o, log = f.result()
            objs.append(o)
            if log:
                print(log)
    so = os.path.join(LIB, "libvllm_b200.so")
    if _stale(so, objs):
        cmd = [NVCC, "-shared", "-o", so] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                     "-cudart", "static", "-Xcompiler", "-fPIC"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return so


def build_torch_shim(names=("_b200_C", "_b200_C_cache_ops", "_b200_moe_C"), out_name="libvllm_b200_torch.so"):
    """Compiles vllm_b200/torch_shim.cpp (TORCH_LIBRARY registrations under the reference's schemas, no kernels)
    against the installed torch and links it to libvllm_b200.so.  `names` are the three library names
    (pass ("_C", "_C_cache_ops", "_moe_C") for a drop-in build)."""
    import sysconfig
    import torch
    tdir = os.path.dirname(torch.__file__)
    src = os.path.join(HERE, "torch_shim.cpp")
    so = os.path.join(LIB, out_name)
    core = build()
    if not _stale(so, [src, core, os.path.join(HERE, "..", "include", "vllm_b200.h")]):
        return so
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [cxx, "-std=c++17", "-O1", "-fPIC", "-shared", f"-D_GLIBCXX_USE_CXX11_ABI={abi}", "-DUSE_CUDA",
           f"-DVB200_LIB_C={names[0]}", f"-DVB200_LIB_CACHE={names[1]}", f"-DVB200_LIB_MOE={names[2]}",
           src, "-o", so, "-isystem", f"{tdir}/include", "-isystem", f"{tdir}/include/torch/csrc/api/include",
           "-isystem", "/usr/local/cuda/include", "-isystem", sysconfig.get_paths()["include"],
           f"-L{tdir}/lib", "-lc10", "-ltorch_cpu", "-ltorch", "-ltorch_cuda", "-lc10_cuda", f"-L{LIB}", "-lvllm_b200",
           "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{tdir}/lib"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"torch shim build failed:\n{r.stdout}\n{r.stderr}")
    return so


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
    if "--shim" in sys.argv:
        print(build_torch_shim())
