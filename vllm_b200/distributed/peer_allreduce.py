"""Row-parallel all-reduce over NVLink peer memory, fused with the residual add + RMSNorm (+ per-token fp8 quant)
that consumes it (kernel: csrc/peer_allreduce.cu).

Stands where the reference's `CustomAllreduce` stands
(/root/reference/vllm/distributed/device_communicators/custom_all_reduce.py:49-300: IPC-registered buffers, flag
barriers, `should_custom_ar` size/world-size gate) together with the all-reduce + RMSNorm fusion pass
(/root/reference/vllm/compilation/collective_fusion.py).  One process per GPU: every rank allocates a staging buffer
and a signal block (cudaMalloc through the C ABI), exports them through CUDA IPC, all-gathers the 64-byte handles over
the existing process group, and opens the peers' handles on its own device (lazy NVLink peer access).  The row-parallel GEMM writes its partial sum
straight into the staging buffer (`staging(rows, hidden)` is the GEMM's `out`), then ONE kernel does barrier ->
pull + sum -> residual add -> RMSNorm (-> quant) -> barrier.  Everything is issued on the current stream and is
CUDA-graph capturable (the barrier counters live in device memory).
"""
import ctypes
import os
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from .._lib import check, dtype_enum, load, ptr, stream

_SUPPORTED_WORLD_SIZES = (2, 4, 8)


class _RawCudaBuffer:
    """A raw device allocation seen through __cuda_array_interface__ (uint8, 1-D) so torch can wrap it."""

    def __init__(self, ptr_: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr_, False), "version": 3,
                                         "strides": None}


class PeerAllReduce:
    def __init__(self, group: Optional["dist.ProcessGroup"], device: torch.device, max_bytes: int = 8 << 20,
                 algo: Optional[str] = None):
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        self.device = torch.device(device)
        self.max_bytes = int(max_bytes)
        if self.world not in _SUPPORTED_WORLD_SIZES:
            raise RuntimeError(f"PeerAllReduce: world size {self.world} not in {_SUPPORTED_WORLD_SIZES}")
        lib = load()
        sig_bytes = int(lib.vb200_peer_allreduce_signal_bytes())
        # one allocation per rank: [signal block | staging buffer]; zeroed once (the counters start at 0 everywhere)
        self._sig_bytes = (sig_bytes + 255) // 256 * 256
        # Set-up is collective: every step that can fail locally (allocation, IPC export, opening a peer's handle)
        # is followed by an agreement step, so either every rank ends up with a working object or every rank raises
        # (and the caller falls back to NCCL on all ranks together).
        payload, err = None, None
        self._own_ptr, self._peer_ptrs = None, []
        # layout of the allocation: [signal block | staging (partial sums) | gather (two-shot: reduced shards; split:
        # finished rows + residual rows + scales)]
        self._gather_bytes = 2 * self.max_bytes + (64 << 10)
        total = self._sig_bytes + self.max_bytes + self._gather_bytes
        try:
            with torch.cuda.device(self.device):
                p, h = ctypes.c_void_p(), ctypes.create_string_buffer(64)
                check(lib.vb200_peer_buffer_alloc(total, ctypes.byref(p), h))
            self._own_ptr = p.value
            payload = bytes(h.raw)
        except Exception as e:  # noqa: BLE001
            err = f"allocation / IPC export failed: {type(e).__name__}: {e}"
        handles: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(handles, payload, group=self.group)
        bases = []
        if err is None and all(h is not None for h in handles):
            try:
                with torch.cuda.device(self.device):
                    for r, h in enumerate(handles):
                        if r == self.rank:
                            bases.append(self._own_ptr)
                            continue
                        q = ctypes.c_void_p()
                        check(lib.vb200_peer_buffer_open(ctypes.create_string_buffer(h, 64), ctypes.byref(q)))
                        self._peer_ptrs.append(q.value)
                        bases.append(q.value)
            except Exception as e:  # noqa: BLE001
                err = f"opening a peer handle failed: {type(e).__name__}: {e}"
        elif err is None:
            err = "a peer could not export its buffer"
        ok = torch.tensor([0 if err else 1], dtype=torch.int32, device=self.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)   # also: nobody launches before all mappings exist
        if int(ok.item()) == 0:
            raise RuntimeError(f"PeerAllReduce set-up failed on some rank ({err or 'peer failure'})")
        self._sig_ptrs = (ctypes.c_void_p * self.world)(*bases)
        self._data_ptrs = (ctypes.c_void_p * self.world)(*[b + self._sig_bytes for b in bases])
        self._gather_ptrs = (ctypes.c_void_p * self.world)(*[b + self._sig_bytes + self.max_bytes for b in bases])
        # "auto": the reference's switch (custom_all_reduce.cuh:560-572): one-shot for 2 ranks and small messages,
        # two-shot (reduce-scatter + all-gather) from 512 KB at 4 ranks / 256 KB at 8 ranks.  VB200_PEER_AR_ALGO / `algo` pins it (same on every rank);
        # the choice depends only on (world, bytes), so every rank takes the same kernel.
        self.algo = algo if algo is not None else os.environ.get("VB200_PEER_AR_ALGO", "auto")
        if self.algo not in ("auto", "1shot", "2shot", "split"):
            raise RuntimeError(f"VB200_PEER_AR_ALGO={self.algo}: expected auto, 1shot, 2shot or split")
        # this rank's staging area as a torch tensor over the raw allocation (the GEMM writes into it)
        self._own = torch.as_tensor(_RawCudaBuffer(self._own_ptr + self._sig_bytes, self.max_bytes), device=self.device)

    # ---- the GEMM's output buffer ------------------------------------------------------------------------------
    def staging(self, rows: int, hidden: int, dtype: torch.dtype) -> torch.Tensor:
        """[rows, hidden] view of this rank's peer-visible buffer: pass it as the row-parallel GEMM's `out`."""
        nbytes = rows * hidden * torch.empty((), dtype=dtype).element_size()
        if nbytes > self.max_bytes:
            raise RuntimeError(f"PeerAllReduce: {nbytes} bytes exceed the staging buffer ({self.max_bytes})")
        return self._own[:nbytes].view(dtype).view(rows, hidden)

    def fits(self, rows: int, hidden: int, dtype: torch.dtype) -> bool:
        """`should_custom_ar` (custom_all_reduce.py:228-243): size gate; hidden % 8 for the 16-byte loads."""
        return hidden % 8 == 0 and rows * hidden * torch.empty((), dtype=dtype).element_size() <= self.max_bytes

    def _check_staged(self, x: torch.Tensor) -> Tuple[int, int]:
        if x.data_ptr() != self._own.data_ptr() or not x.is_contiguous() or x.dim() != 2:
            raise RuntimeError("PeerAllReduce: the input must be the tensor returned by staging()")
        return x.shape[0], x.shape[1]

    def pick(self, rows: int, hidden: int, dtype: torch.dtype) -> str:
        """Which kernel(s) a [rows, hidden] message takes: "1shot", "2shot" or "split" (same answer on every rank: it
        depends only on world size and shape).  auto = what the 8-GPU measurements say
        (profiles/r02_multi_gpu_tp8_s44.txt, r02_peer_allreduce_split_*.txt): one-shot at 2 and 4 ranks and for small
        messages; at 8 ranks from 256 KB on, the split form (reduce-scatter + consumer | all-gather, two launches).  The
        single-kernel two-shot never won (its fenced middle barrier) and is kept for comparison only."""
        nbytes = rows * hidden * torch.empty((), dtype=dtype).element_size()
        split_ok = hidden % 16 == 0 and rows >= self.world and 2 * nbytes + 4 * rows <= self._gather_bytes
        if self.algo == "split":
            return "split" if split_ok else "1shot"
        if self.algo != "auto":
            return self.algo
        # measured per fused op at [256 x 4096] bf16 (us, one-shot / split): 2 ranks 14.2 / 18.3, 4 ranks 19.6 / 23.9,
        # 8 ranks 33.3 / 24.6 -- the split form pays a second launch + barrier and wins once (N-1) copies of the
        # tensor per GPU outweigh them
        if self.world >= 8 and nbytes >= 256 * 1024 and split_ok:
            return "split"
        return "1shot"

    def two_shot(self, nbytes: int) -> bool:
        return self.algo == "2shot"

    def algo_name(self, rows: int, hidden: int, dtype: torch.dtype) -> str:
        return "peer-" + self.pick(rows, hidden, dtype)

    def _call(self, mode, out, rows, hidden, dtype, residual, weight, eps, scales, scale_ub):
        lib = load()
        algo = self.pick(rows, hidden, dtype)
        if algo == "split":
            check(lib.vb200_peer_allreduce_split(mode, ptr(out), self._data_ptrs, self._gather_ptrs, self._sig_ptrs,
                                                 self.rank, self.world, rows, hidden, dtype_enum(dtype), ptr(residual),
                                                 ptr(weight), float(eps), ptr(scales), ptr(scale_ub), stream()))
            return
        if algo == "2shot":
            check(lib.vb200_peer_allreduce_2shot(mode, ptr(out), self._data_ptrs, self._gather_ptrs, self._sig_ptrs,
                                                 self.rank, self.world, rows, hidden, dtype_enum(dtype), ptr(residual),
                                                 ptr(weight), float(eps), ptr(scales), ptr(scale_ub), stream()))
            return
        check(lib.vb200_peer_allreduce(mode, ptr(out), self._data_ptrs, self._sig_ptrs, self.rank, self.world, rows,
                                       hidden, dtype_enum(dtype), ptr(residual), ptr(weight), float(eps),
                                       ptr(scales), ptr(scale_ub), stream()))

    # ---- the three fused forms ---------------------------------------------------------------------------------
    def all_reduce(self, staged: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """out = sum over ranks of their staged tensors (tensor_model_parallel_all_reduce semantics)."""
        rows, hidden = self._check_staged(staged)
        out = torch.empty(rows, hidden, dtype=staged.dtype, device=staged.device) if out is None else out
        self._call(0, out, rows, hidden, staged.dtype, None, None, 0.0, None, None)
        return out

    def all_reduce_add_rms_norm(self, staged: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                                epsilon: float) -> torch.Tensor:
        """all_reduce -> fused_add_rms_norm(x, residual, weight, eps): residual updated in place, returns the normed
        activations (csrc/layernorm_kernels.cu:55-160 arithmetic)."""
        rows, hidden = self._check_staged(staged)
        out = torch.empty(rows, hidden, dtype=staged.dtype, device=staged.device)
        self._call(1, out, rows, hidden, staged.dtype, residual, weight, epsilon, None, None)
        return out

    def all_reduce_add_rms_norm_quant(self, staged: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                                      epsilon: float, scale_ub: Optional[torch.Tensor] = None
                                      ) -> Tuple[torch.Tensor, torch.Tensor]:
        """all_reduce -> rms_norm_dynamic_per_token_quant(x, weight, eps, fp8, scale_ub, residual): returns
        (fp8 [rows, hidden], scales [rows, 1])."""
        rows, hidden = self._check_staged(staged)
        out = torch.empty(rows, hidden, dtype=torch.float8_e4m3fn, device=staged.device)
        scales = torch.empty(rows, 1, dtype=torch.float32, device=staged.device)
        self._call(2, out, rows, hidden, staged.dtype, residual, weight, epsilon, scales, scale_ub)
        return out, scales
