"""One process per GPU; torch.distributed (NCCL over NVLink/NVSwitch on the GPU box, gloo in the CPU tests) is
the plumbing.  Mirrors the call site of /root/reference/vllm/distributed/communication_op.py:12-14 and the
GroupCoordinator.all_reduce contract (parallel_state.py: world size 1 bypasses, the result is the summed tensor).
The all-reduce is issued on the current stream so it is captured inside the decode CUDA graph, as pynccl does in
the reference (device_communicators/cuda_communicator.py:93-118)."""
from typing import Optional

import torch
import torch.distributed as dist

_TP_GROUP: Optional["dist.ProcessGroup"] = None
_TP_WORLD = 1
_TP_RANK = 0
_PEER_AR = None   # PeerAllReduce (NVLink peer-memory all-reduce fused with its consumer) or None
_STAGE_THROUGH_HOST = False   # gloo group carrying CUDA tensors (tests: several ranks sharing one GPU)


def init_tensor_parallel(group: Optional["dist.ProcessGroup"] = None) -> None:
    """Use `group` (default: the world group, if torch.distributed is initialised) as the TP group."""
    global _TP_GROUP, _TP_WORLD, _TP_RANK, _STAGE_THROUGH_HOST
    if group is None and not (dist.is_available() and dist.is_initialized()):
        _TP_GROUP, _TP_WORLD, _TP_RANK, _STAGE_THROUGH_HOST = None, 1, 0, False
        return
    _TP_GROUP = group if group is not None else dist.group.WORLD
    _TP_WORLD = dist.get_world_size(_TP_GROUP)
    _TP_RANK = dist.get_rank(_TP_GROUP)
    _STAGE_THROUGH_HOST = dist.get_backend(_TP_GROUP) == "gloo"


def init_peer_allreduce(device, max_bytes: int = 8 << 20):
    """Set up the peer-memory all-reduce for the TP group (CUDA devices of ONE node, world size 2/4/8, NCCL group
    already initialised).  Returns the PeerAllReduce or None when it does not apply (world 1, unsupported size, or
    VB200_PEER_AR=0)."""
    global _PEER_AR
    import os
    _PEER_AR = None
    if _TP_WORLD not in (2, 4, 8) or torch.device(device).type != "cuda":
        return None
    # the environment is read per process: agree on it before anything collective is built (a rank that skipped the
    # constructor would leave the others hanging in it).  Off if ANY rank says off; the algorithm is rank 0's.
    mine = (os.environ.get("VB200_PEER_AR", "1") != "0", os.environ.get("VB200_PEER_AR_ALGO", "auto"))
    votes = [None] * _TP_WORLD
    dist.all_gather_object(votes, mine, group=_TP_GROUP)
    if all(v[0] for v in votes):
        from .peer_allreduce import PeerAllReduce
        try:
            _PEER_AR = PeerAllReduce(_TP_GROUP, device, max_bytes, algo=votes[0][1])
        except RuntimeError as e:   # raised on EVERY rank together (see PeerAllReduce.__init__): NCCL carries on
            import sys
            print(f"[vllm_b200] peer all-reduce unavailable, using NCCL: {e}", file=sys.stderr, flush=True)
            _PEER_AR = None
    return _PEER_AR


def get_peer_allreduce():
    return _PEER_AR


def get_tensor_model_parallel_world_size() -> int:
    return _TP_WORLD


def get_tensor_model_parallel_rank() -> int:
    return _TP_RANK


def tensor_model_parallel_all_reduce(input_: torch.Tensor) -> torch.Tensor:
    """All-reduce (sum) across the tensor-parallel group; in place, returns its argument."""
    if _TP_WORLD == 1:
        return input_
    if _STAGE_THROUGH_HOST and input_.is_cuda:
        # test plumbing only (gloo ranks sharing one GPU): fp32 sum on the host, rounded once
        host = input_.float().cpu()
        dist.all_reduce(host, op=dist.ReduceOp.SUM, group=_TP_GROUP)
        input_.copy_(host)
        return input_
    dist.all_reduce(input_, op=dist.ReduceOp.SUM, group=_TP_GROUP)
    return input_
