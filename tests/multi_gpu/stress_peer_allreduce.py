"""Run under torchrun (>= 2 GPUs of one node): stress run of the NVLink peer-memory all-reduce kernels
(csrc/peer_allreduce.cu), modelled on the reference's /root/reference/csrc/custom_all_reduce_test.cu (many back-to-back
calls with changing data, checked against the exact answer) and tests/distributed/test_custom_all_reduce.py (eager and
graph).  What it exercises that the parity check does not:
  * >= 10 000 back-to-back all-reduces WITHOUT a host sync or an NCCL call between them (nothing re-aligns the ranks),
  * random per-rank device-side delays (torch.cuda._sleep) before random calls, so ranks enter the flag barriers skewed
    by up to ~1 ms in either direction,
  * message sizes and kernels (one-shot / two-shot, plain / fused-norm / fused-norm-quant) alternating call by call,
  * the same mix captured in a CUDA graph and replayed with the inputs changing between replays.
Every result is compared ON THE DEVICE with the exact expected value (inputs are small integers: sums are exact in
bf16, so any stale or torn read shows up as a mismatch); one counter is read back at the end.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/multi_gpu/stress_peer_allreduce.py
"""
import os
import random
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
ITERS = int(os.environ.get("PEER_AR_STRESS_ITERS", "10000"))
REPLAYS = int(os.environ.get("PEER_AR_STRESS_REPLAYS", "400"))
SHAPES = [(256, 4096), (1, 4096), (64, 8192), (300, 4096), (17, 1024), (512, 4096)]


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    from vllm_b200 import _custom_ops as ops
    from vllm_b200.distributed import init_peer_allreduce, init_tensor_parallel
    init_tensor_parallel()
    par = init_peer_allreduce(dev)
    assert par is not None, "peer all-reduce did not initialise"
    dt = torch.bfloat16
    rng_delay = random.Random(1234 + rank)     # per-rank: who sleeps, and for how long, differs between ranks
    rng_plan = random.Random(99)               # shared: the sequence of shapes / kernels is the same everywhere
    bad = torch.zeros((), dtype=torch.int64, device=dev)
    base = {s: (torch.arange(s[0] * s[1], device=dev) % 7).to(dt).view(s) for s in SHAPES}
    w = {h: torch.ones(h, device=dev, dtype=dt) for h in {s[1] for s in SHAPES}}

    def one(shape, algo, mode, it):
        """x_r = base + (it % 5) + r  =>  sum = world * (base + it % 5) + world (world - 1) / 2 (exact in bf16)."""
        T, H = shape
        par.algo = algo
        st = par.staging(T, H, dt)
        torch.add(base[shape], float(it % 5 + rank), out=st)
        want = base[shape] * world + float(world * (it % 5) + world * (world - 1) // 2)
        if mode == 0:
            got = par.all_reduce(st)
            bad.add_((got != want).any().to(torch.int64))
        else:
            res = torch.full((T, H), 2.0, device=dev, dtype=dt)
            res_want = res.clone()
            h = want.clone()
            if mode == 1:
                o = par.all_reduce_add_rms_norm(st, res, w[H], 1e-5)
                ops.fused_add_rms_norm(h, res_want, w[H], 1e-5)
                bad.add_(((o != h).float().mean() > 1e-3).to(torch.int64))
            else:
                q, s = par.all_reduce_add_rms_norm_quant(st, res, w[H], 1e-5)
                q2, s2 = ops.rms_norm_dynamic_per_token_quant(h, w[H], 1e-5, torch.float8_e4m3fn, None, res_want)
                bad.add_(((q.view(torch.uint8) != q2.view(torch.uint8)).float().mean() > 5e-3).to(torch.int64))
            bad.add_((res != res_want).any().to(torch.int64))     # the residual stream is exact

    # ---- eager: ITERS calls, no host sync in between ------------------------------------------------------------
    t0 = time.perf_counter()
    for it in range(ITERS):
        shape = SHAPES[rng_plan.randrange(len(SHAPES))]
        algo = ("1shot", "2shot", "split")[rng_plan.randrange(3)]
        mode = (0, 0, 0, 1, 2)[rng_plan.randrange(5)]
        if rng_delay.random() < 0.03:
            torch.cuda._sleep(rng_delay.randrange(1, 2_000_000))   # up to ~1 ms at 1.9 GHz
        one(shape, algo, mode, it)
    torch.cuda.synchronize()
    eager_s = time.perf_counter() - t0
    eager_bad = int(bad.item())
    # ---- graph: 8 mixed calls per replay, REPLAYS replays, a replicated scalar input changes between replays --------
    bad.zero_()
    shift = torch.zeros((), device=dev, dtype=dt)
    plan = [(SHAPES[rng_plan.randrange(len(SHAPES))], ("1shot", "2shot", "split")[i % 3]) for i in range(8)]

    def body():
        for shape, algo in plan:
            par.algo = algo
            st = par.staging(shape[0], shape[1], dt)
            torch.add(base[shape] + float(rank), shift, out=st)
            got = par.all_reduce(st)
            want = base[shape] * world + shift * world + float(world * (world - 1) // 2)
            bad.add_((got != want).any().to(torch.int64))
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        body()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            body()
    torch.cuda.synchronize()
    for it in range(REPLAYS):
        shift.fill_(float(it % 4))
        if rng_delay.random() < 0.1:
            torch.cuda._sleep(rng_delay.randrange(1, 2_000_000))
        g.replay()
    torch.cuda.synchronize()
    graph_bad = int(bad.item())
    par.algo = "auto"
    tot = torch.tensor([eager_bad, graph_bad], device=dev)
    dist.all_reduce(tot)
    ok = int(tot.sum().item()) == 0
    if rank == 0:
        print(f"[stress] world {world}: {ITERS} eager calls in {eager_s:.1f} s (mismatching calls: {int(tot[0])}), "
              f"{REPLAYS} graph replays x {len(plan)} calls (mismatching calls: {int(tot[1])})", flush=True)
        print(f"peer all-reduce stress: {'OK' if ok else 'FAILED'}", flush=True)
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0 if ok else 1)


if __name__ == "__main__":
    main()
