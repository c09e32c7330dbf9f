"""Run under torchrun (N = 2, 4 or 8 GPUs of one node): the tensor-parallel decode step computes the SAME function as
the single-GPU step.  Every rank builds the TP-N decoder (its shard of the one seeded weight set, decode_step.py
`_dummy_weight`) and, with the TP state switched to world 1, the unsharded TP-1 decoder from the same seed; both start
from an empty paged KV cache and decode the same token ids for STEPS steps (the cache content is produced by the model
itself, so the caches stay comparable), and the logits of every step must agree up to the fp8 activation-rounding
noise of the two runs (see the bound in the loop below).  Runs each model with the NCCL
all-reduce and with the fused peer all-reduce (one-shot and two-shot).
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/multi_gpu/check_tp_parity.py
TP_PARITY_BACKEND=gloo: the N ranks share ONE GPU (cuda:0) and all-reduce through gloo on the host -- no NVLink, no
peer kernel (its spin barriers need the ranks' kernels co-resident), but the sharding, the shard-shaped kernels and
the host logic are the real ones: this is the variant a 1-GPU box can run.
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
STEPS = 12


def run_decoder(dec, cfg, B, bs, nb, dev, vocab_shard_gather):
    from vllm_b200.attention import B200PagedAttentionMetadata
    g = torch.Generator().manual_seed(11)
    bps = (STEPS + bs - 1) // bs + 1
    bt = torch.randperm(nb, generator=g)[:B * bps].to(torch.int32).reshape(B, bps).to(dev)
    ids_all = torch.randint(0, cfg.vocab_size, (STEPS, B), generator=g).to(dev)
    outs = []
    for t in range(STEPS):
        sl = torch.full((B,), t + 1, dtype=torch.int32, device=dev)
        pos = torch.full((B,), t, dtype=torch.int64, device=dev)
        slots = bt[torch.arange(B, device=dev), t // bs].to(torch.int64) * bs + t % bs
        meta = B200PagedAttentionMetadata(seq_lens_tensor=sl, max_decode_seq_len=t + 1, block_tables=bt,
                                          num_decode_tokens=B, slot_mapping=slots)
        outs.append(vocab_shard_gather(dec.forward(ids_all[t], pos, meta).float()))
    return torch.stack(outs)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    one_gpu = os.environ.get("TP_PARITY_BACKEND", "nccl") == "gloo"
    dev = torch.device("cuda", 0 if one_gpu else local)
    torch.cuda.set_device(dev)
    if one_gpu:
        dist.init_process_group("gloo")
    else:
        dist.init_process_group("nccl", device_id=dev)
    from vllm_b200.decode_step import CONFIGS, Decoder
    from vllm_b200.distributed import communication_op as co
    from vllm_b200.distributed import init_peer_allreduce, init_tensor_parallel
    init_tensor_parallel()
    par = None if one_gpu else init_peer_allreduce(dev)
    assert one_gpu or par is not None, "peer all-reduce did not initialise"
    fails = []
    B, bs = 8, 16
    nb = B * 4 + 4

    def gather_vocab(x):   # [T, V / world] per rank -> [T, V]
        x = x.cpu() if one_gpu else x.contiguous()
        parts = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(parts, x)
        return torch.cat(parts, dim=-1).to(dev)

    for name in ("small-llama", "small-mixtral") if world > 2 or one_gpu else ("tiny-llama", "tiny-mixtral", "small-llama", "small-mixtral"):
        cfg = CONFIGS[name]
        for kvd in ("fp8", "auto"):
            # ---- TP-1 reference on this GPU (same seed => same unsharded weights) ----
            saved = (co._TP_GROUP, co._TP_WORLD, co._TP_RANK, co._PEER_AR)
            co._TP_GROUP, co._TP_WORLD, co._TP_RANK, co._PEER_AR = None, 1, 0, None
            try:
                ref = run_decoder(Decoder(cfg, nb, bs, kvd, "fp8", device=dev, seed=5), cfg, B, bs, nb, dev, lambda x: x)
            finally:
                co._TP_GROUP, co._TP_WORLD, co._TP_RANK, co._PEER_AR = saved
            scale = ref.abs().max().item()
            for ar in ("gloo",) if one_gpu else ("nccl", "1shot", "2shot", "split"):
                co._PEER_AR = None if ar in ("nccl", "gloo") else par
                if par is not None:
                    par.algo = ar if ar in ("1shot", "2shot", "split") else "auto"
                got = run_decoder(Decoder(cfg, nb, bs, kvd, "fp8", device=dev, seed=5), cfg, B, bs, nb, dev, gather_vocab)
                err = (got - ref).abs()
                # TP-N is not bit-identical to TP-1 even with a perfect all-reduce: the row-parallel GEMMs (o_proj,
                # down_proj) quantise their activations per token over each rank's K SHARD (the reference does the
                # same: Fp8LinearOp quantises the local input), so the fp8 rounding noise of the two runs is
                # independent.  Its measured size after 2 layers x 12 steps of cache feedback is ~0.7 % of the logit
                # scale (mean) / ~4 % (max); a sharding or all-reduce bug moves the logits by the scale itself.
                # Bound: every (step, sequence) row within 6 % of the scale, mean within 1.2 %; the MoE model may
                # route a borderline token to a different expert (a discontinuity), so 85 % of its rows suffice.
                row_ok = (err.amax(dim=-1) <= 6e-2 * scale)
                need = 1.0 if not cfg.num_experts else 0.85
                ok = row_ok.float().mean().item() >= need and err.mean().item() <= 1.2e-2 * scale
                if rank == 0:
                    print(f"[tp-parity] {name} kv={kvd} tp{world} allreduce={ar}: max|err| {err.max().item():.4f} "
                          f"mean {err.mean().item():.5f} (logit scale {scale:.3f}) rows within bound "
                          f"{row_ok.float().mean().item():.3f} -> {'ok' if ok else 'FAIL'}", flush=True)
                if not ok:
                    fails.append((name, kvd, ar))
            co._PEER_AR = par
            if par is not None:
                par.algo = "auto"
    dist.barrier()
    torch.cuda.synchronize()
    if rank == 0:
        print(f"tp parity check: {'OK' if not fails else 'FAILED ' + str(fails)}", flush=True)
    sys.stdout.flush()
    os._exit(1 if fails else 0)


if __name__ == "__main__":
    main()
