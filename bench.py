#!/usr/bin/env python3
"""bench.py -- decode throughput of the hot path on BASELINE.json's metric
("decode tokens/s/GPU Llama-3.1-8B FP8; paged-attn HBM GB/s vs peak").

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--config cfg3|cfg2|cfg4|cfg5|tiny]

One "step" = one decode step (all 32 layers + lm_head + greedy arg-max) of Llama-3.1-8B, FP8 w8a8 scaled_mm
(dynamic per-token activation scales, per-channel weight scales) + fp8_e4m3 paged KV cache (block 16), batch 256
at 4k context (BASELINE configs[2], the configuration the metric is quoted on), synthetic token ids, dummy
weights, random cache contents.  N > 1 = tensor parallel over N ranks (heads / columns sharded, NCCL all-reduce
after the row-parallel GEMMs), same global batch => "strong" scaling.  Printed keys: see the task contract;
  value     tokens/s with the step's inputs resident in HBM (CUDA-graph replay, CUDA-event timed, max over ranks)
  e2e       tokens/s through the public API with HOST inputs: per step pinned-host -> device copies of
            token ids / positions / slot mapping / seq lens / block tables, the step, device -> host copy of the
            sampled token ids, host waits for them (the next step's input depends on them)
  roofline  the dominant kernel (paged_attention_kernel): algorithmic bytes / its CUDA-event duration measured
            live in an eager replay of the same step right after the timed region, vs MEASURED_PEAKS.json
  cpu_baseline  the reference's own CPU kernels (oracle/_ref: csrc/cpu attention + cache) + oneDNN bf16 GEMMs
            on a bounded sample of the same step, timed on this box's host cores (N=1, rank 0 only)
`--impl reference` times that CPU implementation as the measured arm.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "decode tokens/s Llama-3.1-8B FP8 (w8a8 scaled_mm + fp8_e4m3 paged KV, batch 256 @ 4k ctx)"

WORKLOADS = {
    # name: (model, batch, ctx, quant, kv_cache_dtype)
    "cfg3": ("Llama-3.1-8B", 256, 4096, "fp8", "fp8"),
    "cfg2": ("Llama-3.1-8B", 64, 2048, "bf16", "auto"),
    "cfg4": ("Mixtral-8x7B", 128, 2048, "fp8", "fp8"),   # BASELINE configs[3] (context not named there: 2k)
    "cfg5": ("Llama-3.1-70B", 128, 8192, "fp8", "fp8"),  # BASELINE configs[4]: needs TP >= 4 (242 GB of weights + KV)
    "tiny": ("tiny-llama", 8, 256, "fp8", "fp8"),
}


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def kernel_traffic(config, world):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel from the committed
    `ncu --set full` capture of this command (profiles/r02_kernel_traffic.json); None when no capture exists for
    the configuration (ncu cannot run inside a timed bench)."""
    if world != 1:
        return None
    for name in ("r02_kernel_traffic.json", "r01_kernel_traffic.json"):   # newest capture first
        try:
            return json.load(open(os.path.join(ROOT, "profiles", name)))[config]["traffic"]
        except Exception:
            continue
    return None


# ---------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------
# CPU arm: the reference's own CPU kernels on a bounded sample of the same step
# ---------------------------------------------------------------------------------------------------
def cpu_reference_step_sample(workload: str, budget_s: float = 12.0, max_reps: int = 8):
    """Times ONE decoder layer of the workload on the host cores for a sample of the batch:
      reshape_and_cache + paged_attention_v1 = the reference's csrc/cpu/{cache,attention}.cpp compiled into
      oracle/_ref/libvllm_ref_cpu.so (kind "reference"; bf16 KV -- that kernel has no fp8 path), or the oracle's
      torch restatement when the .so is absent (kind "port");
      the four projections = torch bf16 matmul (oneDNN / MKL-DNN inside torch: the reference's CPU fp8 GEMM does
      not exist, its CPU w8a8 path is int8 oneDNN).
    tokens/s = sample_seqs / (num_layers * t_layer).  Returns (tokens_per_s, info dict, seconds per rep)."""
    from vllm_b200.decode_step import CONFIGS
    model, batch, ctx, _, _ = WORKLOADS[workload]
    cfg = CONFIGS[model]
    S = min(batch, 64)
    bs, D, Hkv, H = 16, cfg.head_dim, cfg.num_kv_heads, cfg.num_heads
    ref, kind = None, "port"
    so = os.path.join(ROOT, "oracle", "_ref", "libvllm_ref_cpu.so")
    if os.path.exists(so):
        try:
            torch.ops.load_library(so)
            ref, kind = torch.ops._ref_cpu, "reference"
        except Exception:
            ref = None
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    dt = torch.bfloat16
    blocks_per_seq = (ctx + bs - 1) // bs
    nb = S * blocks_per_seq
    g = torch.Generator().manual_seed(0)
    tile = torch.empty(1 << 22, dtype=dt).uniform_(-1, 1, generator=g)          # 8 MiB of random bf16, tiled
    def big(shape):
        n = 1
        for s in shape:
            n *= s
        return tile.repeat((n + tile.numel() - 1) // tile.numel())[:n].reshape(shape).contiguous()
    kc = big((nb, Hkv, D // 8, bs, 8))
    vc = big((nb, Hkv, D, bs))
    bt = torch.randperm(nb, generator=g).to(torch.int32).reshape(S, blocks_per_seq)
    sl = torch.full((S,), ctx, dtype=torch.int32)
    slots = (bt[:, (ctx - 1) // bs].to(torch.int64) * bs + (ctx - 1) % bs)
    hid, inter = cfg.hidden_size, cfg.intermediate_size
    w_qkv = big(((H + 2 * Hkv) * D, hid)) * 0.02
    w_o = big((hid, H * D)) * 0.02
    w_gu = big((2 * inter, hid)) * 0.02
    w_dn = big((hid, inter)) * 0.02
    x = torch.randn(S, hid, generator=g).to(dt)
    one = torch.tensor(1.0)
    scale = D ** -0.5

    def layer():
        qkv = x @ w_qkv.t()
        q = qkv[:, :H * D].reshape(S, H, D).contiguous()
        k = qkv[:, H * D:(H + Hkv) * D].reshape(S, Hkv, D).contiguous()
        v = qkv[:, (H + Hkv) * D:].reshape(S, Hkv, D).contiguous()
        out = torch.empty_like(q)
        if ref is not None:
            ref.reshape_and_cache(k, v, kc, vc, slots, "auto", one, one)
            ref.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, bs, ctx, None, "auto", one, one, 0, 0, 0, 64, 0)
        else:
            from oracle import attention as oattn
            from oracle import cache as ocache
            ocache.reshape_and_cache(k, v, kc, vc, slots)
            out = oattn.paged_attention(q, kc, vc, Hkv, scale, bt, sl)
        h = out.reshape(S, H * D) @ w_o.t()
        gu = h @ w_gu.t()
        act = torch.nn.functional.silu(gu[:, :inter]) * gu[:, inter:]
        return act @ w_dn.t()

    layer()  # warm-up (page-in, thread pool)
    ts = []
    t_start = time.perf_counter()
    while len(ts) < max_reps and (time.perf_counter() - t_start) < budget_s:
        t0 = time.perf_counter()
        layer()
        ts.append(time.perf_counter() - t0)
    t_layer = statistics.median(ts)
    tps = S / (cfg.num_layers * t_layer)
    info = {"value": round(tps, 3), "unit": "tokens/s", "cores": cores, "kind": kind,
            "sample": (f"PARTIAL LAYER, extrapolated: 1 of {cfg.num_layers} layers x {S} of {batch} sequences at ctx "
                       f"{ctx}: reference csrc/cpu reshape_and_cache + paged_attention_v1 (bf16 KV; the CPU kernel has "
                       f"no fp8 path) + the 4 projections as torch bf16 matmul (oneDNN) + silu-mul; norms, RoPE, "
                       f"activation quant, lm_head and sampling are NOT in the sample (they would only slow it "
                       f"further); {len(ts)} reps, median {t_layer * 1e3:.1f} ms/layer; tokens/s = {S} / "
                       f"({cfg.num_layers} x t_layer)"),
            "sample_wall_s": round(sum(ts), 2)}
    return tps, info, t_layer


def run_reference_arm(args, rank: int):
    """bench.py --impl reference: the CPU implementation is the measured arm (rank 0 only)."""
    if rank != 0:
        return
    model, batch, ctx, quant, kvd = WORKLOADS[args.config]
    for _ in range(max(args.warmup, 0) and 1):
        cpu_reference_step_sample(args.config, budget_s=2.0, max_reps=1)
    vals, t_layers, walls = [], [], []
    t_run = time.perf_counter()
    for _ in range(max(args.steps, 1)):
        t_s = time.perf_counter()
        tps, info, t_layer = cpu_reference_step_sample(args.config, budget_s=6.0, max_reps=3)
        walls.append(time.perf_counter() - t_s)
        vals.append(tps)
        t_layers.append(t_layer)
        if time.perf_counter() - t_run > 150:   # keep the whole run within a few minutes
            break
    v = statistics.median(vals)
    from vllm_b200.decode_step import CONFIGS
    nl = CONFIGS[model].num_layers
    info["value"] = round(v, 3)
    # ms_per_step is what one step of THIS arm really took (a bounded sample: set-up + up to 3 reps of one layer on
    # 64 sequences), so steps x ms_per_step is the arm's wall time; the full-step figure behind `value` is separate
    line = {"impl": "reference", "metric": METRIC if args.config == "cfg3" else f"decode tokens/s ({args.config})",
            "value": round(v, 3), "unit": "tokens/s", "n_gpus": args.gpus,
            "steps": len(vals), "warmup": args.warmup,
            "ms_per_step": round(statistics.mean(walls) * 1e3, 1),
            "extrapolated_full_step_ms": round(statistics.median(t_layers) * nl * batch / min(batch, 64) * 1e3, 1),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16 (CPU reference path)",
            "data": "synthetic", "config": workload_config(args.config, 1),
            "cpu_baseline": info,
            "e2e": {"value": round(v, 3), "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(name: str, n_gpus: int):
    model, batch, ctx, quant, kvd = WORKLOADS[name]
    return {"workload": f"BASELINE configs[{ {'cfg3': 2, 'cfg2': 1, 'cfg4': 3, 'cfg5': 4}.get(name, 0) }]: {model} {quant} "
                        f"{'w8a8 scaled_mm ' if quant == 'fp8' else ''}+ {'fp8_e4m3' if kvd == 'fp8' else 'bf16'} "
                        f"paged KV block 16, batch {batch} decode at {ctx} context",
            "global_batch": batch, "context_len": ctx, "block_size": 16, "parallelism": f"tp{n_gpus}",
            "weights": "dummy (random)", "l2": "inputs larger than L2 (KV cache + weights >> 126 MB per step)",
            "context": "advances on the device: every step writes its token's K/V, attends over all tokens so far and "
                       "advance_step_flashattn (inside the captured graph) moves ids / positions / seq_lens / slots on; "
                       f"{ctx} tokens at the first step"}


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
class StepInputs:
    """Host (pinned) and device copies of the decode inputs.  The context ADVANCES: every step appends one token per
    sequence (device side: advance_step_flashattn inside the captured graph; host side: `set_host_context` for the e2e
    leg, which uploads the step's inputs), starting at `ctx` tokens; block tables hold `headroom` extra tokens.

    All five inputs live in ONE packed buffer (ids | positions | slots | seq_lens | block tables) so that a step's
    upload is one pinned-host -> device copy; the host side is double-buffered so that the next step's positions /
    seq_lens / slots can be written while the device still runs the current step."""

    def __init__(self, batch, ctx, block_size, num_blocks, vocab, device, headroom):
        g = torch.Generator().manual_seed(0)
        bps = (ctx + headroom + block_size - 1) // block_size
        self.bs, self.bps = block_size, bps
        self.batch, self.ctx, self.max_len = batch, ctx, ctx + headroom
        o_ids, o_pos, o_slots = 0, 8 * batch, 16 * batch
        o_sl, o_bt = 24 * batch, 28 * batch
        nbytes = o_bt + 4 * batch * bps

        def views(buf):
            return (buf[o_ids:o_pos].view(torch.int64), buf[o_pos:o_slots].view(torch.int64),
                    buf[o_slots:o_sl].view(torch.int64), buf[o_sl:o_bt].view(torch.int32),
                    buf[o_bt:nbytes].view(torch.int32).view(batch, bps))
        self._host = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(2)]
        self._hviews = [views(h) for h in self._host]
        self._cur_buf = 0
        ids = torch.randint(0, vocab, (batch,), generator=g, dtype=torch.int64)
        bt = torch.randperm(num_blocks, generator=g)[:batch * bps].to(torch.int32).reshape(batch, bps)
        for v in self._hviews:
            v[0].copy_(ids)
            v[4].copy_(bt)
        self.bt64 = bt.to(torch.int64)
        self.h_out = torch.empty(batch, dtype=torch.int64).pin_memory()
        self.set_host_context(ctx)
        self._dev = torch.empty(nbytes, dtype=torch.uint8, device=device)
        self.d_ids, self.d_pos, self.d_slots, self.d_sl, self.d_bt = views(self._dev)
        self.upload()
        self.h2d_bytes = nbytes
        self.d2h_bytes = self.h_out.numel() * self.h_out.element_size()

    @property
    def h_ids(self):
        return self._hviews[self._cur_buf][0]

    def flip(self):
        """Switch to the other host buffer (the previous one may still be in flight to the device)."""
        self._cur_buf ^= 1

    def set_host_context(self, seq_len):
        """Host copies for a step whose sequences hold `seq_len` tokens (the last one is the token being decoded)."""
        assert seq_len <= self.max_len, "bench context outgrew the block tables"
        self.cur = seq_len
        _, pos, slots, sl, _ = self._hviews[self._cur_buf]
        pos.fill_(seq_len - 1)
        sl.fill_(seq_len)
        torch.add(self.bt64[:, (seq_len - 1) // self.bs] * self.bs, (seq_len - 1) % self.bs, out=slots)

    def upload(self):
        self._dev.copy_(self._host[self._cur_buf], non_blocking=True)


def attention_algorithmic_bytes(batch, ctx, H, Hkv, D, cache_bytes, q_bytes=2, block_size=16):
    """K and V read once + q in + out + block table + seq_lens."""
    kv = batch * ctx * Hkv * D * 2 * cache_bytes
    qo = 2 * batch * H * D * q_bytes
    bt = batch * ((ctx + block_size - 1) // block_size) * 4 + 4 * batch
    return kv + qo + bt


def gemm_rooflines(dec, batch, hbm_peak, hbm_src):
    """Second roofline block: the four FP8 scaled_mm GEMMs of one layer at this batch, each timed live with CUDA events
    over one back-to-back pass through ALL layers' weights (every launch streams a different weight matrix from HBM:
    layers x bytes >> L2), per-launch average.  Two roofs per GEMM: HBM (weights + activations + output bytes) against
    the measured copy bandwidth, tensor (2 M N K) against 2 x the measured dense bf16 peak (MEASURED_PEAKS.json has no
    fp8 entry; cuBLASLt fp8 8192^3 measured 3.14 PF on this pool, profiles/r01_fp8_peak_probe_s30.json)."""
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        tensor_peak, tensor_src = 2.0 * float(pk["bf16_tflops"]), "2 x measured dense bf16 burst (MEASURED_PEAKS.json)"
    except Exception:
        tensor_peak, tensor_src = 4500.0, "fallback (nominal fp8 dense)"
    from vllm_b200 import _custom_ops as ops
    layers = dec.layers
    if not hasattr(layers[0], "w_gate_up") or layers[0].quant != "fp8":
        return None
    dev = layers[0].w_qkv.device
    out = []
    for name, wattr, sattr in (("qkv_proj", "w_qkv", "s_qkv"), ("o_proj", "w_o", "s_o"),
                               ("gate_up_proj", "w_gate_up", "s_gate_up"), ("down_proj", "w_down", "s_down")):
        w0 = getattr(layers[0], wattr)
        N, K = w0.shape
        a = torch.randn(batch, K, device=dev).clamp_(-3, 3).to(torch.float8_e4m3fn)
        a_s = torch.rand(batch, 1, device=dev) + 0.5
        c = torch.empty(batch, N, device=dev, dtype=torch.bfloat16)

        def sweep():
            for l in layers:
                ops.cutlass_scaled_mm_out(c, a, getattr(l, wattr).t(), a_s, getattr(l, sattr).view(1, -1))
        # one pass captured in a CUDA graph (as in the timed step): eager launches of a ~10 us kernel would time the
        # host's launch path (ctypes + two tensor-map encodes, ~16 us per call), not the kernel
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            sweep()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            sweep()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (3 * len(layers))
        del g
        byts = N * K + batch * K + batch * N * 2 + 4 * (batch + N)
        flops = 2.0 * batch * N * K
        gbs, tfs = byts / us / 1e3, flops / us / 1e6
        bound = "hbm" if byts / hbm_peak / 1e3 >= flops / tensor_peak / 1e6 else "tensor"
        out.append({"kernel": f"scaled_mm_kernel {name} [M={batch}, N={N}, K={K}]", "bound": bound,
                    "avg_launch_us": round(us, 2), "launches_timed": 3 * len(layers),
                    "achieved": round(gbs if bound == "hbm" else tfs, 1),
                    "peak": hbm_peak if bound == "hbm" else tensor_peak,
                    "unit": "GB/s" if bound == "hbm" else "TFLOP/s",
                    "frac": round((gbs / hbm_peak) if bound == "hbm" else (tfs / tensor_peak), 4),
                    "hbm_GBs": round(gbs, 1), "hbm_frac": round(gbs / hbm_peak, 4), "tensor_TFLOPs": round(tfs, 1),
                    "tensor_frac": round(tfs / tensor_peak, 4), "algorithmic_bytes_per_launch": byts,
                    "flops_per_launch": flops, "traffic": None})
    return {"gemms": out, "sum_us_per_layer": round(sum(g["avg_launch_us"] for g in out), 2),
            "hbm_peak_source": hbm_src, "tensor_peak_source": tensor_src,
            "how": "CUDA events around 3 replays of one CUDA graph holding back-to-back launches over all layers' weights "
                   "(PDL on, as in the captured step), per-launch average"}



def _finish_multi_rank(world):
    """N > 1: leave without tearing NCCL down.  The captured CUDA graph holds NCCL kernels; destroying the
    communicator under it (explicitly or from interpreter shutdown) can block for minutes waiting on the peer, which
    keeps torchrun -- and the box -- alive long after the JSON line is out.  All ranks meet at a barrier (so nobody
    exits while a peer still needs the communicator), flush, and exit 0 directly."""
    if world <= 1:
        return
    import torch.distributed as dist
    try:
        dist.barrier()
        torch.cuda.synchronize()
    finally:
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def run_gpu_arm(args, rank, world, local_rank):
    import torch.distributed as dist
    from vllm_b200 import _custom_ops as ops
    from vllm_b200 import _lib
    from vllm_b200.attention import B200PagedAttentionMetadata
    from vllm_b200.decode_step import CONFIGS, Decoder
    from vllm_b200.distributed import init_peer_allreduce, init_tensor_parallel
    _lib.load()  # fail loudly when the CUDA library is missing
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    init_tensor_parallel()
    # N > 1: row-parallel all-reduces go over NVLink peer memory, fused with the norm that consumes them
    # (csrc/peer_allreduce.cu); VB200_PEER_AR=0 keeps NCCL for every all-reduce
    peer_ar = init_peer_allreduce(dev) if world > 1 else None
    model, batch, ctx, quant, kvd = WORKLOADS[args.config]
    cfg = CONFIGS[model]
    bs = 16
    # the context advances one token per step: room for every step this run makes (eager warm-up, graph warm-up,
    # the timed steps, the e2e leg, the eager roofline replay) plus slack
    headroom = ((args.warmup + 2 * args.steps + 24 + bs - 1) // bs + 1) * bs
    bps = (ctx + headroom + bs - 1) // bs
    num_blocks = batch * bps + 64
    dec = Decoder(cfg, num_blocks, bs, kvd, quant, device=dev)
    dec.fill_kv_cache_random()
    inp = StepInputs(batch, ctx, bs, num_blocks, cfg.vocab_size, dev, headroom)
    # max_decode_seq_len bounds EVERY replayed step (it sizes the split-K grid and partial buffers at capture)
    meta = B200PagedAttentionMetadata(seq_lens_tensor=inp.d_sl, max_decode_seq_len=ctx + headroom,
                                      block_tables=inp.d_bt, num_decode_tokens=batch, slot_mapping=inp.d_slots)
    d_next = torch.empty(batch, 1, dtype=torch.int64, device=dev)

    # ---- count launches of OUR kernels in one step (host-side counter around the C ABI) -------------
    lib = _lib.load()
    counter = {"n": 0}
    per_call = {"vb200_paged_attention_v1": 2, "vb200_paged_attention_v2": 2, "vb200_paged_attention_v1_quant": 2,
                "vb200_dynamic_scaled_fp8_quant": 2,
                "vb200_moe_grouped_gemm_fp8": 2, "vb200_moe_grouped_gemm_fp8_q16": 2, "vb200_moe_grouped_gemm_16": 2}
    wrapped = {}
    for name in _lib.SIGNATURES:
        if name.startswith("vb200_") and not any(s in name for s in ("size", "error", "version", "supports")):
            fn = getattr(lib, name)
            wrapped[name] = fn

            def mk(fn, name):
                def w(*a):
                    counter["n"] += per_call.get(name, 1)
                    return fn(*a)
                return w
            setattr(lib, name, mk(fn, name))

    def step():
        """One decode step + the on-device advance to the next one (sampled ids become the next input ids,
        positions / seq_lens += 1, slots follow the block tables): replaying this graph IS the decode loop."""
        d_next.copy_(dec.greedy_step(inp.d_ids, inp.d_pos, meta).view(batch, 1))
        ops.advance_step_flashattn(batch, batch, bs, inp.d_ids, d_next, inp.d_pos, inp.d_sl, inp.d_slots, inp.d_bt)

    # eager warm-up (allocator, workspaces, NCCL), then capture the step in a CUDA graph
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        counter["n"] = 0
        step()
        launches_per_step = counter["n"]
        step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    for name, fn in wrapped.items():
        setattr(lib, name, fn)
    inp.set_host_context(ctx)   # the two eager steps advanced the device state: start the captured loop at `ctx`
    inp.upload()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: device-resident inputs, graph replay ------------------------------------------------
    for _ in range(args.warmup):
        graph.replay()
    barrier()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if args.profile_window:   # ncu --profile-from-start off: only the timed steps are profiled
        torch.cuda.profiler.start()
    e0.record()
    for _ in range(args.steps):
        graph.replay()
    e1.record()
    barrier()
    if args.profile_window:
        torch.cuda.profiler.stop()
    ms = e0.elapsed_time(e1) / args.steps
    # ---- e2e: host inputs every step, sampled ids back to the host every step -------------------------
    dev_evs = []
    host_t = {"upload_call": 0.0, "replay_call": 0.0, "d2h_call": 0.0, "next_context": 0.0, "wait": 0.0, "ids": 0.0}

    def e2e_step():
        pc = time.perf_counter
        evs4 = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        t_0 = pc()
        evs4[0].record()
        inp.upload()                       # H2D: ids, positions, seq_lens, block tables, slots of THIS step (one copy)
        evs4[1].record()
        t_1 = pc()
        graph.replay()
        t_2 = pc()
        evs4[2].record()
        inp.h_out.copy_(d_next.view(-1), non_blocking=True)   # D2H: sampled ids
        evs4[3].record()
        t_3 = pc()
        dev_evs.append(evs4)
        # while the device runs: the host writes the NEXT step's context into its other pinned buffer
        inp.flip()
        inp.set_host_context(inp.cur + 1)
        t_4 = pc()
        torch.cuda.current_stream().synchronize()
        t_5 = pc()
        inp.h_ids.copy_(inp.h_out)         # the host owns the loop: next ids are what came back
        t_6 = pc()
        for k, v in zip(host_t, (t_1 - t_0, t_2 - t_1, t_3 - t_2, t_4 - t_3, t_5 - t_4, t_6 - t_5)):
            host_t[k] += v
    inp.set_host_context(ctx + args.warmup + args.steps)   # carry on where the device loop stopped
    for _ in range(min(args.warmup, 3)):
        e2e_step()
    barrier()
    for k in host_t:
        host_t[k] = 0.0
    dev_evs.clear()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    e2e_s = (time.perf_counter() - t0) / args.steps
    host_us = {k: round(v / args.steps * 1e6, 1) for k, v in host_t.items()}
    # where the device spent the step (CUDA events on the stream): the H2D copy, the captured step, the D2H copy
    dev_us = {name: round(statistics.mean(e[i].elapsed_time(e[i + 1]) for e in dev_evs) * 1e3, 1)
              for i, name in enumerate(("h2d", "graph", "d2h"))}
    barrier()
    if os.environ.get("VB200_E2E_PROBE") == "1" and rank == 0:
        # what an isolated replay (the device idle before it) costs over a queued one
        def timed(n_per_sync, reps, sleep_s=0.0):
            tot = 0.0
            for _ in range(reps):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                if sleep_s:
                    time.sleep(sleep_s)
                a.record()
                for _ in range(n_per_sync):
                    graph.replay()
                b.record()
                torch.cuda.synchronize()
                tot += a.elapsed_time(b)
            return tot / (reps * n_per_sync)
        inp.set_host_context(ctx)
        inp.upload()
        for name, n, r, sl in (("queued x10", 10, 1, 0.0), ("1 per sync", 1, 8, 0.0), ("2 per sync", 2, 4, 0.0),
                               ("4 per sync", 4, 2, 0.0), ("1 per sync after 2 ms idle", 1, 6, 0.002)):
            inp.upload()   # reset the context so that every probe attends over the same lengths
            torch.cuda.synchronize()
            print(f"[e2e-probe] {name}: {timed(n, r, sl):.3f} ms per replay", file=sys.stderr, flush=True)
        inp.set_host_context(ctx + args.warmup + 2 * args.steps + 3)
        inp.upload()
    clk = clocks.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms, e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_s = float(t[0]), float(t[1])

    # ---- roofline of the dominant kernel: eager replay of the step with events around every attention call ----
    layer0 = dec.layers[0]
    attn_ms = []
    real_fd = __import__("vllm_b200.attention.paged_attn", fromlist=["PagedAttention"]).PagedAttention
    orig = real_fd.forward_decode
    evs = []

    def timed_forward_decode(*a, **k):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        r = orig(*a, **k)
        e.record()
        evs.append((s, e))
        return r
    real_fd.forward_decode = staticmethod(timed_forward_decode)
    # fp8 layers take the attention through forward_quant (output quantisation fused into the split-K reduce kernel)
    impl_cls = __import__("vllm_b200.attention.backend", fromlist=["B200PagedAttentionImpl"]).B200PagedAttentionImpl
    orig_fq = impl_cls.forward_quant

    def timed_forward_quant(self, *a, **k):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        r = orig_fq(self, *a, **k)
        e.record()
        evs.append((s, e))
        return r
    impl_cls.forward_quant = timed_forward_quant
    try:
        inp.upload()
        for _ in range(max(2, min(args.steps, 5))):
            step()
        torch.cuda.synchronize()
    finally:
        real_fd.forward_decode = staticmethod(orig)
        impl_cls.forward_quant = orig_fq
    attn_ms = [s.elapsed_time(e) for s, e in evs[len(dec.layers):]]  # drop the first step
    attn_s = statistics.mean(attn_ms) * 1e-3
    cache_bytes = 1 if kvd != "auto" else 2
    n_eager = len(evs) // len(dec.layers)
    ctx_eager = inp.cur + (1 + n_eager - 1) / 2.0   # mean context of the timed eager steps (first one dropped)
    algo = int(attention_algorithmic_bytes(batch, ctx_eager, layer0.nh, layer0.nkv, cfg.head_dim, cache_bytes))
    peak, peak_src = peaks()
    achieved = algo / attn_s / 1e9
    gemm_rf = gemm_rooflines(dec, batch, peak, peak_src) if rank == 0 else None
    if rank != 0:
        _finish_multi_rank(world)
        return
    tps = batch / (ms * 1e-3)
    line = {
        "metric": METRIC if args.config == "cfg3" else f"decode tokens/s ({args.config})",
        "value": round(tps, 1), "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "fp8_e4m3 x fp8_e4m3 -> f32 accumulate (GEMMs), bf16 activations, fp8_e4m3 KV" if quant == "fp8" else "bf16",
        "data": "synthetic", "config": workload_config(args.config, world),
        "tokens_per_s_per_gpu": round(tps / world, 1),
        "e2e": {"value": round(batch / e2e_s, 1), "unit": "tokens/s", "h2d_bytes_per_step": inp.h2d_bytes,
                "d2h_bytes_per_step": inp.d2h_bytes, "ms_per_step": round(e2e_s * 1e3, 4),
                "host_us_per_step": host_us, "device_us_per_step": dev_us},
        "gpu_launches": launches_per_step * args.steps,
        "gpu_launches_per_step": launches_per_step,
        "roofline": {"bound": "hbm", "kernel": "paged_attention_kernel (+ split-K reduce with the fused per-token fp8 output quant)" if quant == "fp8" else "paged_attention_kernel (+ split-K reduce)",
                     "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                     "traffic": kernel_traffic(args.config, world), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": algo, "avg_launch_us": round(attn_s * 1e6, 2),
                     "launches_timed": len(attn_ms),
                     "share_of_step": round(attn_s * 1e3 * len(dec.layers) / ms, 4),
                     "how": "CUDA events around each layer's paged_attention call in an eager replay of the same "
                            "step right after the timed region (the timed region itself is one CUDA graph)"},
        "clocks": clk,
    }
    if gemm_rf is not None:
        line["roofline_gemm"] = gemm_rf
    line["config"]["allreduce"] = ("none" if world == 1 else "nccl" if peer_ar is None else
                                   peer_ar.algo_name(batch, cfg.hidden_size, torch.bfloat16))
    line["config"]["pdl"] = os.environ.get("VB200_PDL", "1") != "0"
    if world == 1 and not args.no_cpu_baseline:
        del dec
        torch.cuda.empty_cache()
        try:
            _, info, _ = cpu_reference_step_sample(args.config)
            line["cpu_baseline"] = info
        except Exception as ex:  # the CPU leg must not lose the GPU numbers
            line["cpu_baseline"] = {"value": None, "unit": "tokens/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": f"failed: {type(ex).__name__}: {ex}"}
    print(json.dumps(line), flush=True)
    _finish_multi_rank(world)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg3", choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile-window", action="store_true",
                    help="cudaProfilerStart/Stop around the timed steps (for ncu --profile-from-start off)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback); use --impl reference "
                         "for the CPU arm")
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit(f"bench.py --gpus {args.gpus}: launch with torchrun --nproc-per-node {args.gpus}")
    run_gpu_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
