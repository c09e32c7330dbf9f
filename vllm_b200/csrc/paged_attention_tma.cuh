// Paged-attention decode, bulk-copy (TMA engine) load path for BLOCK_SIZE == 16.
//
// Why a second load path: a read-only probe (benchmarks/hbm_read_probe.cu, profiles/r01_hbm_read_probe.txt)
// shows that on B200 the HBM read rate of a kernel that loads with ld.global is set by the number of resident
// WARPS, not by the bytes each has in flight: 12 warps/SM (what the register-staged kernel above can keep
// resident at 150+ registers/thread) top out at 5.7 TB/s, 32 warps reach 7.1-7.3 TB/s.  cp.async.bulk requests
// are issued by the TMA engine and do not occupy per-warp load slots, so ONE producer warp can keep > 100 KB
// per SM in flight.  In the vLLM cache layout the K bytes of one (block, kv head) are contiguous
// ([head_size/x][16][x], head_size * 16 * sizeof(cache_t) bytes) and so are the V bytes ([head_size][16]):
// two 1-D bulk copies per 16-token block, no tensor map.
//
//   warp NWARPS (producer, one lane): for every block of the partition: wait empty[stage] -> expect_tx ->
//        bulk copy K block + V block into the stage (L2 evict-first: the cache is streamed once per step)
//   warps 0..NWARPS-1 (consumers): tiles strided over the warps; wait full[stage]; read MMA fragments straight
//        from the stage with 4-byte (K: word c of run X of token row) / 4- or 8-byte (V) shared loads; same
//        mma.sync m16n8k16 math, online softmax and partial formats as the register-staged kernel; arrive
//        empty[stage] as soon as the tile's fragments have been consumed.
// Bank behaviour: a K word load hits bank (4 * token + c) mod 32; the token permutation the PV fragment
// layout needs (rows r -> tokens 4(r/2) + (r&1), +2) makes it a 2-way conflict, V loads are conflict-free:
// 48 shared-memory wavefronts per 4 KiB tile, ~1/4 of the shared-memory bandwidth at 7 TB/s.
#pragma once

namespace vb200 {

__device__ __forceinline__ uint32_t pa_smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void pa_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(pa_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void pa_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(pa_smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void pa_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(pa_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void pa_mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(pa_smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!ok);
}
// 1-D bulk copy global -> shared with an L2 evict-first hint, completion (bytes) on an mbarrier
__device__ __forceinline__ void pa_bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
      ::"r"(pa_smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(pa_smem_u32(bar)),
      "l"(0x12F0000000000000ull)
      : "memory");
}
__device__ __forceinline__ uint32_t pa_lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint2 pa_lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
  return v;
}

constexpr int kPaMaxStages = 32;

// QT: query/output element type.  FP8: cache holds e4m3 bytes, else the cache element type is QT.
// HEAD_SIZE must be a multiple of 16.  Block size is 16 (one tile == one cache block).
template <typename QT, bool FP8, int HEAD_SIZE, int NWARPS, int MIN_CTAS>
__global__ void __launch_bounds__((NWARPS + 1) * 32, MIN_CTAS)
    paged_attention_tma_kernel(const PagedAttnParams p, const int num_stages) {
  pdl_wait();
  pdl_trigger();
  using MT = typename std::conditional<FP8, __half, QT>::type;  // tensor-core operand type
  using M = MmaType<MT>;
  constexpr int CE = FP8 ? 1 : 2;                  // cache element bytes
  constexpr int X = 16 / CE;                       // elements per 16-byte K run
  constexpr int NCHUNK = HEAD_SIZE / X;            // K runs per token
  constexpr int NQK = FP8 ? NCHUNK : NCHUNK / 2;   // m16n8k16 MMAs of one QK^T tile
  constexpr int NJ = HEAD_SIZE / 16;               // 16-row d tiles for PV
  constexpr int KBYTES = HEAD_SIZE * 16 * CE;      // bytes of K (and of V) of one block and kv head
  constexpr int STAGE = 2 * KBYTES;
  static_assert(HEAD_SIZE % 16 == 0, "head size must be a multiple of 16");

  const int seq_idx = blockIdx.y;
  const int part_idx = blockIdx.z;
  const int num_q_per_kv = p.num_heads / p.num_kv_heads;
  const int groups_per_kv = (num_q_per_kv + 7) >> 3;
  const int kv_head = blockIdx.x / groups_per_kv;
  const int hgroup = blockIdx.x - kv_head * groups_per_kv;
  const int head0 = kv_head * num_q_per_kv + hgroup * 8;
  const int nheads = min(8, num_q_per_kv - hgroup * 8);

  const int seq_len = p.seq_lens[seq_idx];
  const int tok_begin = part_idx * p.partition_size;
  if (tok_begin >= seq_len && !(seq_len <= 0 && part_idx == 0)) return;
  const int tok_end = min(tok_begin + p.partition_size, seq_len);
  const int ntok = max(tok_end - tok_begin, 0);
  const int ntiles = (ntok + 15) >> 4;

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int r = lane >> 2;  // 0..7
  const int c = lane & 3;   // 0..3

  extern __shared__ __align__(128) unsigned char smem_raw[];
  // layout: [ring: num_stages x (K block | V block)]  (re-used for the warp merge: [NWARPS][8][HEAD_SIZE] fp32 O,
  //         [NWARPS][8] m, [NWARPS][8] l) | full[32] empty[32] | block table ints
  const int ring_bytes = num_stages * STAGE;
  constexpr int MERGE_BYTES = NWARPS * 8 * HEAD_SIZE * 4 + 2 * NWARPS * 8 * 4;
  const int data_bytes = ((ring_bytes > MERGE_BYTES ? ring_bytes : MERGE_BYTES) + 127) & ~127;
  unsigned char* ring = smem_raw;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_raw + data_bytes);
  uint64_t* empty_bar = full_bar + kPaMaxStages;
  int* sm_bt = reinterpret_cast<int*>(empty_bar + kPaMaxStages);

  const int* bt = p.block_tables + (int64_t)seq_idx * p.max_num_blocks_per_seq;
  const int blk_begin = tok_begin >> 4;
  for (int i = tid; i < ntiles; i += (NWARPS + 1) * 32) sm_bt[i] = bt[blk_begin + i];
  if (tid == 0) {
    for (int s = 0; s < num_stages; ++s) {
      pa_mbar_init(&full_bar[s], 1);
      pa_mbar_init(&empty_bar[s], 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  float o[NJ][4];
#pragma unroll
  for (int j = 0; j < NJ; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};

  if (warp == NWARPS) {
    // ===== producer: one lane issues two bulk copies per block =====
    if (lane == 0) {
      const unsigned char* kbase = reinterpret_cast<const unsigned char*>(p.k_cache) +
                                   (int64_t)kv_head * p.kv_head_stride * CE;
      const unsigned char* vbase = reinterpret_cast<const unsigned char*>(p.v_cache) +
                                   (int64_t)kv_head * p.kv_head_stride * CE;
      const int64_t blk_stride_b = p.kv_block_stride * CE;
      for (int t = 0; t < ntiles; ++t) {
        const int stage = t % num_stages;
        const uint32_t phase = (t / num_stages) & 1;
        pa_mbar_wait(&empty_bar[stage], phase ^ 1);
        const int64_t off = (int64_t)sm_bt[t] * blk_stride_b;
        unsigned char* dst = ring + stage * STAGE;
        pa_mbar_expect_tx(&full_bar[stage], STAGE);
        pa_bulk_load(dst, kbase + off, KBYTES, &full_bar[stage]);
        pa_bulk_load(dst + KBYTES, vbase + off, KBYTES, &full_bar[stage]);
      }
    }
  } else {
    // ===== consumers =====
    // query fragments (B operand of QK^T): MMA i covers d = 16 i .. 16 i + 15; thread (r, c) holds head r,
    // fp8 cache: d = 16 i + 4 c + {0,1} | {2,3};  16-bit cache: d = 16 i + 2 c + {0,1} | 16 i + 8 + 2 c + {0,1}
    uint32_t qf[NQK][2];
    {
      const QT* qp = reinterpret_cast<const QT*>(p.q) + (int64_t)seq_idx * p.q_stride +
                     (int64_t)(head0 + r) * HEAD_SIZE;
      const bool hvalid = r < nheads;
#pragma unroll
      for (int i = 0; i < NQK; ++i) {
        uint32_t w0 = 0, w1 = 0;
        if (hvalid) {
          if (FP8) {
            const QT* e = qp + 16 * i + 4 * c;
            if (std::is_same<QT, MT>::value) {
              const uint2 v = *reinterpret_cast<const uint2*>(e);
              w0 = v.x;
              w1 = v.y;
            } else {
              w0 = M::pack(to_float<QT>(e[0]), to_float<QT>(e[1]));
              w1 = M::pack(to_float<QT>(e[2]), to_float<QT>(e[3]));
            }
          } else {
            w0 = *reinterpret_cast<const uint32_t*>(qp + 16 * i + 2 * c);
            w1 = *reinterpret_cast<const uint32_t*>(qp + 16 * i + 8 + 2 * c);
          }
        }
        qf[i][0] = w0;
        qf[i][1] = w1;
      }
    }
    float sl2 = p.scale * kLog2e;
    if (FP8) sl2 *= *p.k_scale;
    float alibi2[2] = {0.f, 0.f};
    if (p.alibi_slopes != nullptr) {
#pragma unroll
      for (int h = 0; h < 2; ++h)
        if (2 * c + h < nheads) alibi2[h] = p.alibi_slopes[head0 + 2 * c + h] * kLog2e;
    }
    // token slots of this thread inside a tile: QK^T rows r / r+8 <-> tokens t0 / t0+2 (see the header of
    // paged_attention.cu); PV k-slots of lane c <-> tokens 4c .. 4c+3
    const int t0 = ((r >> 1) << 2) + (r & 1);
    const uint32_t ring_u32 = pa_smem_u32(ring);

    for (int tile = warp; tile < ntiles; tile += NWARPS) {
      const int stage = tile % num_stages;
      const uint32_t phase = (tile / num_stages) & 1;
      pa_mbar_wait(&full_bar[stage], phase);
      const uint32_t ks = ring_u32 + stage * STAGE;
      const uint32_t vs = ks + KBYTES;
      const int T0 = tile << 4;  // token offset inside the partition
      // ---- S^T = K Q^T : rows = token slots, cols = heads ---------------------------------------------
      float s[4] = {0.f, 0.f, 0.f, 0.f};
      {
        const uint32_t ka = ks + t0 * 16 + 4 * c;  // word c of the runs of token t0; token t0 + 2 is 32 B on
#pragma unroll
        for (int i = 0; i < NQK; ++i) {
          uint32_t a0, a1, a2, a3;
          if (FP8) {
            const uint32_t wa = pa_lds32(ka + i * 256);
            const uint32_t wb = pa_lds32(ka + i * 256 + 32);
            a0 = e4m3x2_to_f16x2((uint16_t)(wa & 0xffffu));
            a2 = e4m3x2_to_f16x2((uint16_t)(wa >> 16));
            a1 = e4m3x2_to_f16x2((uint16_t)(wb & 0xffffu));
            a3 = e4m3x2_to_f16x2((uint16_t)(wb >> 16));
          } else {
            a0 = pa_lds32(ka + (2 * i) * 256);
            a2 = pa_lds32(ka + (2 * i + 1) * 256);
            a1 = pa_lds32(ka + (2 * i) * 256 + 32);
            a3 = pa_lds32(ka + (2 * i + 1) * 256 + 32);
          }
          M::mma(s, a0, a1, a2, a3, qf[i][0], qf[i][1]);
        }
      }
      // ---- scale, alibi, mask ---------------------------------------------------------------------------
      const int tok_a = T0 + t0, tok_b = tok_a + 2;  // partition-relative
      {
        const float pa = (float)(tok_begin + tok_a - seq_len + 1);
        const float pb = (float)(tok_begin + tok_b - seq_len + 1);
        s[0] = (tok_a < ntok) ? fmaf(s[0], sl2, alibi2[0] * pa) : -INFINITY;
        s[1] = (tok_a < ntok) ? fmaf(s[1], sl2, alibi2[1] * pa) : -INFINITY;
        s[2] = (tok_b < ntok) ? fmaf(s[2], sl2, alibi2[0] * pb) : -INFINITY;
        s[3] = (tok_b < ntok) ? fmaf(s[3], sl2, alibi2[1] * pb) : -INFINITY;
      }
      // ---- online softmax over the 16 token rows (8 r-lanes x 2) ------------------------------------------
      float mx0 = fmaxf(s[0], s[2]), mx1 = fmaxf(s[1], s[3]);
#pragma unroll
      for (int off = 4; off < 32; off <<= 1) {
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, off));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, off));
      }
      const float mn0 = fmaxf(m_run[0], mx0), mn1 = fmaxf(m_run[1], mx1);
      const float corr0 = fast_exp2(m_run[0] - mn0), corr1 = fast_exp2(m_run[1] - mn1);
      m_run[0] = mn0;
      m_run[1] = mn1;
      const float p0 = fast_exp2(s[0] - mn0), p1 = fast_exp2(s[1] - mn1);
      const float p2 = fast_exp2(s[2] - mn0), p3 = fast_exp2(s[3] - mn1);
      l_run[0] = fmaf(l_run[0], corr0, p0 + p2);
      l_run[1] = fmaf(l_run[1], corr1, p1 + p3);
      const uint32_t b0 = movmatrix_trans(M::pack(p0, p1));
      const uint32_t b1 = movmatrix_trans(M::pack(p2, p3));
      // ---- O^T += V^T P^T ; V of tokens past the end is zeroed (it may hold NaN / Inf garbage) -----------
      uint32_t vmask0 = 0xffffffffu, vmask1 = 0xffffffffu;
      if (T0 + 16 > ntok) {
        const int nv = min(max(ntok - (T0 + 4 * c), 0), 4);
        if (FP8) {
          vmask0 = nv >= 4 ? 0xffffffffu : ((1u << (8 * nv)) - 1u);
        } else {
          vmask0 = nv >= 2 ? 0xffffffffu : (nv == 1 ? 0x0000ffffu : 0u);
          vmask1 = nv >= 4 ? 0xffffffffu : (nv == 3 ? 0x0000ffffu : 0u);
        }
      }
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        o[j][0] *= corr0;
        o[j][1] *= corr1;
        o[j][2] *= corr0;
        o[j][3] *= corr1;
        uint32_t a0, a1, a2, a3;
        if (FP8) {
          // V row d = 16 bytes (16 tokens); lane c takes tokens 4c .. 4c+3
          const uint32_t wa = pa_lds32(vs + (16 * j + r) * 16 + 4 * c) & vmask0;
          const uint32_t wb = pa_lds32(vs + (16 * j + r + 8) * 16 + 4 * c) & vmask0;
          a0 = e4m3x2_to_f16x2((uint16_t)(wa & 0xffffu));
          a2 = e4m3x2_to_f16x2((uint16_t)(wa >> 16));
          a1 = e4m3x2_to_f16x2((uint16_t)(wb & 0xffffu));
          a3 = e4m3x2_to_f16x2((uint16_t)(wb >> 16));
        } else {
          const uint2 wa = pa_lds64(vs + (16 * j + r) * 32 + 8 * c);
          const uint2 wb = pa_lds64(vs + (16 * j + r + 8) * 32 + 8 * c);
          a0 = wa.x & vmask0;
          a2 = wa.y & vmask1;
          a1 = wb.x & vmask0;
          a3 = wb.y & vmask1;
        }
        M::mma(o[j], a0, a1, a2, a3, b0, b1);
      }
      // the stage's bytes are in registers / consumed: hand it back to the producer
      __syncwarp();
      if (lane == 0) pa_mbar_arrive(&empty_bar[stage]);
    }
  }

  // ---- per-warp totals -> shared memory (the ring is idle from here on) -----------------------------------
  __syncthreads();
  float* sm_o = reinterpret_cast<float*>(smem_raw);
  float* sm_m = sm_o + NWARPS * 8 * HEAD_SIZE;
  float* sm_l = sm_m + NWARPS * 8;
  if (warp < NWARPS) {
#pragma unroll
    for (int off = 4; off < 32; off <<= 1) {
      l_run[0] += __shfl_xor_sync(0xffffffffu, l_run[0], off);
      l_run[1] += __shfl_xor_sync(0xffffffffu, l_run[1], off);
    }
    if (r == 0) {
      sm_m[warp * 8 + 2 * c] = m_run[0];
      sm_m[warp * 8 + 2 * c + 1] = m_run[1];
      sm_l[warp * 8 + 2 * c] = l_run[0];
      sm_l[warp * 8 + 2 * c + 1] = l_run[1];
    }
    float* so = sm_o + warp * 8 * HEAD_SIZE;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int da = 16 * j + r, db = da + 8;
      so[(2 * c) * HEAD_SIZE + da] = o[j][0];
      so[(2 * c + 1) * HEAD_SIZE + da] = o[j][1];
      so[(2 * c) * HEAD_SIZE + db] = o[j][2];
      so[(2 * c + 1) * HEAD_SIZE + db] = o[j][3];
    }
  }
  __syncthreads();

  // ---- merge warps, normalise, write (identical to the register-staged kernel) ------------------------------
  const float v_scale = FP8 ? *p.v_scale : 1.f;
  const bool partitioned = p.max_num_partitions > 1 || p.force_partials;
  for (int idx = tid; idx < nheads * HEAD_SIZE; idx += (NWARPS + 1) * 32) {
    const int h = idx / HEAD_SIZE;
    const int d = idx - h * HEAD_SIZE;
    float mg = -INFINITY;
#pragma unroll
    for (int w = 0; w < NWARPS; ++w) mg = fmaxf(mg, sm_m[w * 8 + h]);
    float lsum = 0.f, acc = 0.f;
#pragma unroll
    for (int w = 0; w < NWARPS; ++w) {
      const float f = (sm_m[w * 8 + h] == -INFINITY) ? 0.f : fast_exp2(sm_m[w * 8 + h] - mg);
      lsum = fmaf(sm_l[w * 8 + h], f, lsum);
      acc = fmaf(sm_o[(w * 8 + h) * HEAD_SIZE + d], f, acc);
    }
    const float res = acc * v_scale * __fdividef(1.f, lsum + 1e-6f);
    const int head = head0 + h;
    if (partitioned) {
      const int64_t o2 = ((int64_t)seq_idx * p.num_heads + head) * p.max_num_partitions + part_idx;
      reinterpret_cast<QT*>(p.tmp_out)[o2 * HEAD_SIZE + d] = from_float<QT>(res);
      if (d == 0) {
        p.exp_sums[o2] = lsum;
        p.max_logits[o2] = mg * kLn2;  // back to natural-log units, as the reference stores
      }
    } else {
      reinterpret_cast<QT*>(p.out)[((int64_t)seq_idx * p.num_heads + head) * HEAD_SIZE + d] = from_float<QT>(res);
    }
  }
}

// shared-memory bytes of the bulk-copy kernel for a ring of `stages`
template <int HEAD_SIZE, int NWARPS>
static inline size_t pa_tma_smem_bytes(int stages, int cache_elem_bytes, int partition_tokens) {
  const size_t stage = (size_t)2 * HEAD_SIZE * 16 * cache_elem_bytes;
  const size_t merge = (size_t)NWARPS * 8 * HEAD_SIZE * 4 + 2 * NWARPS * 8 * 4;
  size_t data = stages * stage > merge ? stages * stage : merge;
  data = (data + 127) & ~(size_t)127;
  return data + 2 * kPaMaxStages * 8 + (size_t)((partition_tokens + 15) / 16) * 4;
}

}  // namespace vb200
