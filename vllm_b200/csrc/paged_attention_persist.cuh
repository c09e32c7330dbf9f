// Persistent variant of the bulk-copy paged-attention kernel (paged_attention_tma.cuh) for batches whose natural grid
// fills the machine badly.
//
// Why: a CTA of the non-persistent kernel is one (sequence, kv-head group, partition); all CTAs are resident at once
// (4 per SM), so with 512 CTAs on 592 slots -- config 5: 128 sequences x 1 kv head x 4 partitions; the TP-8 shard of
// config 3 -- 68 SMs hold 4 CTAs and 80 hold 3, and the kernel takes as long as the SMs with 4
// (profiles/r01_paged_attention_cfg5_ncu_summary_s28.txt: 0.76 of the HBM peak).  Finer partitions even the load out but
// every CTA pays its own fill / drain bubble (block table, barriers, first bytes ~2 us after the first copy is issued,
// merge), which costs what the balance gains (profiles/r01_paged_attention_cfg5_partition_sweep_s29.txt).
// Here the grid is exactly the resident slots (4 CTAs per SM) and every CTA walks a static round-robin list of
// fixed-size work items (sequence, kv-head group, 512-token partition).  The producer warp does not stop at item
// boundaries: while the consumer warps merge and write item i it is already streaming item i + 1 into the ring, so
// there is one fill bubble per CTA instead of one per item, and per-SM work differs by at most one small item.
//   * ring, barriers, fragment loads and math: identical to paged_attention_tma_kernel (same stage <-> consumer warp
//     binding: global tile g of the CTA belongs to warp g % NWARPS and stage g % num_stages, num_stages % NWARPS == 0);
//   * block ids: the producer warp loads 32 at a time (one coalesced load per 512 tokens) and hands them to the issuing
//     lane by shuffle -- no shared-memory block table, nothing to synchronise with the consumers;
//   * per-item merge of the 4 warps' (m, l, O) through its own shared-memory area (the ring stays live), bracketed by a
//     named barrier of the 128 consumer threads only;
//   * outputs: the reference's v2 partial formats, then paged_attention_reduce_kernel -- unchanged.
#pragma once

namespace vb200 {

__device__ __forceinline__ void pa_named_bar(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

template <typename QT, bool FP8, int HEAD_SIZE, int NWARPS, int MIN_CTAS>
__global__ void __launch_bounds__((NWARPS + 1) * 32, MIN_CTAS)
    paged_attention_persist_kernel(const PagedAttnParams p, const int num_stages, const int total_items) {
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

  const int num_q_per_kv = p.num_heads / p.num_kv_heads;
  const int groups_per_kv = (num_q_per_kv + 7) >> 3;
  const int HG = p.num_kv_heads * groups_per_kv;   // (kv head, head group) pairs per sequence
  const int P = p.max_num_partitions;

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int r = lane >> 2;  // 0..7
  const int c = lane & 3;   // 0..3

  extern __shared__ __align__(128) unsigned char smem_raw_p[];
  // layout: [ring: num_stages x (K block | V block)] | merge: [NWARPS][8][HEAD_SIZE] fp32 O, [NWARPS][8] m, l | barriers
  const int ring_bytes = num_stages * STAGE;
  constexpr int MERGE_BYTES = NWARPS * 8 * HEAD_SIZE * 4 + 2 * NWARPS * 8 * 4;
  unsigned char* ring = smem_raw_p;
  float* sm_o = reinterpret_cast<float*>(smem_raw_p + ring_bytes);
  float* sm_m = sm_o + NWARPS * 8 * HEAD_SIZE;
  float* sm_l = sm_m + NWARPS * 8;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_raw_p + ring_bytes + ((MERGE_BYTES + 127) & ~127));
  uint64_t* empty_bar = full_bar + kPaMaxStages;

  if (tid == 0) {
    for (int s = 0; s < num_stages; ++s) {
      pa_mbar_init(&full_bar[s], 1);
      pa_mbar_init(&empty_bar[s], 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // item -> (sequence, head group, partition); the token range of the partition; 0 tiles = nothing to do
  auto decode = [&](int item, int& seq_idx, int& hg, int& part_idx, int& seq_len, int& tok_begin, int& ntok) {
    part_idx = item % P;
    const int rest = item / P;
    hg = rest % HG;
    seq_idx = rest / HG;
    seq_len = p.seq_lens[seq_idx];
    tok_begin = part_idx * p.partition_size;
    ntok = max(min(tok_begin + p.partition_size, seq_len) - tok_begin, 0);
    // an empty sequence still writes its (all-zero) row through partition 0, as the non-persistent kernel does
    return ntok > 0 || (seq_len <= 0 && part_idx == 0);
  };

  if (warp == NWARPS) {
    // ===== producer warp: all lanes fetch block ids, lane 0 issues two bulk copies per 16-token block =====
    int stage = 0;
    uint32_t phase = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
      int seq_idx, hg, part_idx, seq_len, tok_begin, ntok;
      if (!decode(item, seq_idx, hg, part_idx, seq_len, tok_begin, ntok)) continue;
      const int ntiles = (ntok + 15) >> 4;
      const int kv_head = hg / groups_per_kv;
      const unsigned char* kbase = reinterpret_cast<const unsigned char*>(p.k_cache) + (int64_t)kv_head * p.kv_head_stride * CE;
      const unsigned char* vbase = reinterpret_cast<const unsigned char*>(p.v_cache) + (int64_t)kv_head * p.kv_head_stride * CE;
      const int64_t blk_stride_b = p.kv_block_stride * CE;
      const int* bt = p.block_tables + (int64_t)seq_idx * p.max_num_blocks_per_seq + (tok_begin >> 4);
      for (int t0 = 0; t0 < ntiles; t0 += 32) {
        const int mine = (t0 + lane < ntiles) ? bt[t0 + lane] : 0;
        const int n = min(32, ntiles - t0);
        for (int j = 0; j < n; ++j) {
          const int blk = __shfl_sync(0xffffffffu, mine, j);
          if (lane == 0) {
            pa_mbar_wait(&empty_bar[stage], phase ^ 1);
            const int64_t off = (int64_t)blk * blk_stride_b;
            unsigned char* dst = ring + stage * STAGE;
            pa_mbar_expect_tx(&full_bar[stage], STAGE);
            pa_bulk_load(dst, kbase + off, KBYTES, &full_bar[stage]);
            pa_bulk_load(dst + KBYTES, vbase + off, KBYTES, &full_bar[stage]);
          }
          if (++stage == num_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        __syncwarp();
      }
    }
    return;
  }

  // ===== consumer warps =====
  const int t0s = ((r >> 1) << 2) + (r & 1);   // token slot of this thread inside a tile (see paged_attention.cu)
  const uint32_t ring_u32 = pa_smem_u32(ring);
  const float v_scale = FP8 ? *p.v_scale : 1.f;
  const float k_scale = FP8 ? *p.k_scale : 1.f;
  const bool partitioned = P > 1 || p.force_partials;
  uint32_t g = 0;             // global tile counter of this CTA (same value in every consumer thread)
  uint32_t my_g = warp;       // next global tile this warp owns
  int stage = warp % num_stages;
  uint32_t phase = 0;

  for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
    int seq_idx, hg, part_idx, seq_len, tok_begin, ntok;
    if (!decode(item, seq_idx, hg, part_idx, seq_len, tok_begin, ntok)) continue;
    const int ntiles = (ntok + 15) >> 4;
    const int kv_head = hg / groups_per_kv;
    const int hgroup = hg - kv_head * groups_per_kv;
    const int head0 = kv_head * num_q_per_kv + hgroup * 8;
    const int nheads = min(8, num_q_per_kv - hgroup * 8);

    // query fragments (B operand of QK^T), as in paged_attention_tma_kernel
    uint32_t qf[NQK][2];
    {
      const QT* qp = reinterpret_cast<const QT*>(p.q) + (int64_t)seq_idx * p.q_stride + (int64_t)(head0 + r) * HEAD_SIZE;
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
    const float sl2 = p.scale * kLog2e * k_scale;
    float alibi2[2] = {0.f, 0.f};
    if (p.alibi_slopes != nullptr) {
#pragma unroll
      for (int h = 0; h < 2; ++h)
        if (2 * c + h < nheads) alibi2[h] = p.alibi_slopes[head0 + 2 * c + h] * kLog2e;
    }
    float o[NJ][4];
#pragma unroll
    for (int j = 0; j < NJ; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY};
    float l_run[2] = {0.f, 0.f};

    const uint32_t g_end = g + (uint32_t)ntiles;
    for (; my_g < g_end; my_g += NWARPS) {
      const int tile = (int)(my_g - g);
      pa_mbar_wait(&full_bar[stage], phase);
      const uint32_t ks = ring_u32 + stage * STAGE;
      const uint32_t vs = ks + KBYTES;
      const int T0 = tile << 4;  // token offset inside the partition
      float s[4] = {0.f, 0.f, 0.f, 0.f};
      {
        const uint32_t ka = ks + t0s * 16 + 4 * c;
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
      const int tok_a = T0 + t0s, tok_b = tok_a + 2;  // partition-relative
      {
        const float pa = (float)(tok_begin + tok_a - seq_len + 1);
        const float pb = (float)(tok_begin + tok_b - seq_len + 1);
        s[0] = (tok_a < ntok) ? fmaf(s[0], sl2, alibi2[0] * pa) : -INFINITY;
        s[1] = (tok_a < ntok) ? fmaf(s[1], sl2, alibi2[1] * pa) : -INFINITY;
        s[2] = (tok_b < ntok) ? fmaf(s[2], sl2, alibi2[0] * pb) : -INFINITY;
        s[3] = (tok_b < ntok) ? fmaf(s[3], sl2, alibi2[1] * pb) : -INFINITY;
      }
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
      __syncwarp();
      if (lane == 0) pa_mbar_arrive(&empty_bar[stage]);
      stage += NWARPS;
      if (stage >= num_stages) {
        stage -= num_stages;
        phase ^= 1u;
      }
    }
    g = g_end;

    // ---- per-item merge of the NWARPS partial (m, l, O) and write-out (consumer threads only) ----
#pragma unroll
    for (int off = 4; off < 32; off <<= 1) {
      l_run[0] += __shfl_xor_sync(0xffffffffu, l_run[0], off);
      l_run[1] += __shfl_xor_sync(0xffffffffu, l_run[1], off);
    }
    pa_named_bar(2, NWARPS * 32);   // the previous item's merge reads are over
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
    pa_named_bar(2, NWARPS * 32);
    for (int idx = tid; idx < nheads * HEAD_SIZE; idx += NWARPS * 32) {
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
        const int64_t o2 = ((int64_t)seq_idx * p.num_heads + head) * P + part_idx;
        reinterpret_cast<QT*>(p.tmp_out)[o2 * HEAD_SIZE + d] = from_float<QT>(res);
        if (d == 0) {
          p.exp_sums[o2] = lsum;
          p.max_logits[o2] = mg * kLn2;
        }
      } else {
        reinterpret_cast<QT*>(p.out)[((int64_t)seq_idx * p.num_heads + head) * HEAD_SIZE + d] = from_float<QT>(res);
      }
    }
  }
}

template <int HEAD_SIZE, int NWARPS>
static inline size_t pa_persist_smem_bytes(int stages, int cache_elem_bytes) {
  const size_t stage = (size_t)2 * HEAD_SIZE * 16 * cache_elem_bytes;
  const size_t merge = ((size_t)NWARPS * 8 * HEAD_SIZE * 4 + 2 * NWARPS * 8 * 4 + 127) & ~(size_t)127;
  return stages * stage + merge + 2 * kPaMaxStages * 8;
}

}  // namespace vb200
