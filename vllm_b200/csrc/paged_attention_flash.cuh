// Paged-attention decode over the FLASH cache layout  cache[block][token][kv head][d]  addressed through runtime
// strides, i.e. the NHD layout and the HND-permuted view  [block][kv head][token][d]  that vLLM's V1 backends use
// (/root/reference/vllm/v1/attention/backends/flash_attn.py:79-102, forced to HND on Blackwell:
// /root/reference/vllm/platforms/cuda.py:264-290; written by reshape_and_cache_flash, csrc/cache_kernels.cu:264-306).
// Same math, split-K partials and reduce kernel as paged_attention_tma_kernel; what differs is how a 16-token tile
// gets from HBM into MMA fragments:
//
//   * loads: the K (and V) bytes of one (block, kv head) are 16 rows of HEAD_SIZE * sizeof(cache_t) bytes -- one
//     contiguous slab in HND, rows `num_kv_heads * HEAD_SIZE` elements apart in NHD.  A 4-D tensor map over the cache
//     bytes {d bytes, token, head, block} with the cache's own strides covers both; the producer lane issues one
//     cp.async.bulk.tensor box {128 B, 16 tokens} per 128 bytes of head dimension (1 box for fp8 / head 128, 2 for
//     bf16) with the hardware 128-byte swizzle, completion on the stage's mbarrier.
//   * K fragments (A operand of S^T = K Q^T, row = token, k = d): token rows are 128 bytes apart in a box, so plain
//     word loads would hit one bank 8 ways; the swizzle puts 16-byte chunk c of token t at chunk c ^ (t & 7), which
//     makes the 8 token rows of a fragment load land in 8 different chunk positions: conflict-free 4-byte loads.
//   * V fragments (A operand of O^T = V^T P^T, row = d, k = token): V is [token][d] here, so consecutive k of one row
//     are 128 bytes apart -- the transpose comes from ldmatrix.trans.  16-bit caches: the textbook x4.trans.  fp8: the
//     8-bit tile is read as a b16 matrix of d-PAIRS; after the transpose a register holds (d, d+1) of tokens 2c and
//     2c+1; one prmt regroups the bytes into (token 2c, token 2c+1) of d and of d+1, and MMA rows r / r+8 are bound to
//     d = 2r / 2r+1 of the 16-d tile (legal: the row <-> d binding only decides which accumulator holds which d).
//   * tokens keep their natural order (MMA row r <-> token r), unlike the paged-layout kernel whose V loads dictate a
//     permutation.
#pragma once
#include <cuda.h>

namespace vb200 {

__device__ __forceinline__ void pa_tma_load_4d(uint32_t smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1,
                                               int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1, {%3, %4, %5, %6}], [%2], %7;"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(pa_smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3),
      "l"(0x12F0000000000000ull)   // evict-first: the cache is streamed once per step
      : "memory");
}
__device__ __forceinline__ void pa_ldmatrix_x2_trans(uint32_t addr, uint32_t& r0, uint32_t& r1) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0, %1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
}
__device__ __forceinline__ void pa_ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                                     uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

// QT: query/output element type.  FP8: cache holds e4m3 bytes, else the cache element type is QT.
// HEAD_SIZE * sizeof(cache_t) must be a multiple of 128 (one swizzled box per 128 bytes).  Block size is 16.
template <typename QT, bool FP8, int HEAD_SIZE, int NWARPS, int MIN_CTAS>
__global__ void __launch_bounds__((NWARPS + 1) * 32, MIN_CTAS)
    paged_attention_flash_kernel(const __grid_constant__ CUtensorMap tmap_k, const __grid_constant__ CUtensorMap tmap_v,
                                 const PagedAttnParams p, const int num_stages) {
  pdl_wait();
  pdl_trigger();
  using MT = typename std::conditional<FP8, __half, QT>::type;  // tensor-core operand type
  using M = MmaType<MT>;
  constexpr int CE = FP8 ? 1 : 2;                  // cache element bytes
  constexpr int ROWB = HEAD_SIZE * CE;             // bytes of one token's K (or V) row of one head
  constexpr int NBOX = ROWB / 128;                 // 128-byte-wide boxes per tile
  constexpr int BOXB = 16 * 128;                   // bytes of one box (16 tokens)
  constexpr int KBYTES = NBOX * BOXB;              // K (and V) bytes of one tile
  constexpr int STAGE = 2 * KBYTES;
  constexpr int NQK = HEAD_SIZE / 16;              // m16n8k16 MMAs of one QK^T tile
  constexpr int NJ = HEAD_SIZE / 16;               // 16-row d tiles for PV
  static_assert(ROWB % 128 == 0, "head_size * sizeof(cache element) must be a multiple of 128");

  const int seq_idx = blockIdx.y;
  const int part_idx = blockIdx.z;
  const int num_q_per_kv = p.num_heads / p.num_kv_heads;
  const int groups_per_kv = (num_q_per_kv + 7) >> 3;
  const int kv_head = blockIdx.x / groups_per_kv;
  const int hgroup = blockIdx.x - kv_head * groups_per_kv;
  const int head0 = kv_head * num_q_per_kv + hgroup * 8;
  const int nheads = min(8, num_q_per_kv - hgroup * 8);

  const int seq_len = p.seq_lens[seq_idx];
  // sliding window: the first visible key and the 16-token block that holds it; partition 0 starts at that block
  const int win_start = p.window_left >= 0 ? max(0, seq_len - 1 - p.window_left) : 0;
  const int tok_begin = (win_start & ~15) + part_idx * p.partition_size;
  if (tok_begin >= seq_len && !(seq_len <= 0 && part_idx == 0)) return;
  const int tok_end = min(tok_begin + p.partition_size, seq_len);
  const int ntok = max(tok_end - tok_begin, 0);
  const int ntiles = (ntok + 15) >> 4;

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int r = lane >> 2;  // 0..7
  const int c = lane & 3;   // 0..3

  extern __shared__ __align__(128) unsigned char smem_raw_f[];
  // swizzled boxes need 1024-byte alignment
  unsigned char* smem_al = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw_f) + 1023) &
                                                           ~static_cast<uintptr_t>(1023));
  const int ring_bytes = num_stages * STAGE;
  constexpr int MERGE_BYTES = NWARPS * 8 * HEAD_SIZE * 4 + 2 * NWARPS * 8 * 4;
  const int data_bytes = ((ring_bytes > MERGE_BYTES ? ring_bytes : MERGE_BYTES) + 127) & ~127;
  unsigned char* ring = smem_al;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_al + data_bytes);
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
  const uint32_t ring_u32 = pa_smem_u32(ring);

  if (warp == NWARPS) {
    // ===== producer: one lane issues 2 * NBOX boxes per tile =====
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_k)) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_v)) : "memory");
      int stage = 0;
      uint32_t phase = 0;
      for (int t = 0; t < ntiles; ++t) {
        pa_mbar_wait(&empty_bar[stage], phase ^ 1);
        const int blk = sm_bt[t];
        const uint32_t dst = ring_u32 + stage * STAGE;
        pa_mbar_expect_tx(&full_bar[stage], STAGE);
#pragma unroll
        for (int b = 0; b < NBOX; ++b) {
          pa_tma_load_4d(dst + b * BOXB, &tmap_k, &full_bar[stage], b * 128, 0, kv_head, blk);
          pa_tma_load_4d(dst + KBYTES + b * BOXB, &tmap_v, &full_bar[stage], b * 128, 0, kv_head, blk);
        }
        if (++stage == num_stages) {
          stage = 0;
          phase ^= 1u;
        }
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
    // logits soft-capping (flash_attn.py:371-374): cap * tanh(logit / cap) -- the tanh takes the raw dot product times
    // scale / cap, the exponent scale becomes cap * log2(e); tanh.approx as in the library kernel
    const bool capped = p.softcap > 0.f;
    const float cap_pre = capped ? sl2 / (kLog2e * p.softcap) : 0.f;
    if (capped) sl2 = p.softcap * kLog2e;
    float alibi2[2] = {0.f, 0.f};
    if (p.alibi_slopes != nullptr) {
#pragma unroll
      for (int h = 0; h < 2; ++h)
        if (2 * c + h < nheads) alibi2[h] = p.alibi_slopes[head0 + 2 * c + h] * kLog2e;
    }
    // ldmatrix row address of this lane inside a V box: lanes 0-7 -> tokens 0-7, 8-15 -> tokens 8-15 (fp8, x2);
    // 16-bit (x4): lanes 0-7 tokens 0-7 chunk A, 8-15 tokens 0-7 chunk A+1, 16-23 tokens 8-15 chunk A, 24-31 tokens
    // 8-15 chunk A+1
    const int lm_tok = FP8 ? (lane & 15) : ((lane & 7) + ((lane >> 4) << 3));
    const int lm_chunk_add = FP8 ? 0 : ((lane >> 3) & 1);

    int stage = warp % num_stages;   // tiles are strided over the warps; num_stages is a multiple of NWARPS
    uint32_t phase = 0;
    for (int tile = warp; tile < ntiles; tile += NWARPS) {
      pa_mbar_wait(&full_bar[stage], phase);
      const uint32_t ks = ring_u32 + stage * STAGE;
      const uint32_t vs = ks + KBYTES;
      const int T0 = tile << 4;  // token offset inside the partition
      // ---- S^T = K Q^T : rows = tokens r / r + 8, cols = heads --------------------------------------------
      float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < NQK; ++i) {
        uint32_t a0, a1, a2, a3;
        if (FP8) {
          // 16 d = one 16-byte chunk: chunk i of the row, box i / 8
          const uint32_t base = ks + (i >> 3) * BOXB + r * 128 + ((((i & 7) ^ r) << 4)) + 4 * c;
          const uint32_t wa = pa_lds32(base);
          const uint32_t wb = pa_lds32(base + 1024);   // token r + 8: same swizzle phase, 8 rows further
          a0 = e4m3x2_to_f16x2((uint16_t)(wa & 0xffffu));
          a2 = e4m3x2_to_f16x2((uint16_t)(wa >> 16));
          a1 = e4m3x2_to_f16x2((uint16_t)(wb & 0xffffu));
          a3 = e4m3x2_to_f16x2((uint16_t)(wb >> 16));
        } else {
          // 16 d = 32 bytes = chunks 2i, 2i + 1 of the row, box i / 4
          const uint32_t rowb = ks + (i >> 2) * BOXB + r * 128 + 4 * c;
          const int ch = (2 * i) & 7;
          a0 = pa_lds32(rowb + ((ch ^ r) << 4));
          a2 = pa_lds32(rowb + (((ch + 1) ^ r) << 4));
          a1 = pa_lds32(rowb + 1024 + ((ch ^ r) << 4));
          a3 = pa_lds32(rowb + 1024 + (((ch + 1) ^ r) << 4));
        }
        M::mma(s, a0, a1, a2, a3, qf[i][0], qf[i][1]);
      }
      // ---- scale, alibi, mask -----------------------------------------------------------------------------
      const int tok_a = T0 + r, tok_b = tok_a + 8;  // partition-relative
      {
        const float pa = (float)(tok_begin + tok_a - seq_len + 1);
        const float pb = (float)(tok_begin + tok_b - seq_len + 1);
        // visible: inside the partition / sequence and not left of the window (only the window's first block has such keys)
        if (capped) {
#pragma unroll
          for (int e = 0; e < 4; ++e) asm("tanh.approx.f32 %0, %1;" : "=f"(s[e]) : "f"(s[e] * cap_pre));
        }
        const bool va = tok_a < ntok && tok_begin + tok_a >= win_start;
        const bool vb = tok_b < ntok && tok_begin + tok_b >= win_start;
        s[0] = va ? fmaf(s[0], sl2, alibi2[0] * pa) : -INFINITY;
        s[1] = va ? fmaf(s[1], sl2, alibi2[1] * pa) : -INFINITY;
        s[2] = vb ? fmaf(s[2], sl2, alibi2[0] * pb) : -INFINITY;
        s[3] = vb ? fmaf(s[3], sl2, alibi2[1] * pb) : -INFINITY;
      }
      // ---- online softmax over the 16 token rows (8 r-lanes x 2) --------------------------------------------
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
      const uint32_t b0 = movmatrix_trans(M::pack(p0, p1));   // (k = tokens 2c, 2c+1; n = head r)
      const uint32_t b1 = movmatrix_trans(M::pack(p2, p3));   // (k = tokens 8+2c, 9+2c; n = head r)
      // ---- O^T += V^T P^T ; V of tokens past the end is zeroed (it may hold NaN / Inf garbage) -----------------
      // after ldmatrix.trans a register holds token 2c (+8) in its low half and token 2c+1 (+8) in its high half
      uint32_t vm_lo = 0xffffffffu, vm_hi = 0xffffffffu;   // masks of the tokens 0-7 / 8-15 registers
      if (T0 + 16 > ntok) {
        const int n0 = ntok - (T0 + 2 * c), n1 = n0 - 8;
        vm_lo = n0 >= 2 ? 0xffffffffu : (n0 == 1 ? 0x0000ffffu : 0u);
        vm_hi = n1 >= 2 ? 0xffffffffu : (n1 == 1 ? 0x0000ffffu : 0u);
      }
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        o[j][0] *= corr0;
        o[j][1] *= corr1;
        o[j][2] *= corr0;
        o[j][3] *= corr1;
        uint32_t a0, a1, a2, a3;
        if (FP8) {
          // d tile j = chunk j of every token row (box j / 8)
          uint32_t w0, w1;
          pa_ldmatrix_x2_trans(vs + (j >> 3) * BOXB + lm_tok * 128 + ((((j & 7) ^ (lm_tok & 7)) << 4)), w0, w1);
          w0 &= vm_lo;
          w1 &= vm_hi;
          // bytes (t0:d, t0:d+1, t1:d, t1:d+1) -> (t0:d, t1:d | t0:d+1, t1:d+1)
          w0 = __byte_perm(w0, 0, 0x3120);
          w1 = __byte_perm(w1, 0, 0x3120);
          a0 = e4m3x2_to_f16x2((uint16_t)(w0 & 0xffffu));   // row r   <-> d = 16 j + 2 r,     k = tokens 2c, 2c+1
          a1 = e4m3x2_to_f16x2((uint16_t)(w0 >> 16));       // row r+8 <-> d = 16 j + 2 r + 1
          a2 = e4m3x2_to_f16x2((uint16_t)(w1 & 0xffffu));   // k = tokens 8+2c, 9+2c
          a3 = e4m3x2_to_f16x2((uint16_t)(w1 >> 16));
        } else {
          // d tile j = chunks 2j, 2j+1 of every token row (box j / 4): matrices (d 0-7 | tok 0-7), (d 8-15 | tok 0-7),
          // (d 0-7 | tok 8-15), (d 8-15 | tok 8-15) = a0, a1, a2, a3
          const int ch = ((2 * j) & 7) + lm_chunk_add;
          pa_ldmatrix_x4_trans(vs + (j >> 2) * BOXB + lm_tok * 128 + (((ch ^ (lm_tok & 7)) << 4)), a0, a1, a2, a3);
          a0 &= vm_lo;
          a1 &= vm_lo;
          a2 &= vm_hi;
          a3 &= vm_hi;
        }
        M::mma(o[j], a0, a1, a2, a3, b0, b1);
      }
      // the stage's bytes are consumed: hand it back to the producer
      __syncwarp();
      if (lane == 0) pa_mbar_arrive(&empty_bar[stage]);
      stage += NWARPS;
      if (stage >= num_stages) {
        stage -= num_stages;
        phase ^= 1u;
      }
    }
  }

  // ---- per-warp totals -> shared memory (the ring is idle from here on) -----------------------------------
  __syncthreads();
  float* sm_o = reinterpret_cast<float*>(smem_al);
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
      // accumulator rows r / r + 8 <-> d (fp8: 16j + 2r / + 1;  16-bit: 16j + r / + 8)
      const int da = FP8 ? 16 * j + 2 * r : 16 * j + r;
      const int db = FP8 ? da + 1 : da + 8;
      so[(2 * c) * HEAD_SIZE + da] = o[j][0];
      so[(2 * c + 1) * HEAD_SIZE + da] = o[j][1];
      so[(2 * c) * HEAD_SIZE + db] = o[j][2];
      so[(2 * c + 1) * HEAD_SIZE + db] = o[j][3];
    }
  }
  __syncthreads();

  // ---- merge warps, normalise, write (identical to the paged-layout kernels) --------------------------------
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

// shared-memory bytes of the flash-layout kernel for a ring of `stages` (1 KiB of alignment slack included)
template <int HEAD_SIZE, int NWARPS>
static inline size_t pa_flash_smem_bytes(int stages, int cache_elem_bytes, int partition_tokens) {
  const size_t stage = (size_t)2 * HEAD_SIZE * 16 * cache_elem_bytes;
  const size_t merge = (size_t)NWARPS * 8 * HEAD_SIZE * 4 + 2 * NWARPS * 8 * 4;
  size_t data = stages * stage > merge ? stages * stage : merge;
  data = (data + 127) & ~(size_t)127;
  return 1024 + data + 2 * kPaMaxStages * 8 + (size_t)((partition_tokens + 15) / 16) * 4;
}

}  // namespace vb200
