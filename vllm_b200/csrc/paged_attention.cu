// Paged-attention decode for sm_100a (paged_attention_v1 / v2 semantics of
// /root/reference/csrc/attention/attention_kernels.cuh:81-664), designed for HBM streaming:
//
//  * one CTA per (sequence, kv-head [x group of 8 query heads], split-K partition): all query
//    heads of a GQA group share ONE pass over the KV bytes (the reference re-reads K/V once per
//    query head);
//  * K/V go HBM -> registers with 16-byte / 8-byte streaming loads issued one 16-token tile ahead
//    of the math (register double buffer), several CTAs per SM keep >= 64 KB in flight per SM;
//  * QK^T and PV run on the warp-level tensor-core path (mma.sync m16n8k16, fp32 accumulate) with
//    tokens on the M side and the (<= 8) query heads on the N side, so there is no padding waste
//    for GQA-8 and the fp8 -> f16 dequant (cvt.f16x2.e4m3x2) is the only ALU work per byte;
//    k_scale / v_scale are folded into the fp32 logits / output (exact products, one rounding
//    fewer than the reference's bf16(k * scale));
//  * the vLLM cache layouts are consumed as-is: a K "x-run" [token][x] is one 16-byte load, a V
//    row [d][tokens] gives 4 consecutive tokens per load; a fixed permutation of the token and
//    head-dim order inside a tile (legal because both are summed over) makes those loads line up
//    with the MMA fragment layouts without any shared-memory transpose;
//  * online softmax in the exp2 domain, per-warp (m, l, O) merged through shared memory, then
//    either the final output or the v2 partials (normalised tmp_out, exp_sums, max_logits) with
//    the reference's +1e-6 normalisers (attention_kernels.cuh:337,647).
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "tc_common.cuh"

namespace vb200 {

struct PagedAttnParams {
  void* out;          // [num_seqs, num_heads, head_size]
  float* exp_sums;    // [num_seqs, num_heads, max_num_partitions]   (partitioned only)
  float* max_logits;  // same
  void* tmp_out;      // [num_seqs, num_heads, max_num_partitions, head_size]
  const void* q;
  const void* k_cache;
  const void* v_cache;
  const int* block_tables;
  const int* seq_lens;
  const float* alibi_slopes;
  const float* k_scale;
  const float* v_scale;
  int num_heads, num_kv_heads, max_num_blocks_per_seq;
  int partition_size;      // tokens; multiple of 16 and of BLOCK_SIZE
  int max_num_partitions;  // 1 => write `out` directly (unless force_partials)
  int force_partials = 0;  // paged_attention_v2: the caller's partial buffers are always filled, as the reference does
  int64_t q_stride, kv_block_stride, kv_head_stride;
  float scale;
  int prefetch_dist;  // tiles (per warp) ahead of the register double buffer that are pulled into L2; 0 = off
  // fused output quantisation (vb200_paged_attention_v1_quant): the split-K reduce kernel writes fp8 + per-token scales
  void* out_q = nullptr;
  float* out_scales = nullptr;
  const float* scale_ub = nullptr;
  // flash-layout kernel only: sliding window, a query at position seq_len - 1 attends keys seq_len - 1 - window_left ..
  // seq_len - 1 (flash-attn's window_size = (window_left, 0)); < 0: none.  Partitions are counted from the 16-token
  // block that holds the first visible key.
  int window_left = -1;
  float softcap = 0.f;   // flash-layout kernel only: logits = softcap * tanh(logits / softcap); 0: off
};

template <typename T>
struct MmaType;
template <>
struct MmaType<__half> {
  static __device__ __forceinline__ void mma(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                             uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
    __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
  }
};
template <>
struct MmaType<__nv_bfloat16> {
  static __device__ __forceinline__ void mma(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                             uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
  }
};

__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t v) {
  uint32_t r;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(r) : "r"(v));
  return r;
}

__device__ __forceinline__ float fast_exp2(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// QT: query/output element type (__half / __nv_bfloat16).  FP8: cache holds e4m3 bytes (x = 16),
// otherwise the cache element type is QT (x = 8).
template <typename QT, bool FP8, int HEAD_SIZE, int BLOCK_SIZE, int NWARPS>
__global__ void __launch_bounds__(NWARPS * 32, (HEAD_SIZE <= 128 ? (FP8 ? 3 : 2) : 1))
    paged_attention_kernel(const PagedAttnParams p) {
  pdl_wait();
  pdl_trigger();
  using MT = typename std::conditional<FP8, __half, QT>::type;  // tensor-core operand type
  using M = MmaType<MT>;
  constexpr int X = FP8 ? 16 : 8;                 // elements per 16-byte K run
  constexpr int CE = FP8 ? 1 : 2;                 // cache element bytes
  constexpr int NCHUNK = HEAD_SIZE / X;           // K runs per token
  constexpr int KI = (NCHUNK + 3) / 4;            // runs per thread (chunk = c + 4 i)
  constexpr int KJ = X / 4;                       // MMAs per run
  constexpr int NJ = (HEAD_SIZE + 15) / 16;       // 16-row d tiles for PV
  constexpr int VW = FP8 ? 1 : 2;                 // 32-bit words per V load (4 tokens)
  static_assert(HEAD_SIZE % X == 0, "head size must be a multiple of the K run");
  static_assert(BLOCK_SIZE % 4 == 0, "block size must be a multiple of 4");

  const int seq_idx = blockIdx.y;
  const int part_idx = blockIdx.z;
  const int num_q_per_kv = p.num_heads / p.num_kv_heads;
  const int groups_per_kv = (num_q_per_kv + 7) >> 3;
  const int kv_head = blockIdx.x / groups_per_kv;
  const int hgroup = blockIdx.x - kv_head * groups_per_kv;
  const int head0 = kv_head * num_q_per_kv + hgroup * 8;      // first query head of this CTA
  const int nheads = min(8, num_q_per_kv - hgroup * 8);       // valid heads in this CTA

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

  extern __shared__ __align__(16) unsigned char smem_raw[];
  // layout: [NWARPS][8][HEAD_SIZE] fp32 O | [NWARPS][8] m | [NWARPS][8] l | block table ints
  float* sm_o = reinterpret_cast<float*>(smem_raw);
  float* sm_m = sm_o + NWARPS * 8 * HEAD_SIZE;
  float* sm_l = sm_m + NWARPS * 8;
  int* sm_bt = reinterpret_cast<int*>(sm_l + NWARPS * 8);

  // ---- stage this partition's slice of the block table ---------------------------------------
  const int* bt = p.block_tables + (int64_t)seq_idx * p.max_num_blocks_per_seq;
  const int blk_begin = tok_begin / BLOCK_SIZE;
  const int nblk = (ntok + BLOCK_SIZE - 1) / BLOCK_SIZE;
  for (int i = tid; i < nblk; i += NWARPS * 32) sm_bt[i] = bt[blk_begin + i];

  // ---- query fragments (B operand of QK^T): thread (r, c) holds head r, runs c + 4i ----------
  uint32_t qf[KI][KJ][2];
  {
    const QT* qp = reinterpret_cast<const QT*>(p.q) + (int64_t)seq_idx * p.q_stride +
                   (int64_t)(head0 + r) * HEAD_SIZE;
    const bool hvalid = r < nheads;
#pragma unroll
    for (int i = 0; i < KI; ++i) {
      const int chunk = c + 4 * i;
#pragma unroll
      for (int j = 0; j < KJ; ++j) {
        uint32_t w0 = 0, w1 = 0;
        if (hvalid && chunk < NCHUNK) {
          const QT* e = qp + chunk * X + 4 * j;
          if (std::is_same<QT, MT>::value) {
            uint2 v = *reinterpret_cast<const uint2*>(e);
            w0 = v.x;
            w1 = v.y;
          } else {
            w0 = M::pack(to_float<QT>(e[0]), to_float<QT>(e[1]));
            w1 = M::pack(to_float<QT>(e[2]), to_float<QT>(e[3]));
          }
        }
        qf[i][j][0] = w0;
        qf[i][j][1] = w1;
      }
    }
  }

  // logits are kept in the log2 domain: s2 = (q.k) * scale * k_scale * log2(e) (+ alibi * log2e)
  float sl2 = p.scale * kLog2e;
  if (FP8) sl2 *= *p.k_scale;
  float alibi2[2] = {0.f, 0.f};
  if (p.alibi_slopes != nullptr) {
#pragma unroll
    for (int h = 0; h < 2; ++h)
      if (2 * c + h < nheads) alibi2[h] = p.alibi_slopes[head0 + 2 * c + h] * kLog2e;
  }

  const unsigned char* kbase = reinterpret_cast<const unsigned char*>(p.k_cache) +
                               (int64_t)kv_head * p.kv_head_stride * CE;
  const unsigned char* vbase = reinterpret_cast<const unsigned char*>(p.v_cache) +
                               (int64_t)kv_head * p.kv_head_stride * CE;
  const int64_t blk_stride_b = p.kv_block_stride * CE;

  // per-thread token slots inside a 16-token tile (see header comment):
  //   QK^T rows: slot r <-> token t0 = 4*(r/2) + (r&1); slot r+8 <-> t0 + 2
  //   PV k-slots of lane c <-> tokens 4c .. 4c+3
  const int t0 = ((r >> 1) << 2) + (r & 1);

  __syncthreads();  // block table staged

  struct Tile {
    uint4 ka[KI], kb[KI];
    uint32_t va[NJ][VW], vb[NJ][VW];
  };

  auto load_tile = [&](Tile& t, int tile) {
    const int T0 = tile << 4;  // token offset inside the partition
    // K rows
    const int ta = T0 + t0, tb = ta + 2;
    const bool va_ok = ta < ntok, vb_ok = tb < ntok;
    const int64_t pba = va_ok ? sm_bt[ta / BLOCK_SIZE] : 0;
    const unsigned char* ka = kbase + pba * blk_stride_b + ((ta % BLOCK_SIZE) << 4);
    const unsigned char* kb = ka + 32;  // tb = ta + 2 is in the same block (BLOCK_SIZE % 4 == 0)
#pragma unroll
    for (int i = 0; i < KI; ++i) {
      const int chunk = c + 4 * i;
      const bool cok = chunk < NCHUNK;
      t.ka[i] = (va_ok && cok) ? ldg_stream_v4(ka + chunk * (BLOCK_SIZE * 16)) : make_uint4(0, 0, 0, 0);
      t.kb[i] = (vb_ok && cok) ? ldg_stream_v4(kb + chunk * (BLOCK_SIZE * 16)) : make_uint4(0, 0, 0, 0);
    }
    // V rows: tokens T0 + 4c .. +3 of rows d = 16 j + r and + 8
    const int tv = T0 + 4 * c;
    const bool v_ok = tv < ntok;
    const int64_t pbv = v_ok ? sm_bt[tv / BLOCK_SIZE] : 0;
    const unsigned char* vp = vbase + pbv * blk_stride_b + (tv % BLOCK_SIZE) * CE;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int da = 16 * j + r, db = da + 8;
      if (VW == 2) {
        uint2 a = (v_ok && da < HEAD_SIZE) ? ldg_stream_v2(vp + da * (BLOCK_SIZE * CE)) : make_uint2(0, 0);
        uint2 b = (v_ok && db < HEAD_SIZE) ? ldg_stream_v2(vp + db * (BLOCK_SIZE * CE)) : make_uint2(0, 0);
        t.va[j][0] = a.x;
        t.va[j][VW - 1] = a.y;
        t.vb[j][0] = b.x;
        t.vb[j][VW - 1] = b.y;
      } else {
        t.va[j][0] = (v_ok && da < HEAD_SIZE) ? ldg_stream_u32(vp + da * (BLOCK_SIZE * CE)) : 0u;
        t.vb[j][0] = (v_ok && db < HEAD_SIZE) ? ldg_stream_u32(vp + db * (BLOCK_SIZE * CE)) : 0u;
      }
    }
  };

  // L2 prefetch of a whole 16-token tile (K rows and V rows of this kv head): the tile's bytes are
  // HEAD_SIZE * 16 * CE of K (contiguous in one cache block when BLOCK_SIZE == 16, else strided
  // runs) + the same of V; one 128-byte line per lane and per round.  It raises the bytes in flight
  // per SM beyond what the register double buffer holds, at no register cost.
  auto prefetch_tile = [&](int tile) {
    const int T0 = tile << 4;
    if (T0 >= ntok) return;
    const int64_t pb = sm_bt[T0 / BLOCK_SIZE];
    const int tin = T0 % BLOCK_SIZE;
    constexpr int kRowBytes = BLOCK_SIZE * 16;              // one K run-row [BLOCK_SIZE][x] in bytes
    constexpr int kKLines = (NCHUNK * 16 * 16 + 127) / 128;  // 128-byte lines of K in a 16-token tile
    constexpr int kVLines = (HEAD_SIZE * 16 * CE + 127) / 128;
    const unsigned char* kb_ = kbase + pb * blk_stride_b;
    const unsigned char* vb_ = vbase + pb * blk_stride_b;
    for (int l = lane; l < kKLines + kVLines; l += 32) {
      const unsigned char* a;
      if (l < kKLines) {
        // line l covers bytes [128 l, 128 l + 128) of the tile's K, laid out chunk-major: 256 bytes per chunk
        const int chunk = (l * 128) / 256, within = (l * 128) % 256;
        a = kb_ + chunk * kRowBytes + tin * 16 + within;
      } else {
        const int lv = l - kKLines;
        // V rows are [d][BLOCK_SIZE] elements: 16 tokens of row d = 16 * CE bytes
        const int d = (lv * 128) / (16 * CE), within = (lv * 128) % (16 * CE);
        a = vb_ + (d * BLOCK_SIZE + tin) * CE + within;
      }
      asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
    }
  };

  float o[NJ][4];
#pragma unroll
  for (int j = 0; j < NJ; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY};  // running max (log2 domain) of heads 2c, 2c+1
  float l_run[2] = {0.f, 0.f};              // this thread's partial sums

  auto compute_tile = [&](Tile& t, int tile) {
    const int T0 = tile << 4;
    // ---- S^T = K Q^T ---------------------------------------------------------------------------
    float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < KI; ++i) {
      const uint32_t kaw[4] = {t.ka[i].x, t.ka[i].y, t.ka[i].z, t.ka[i].w};
      const uint32_t kbw[4] = {t.kb[i].x, t.kb[i].y, t.kb[i].z, t.kb[i].w};
#pragma unroll
      for (int j = 0; j < KJ; ++j) {
        uint32_t a0, a1, a2, a3;
        if (FP8) {
          a0 = e4m3x2_to_f16x2((uint16_t)(kaw[j] & 0xffffu));
          a2 = e4m3x2_to_f16x2((uint16_t)(kaw[j] >> 16));
          a1 = e4m3x2_to_f16x2((uint16_t)(kbw[j] & 0xffffu));
          a3 = e4m3x2_to_f16x2((uint16_t)(kbw[j] >> 16));
        } else {
          a0 = kaw[2 * j];
          a2 = kaw[2 * j + 1];
          a1 = kbw[2 * j];
          a3 = kbw[2 * j + 1];
        }
        M::mma(s, a0, a1, a2, a3, qf[i][j][0], qf[i][j][1]);
      }
    }
    // ---- scale, alibi, mask --------------------------------------------------------------------
    const int tok_a = T0 + t0, tok_b = tok_a + 2;  // partition-relative
    {
      const float pa = (float)(tok_begin + tok_a - seq_len + 1);
      const float pb = (float)(tok_begin + tok_b - seq_len + 1);
      s[0] = (tok_a < ntok) ? fmaf(s[0], sl2, alibi2[0] * pa) : -INFINITY;
      s[1] = (tok_a < ntok) ? fmaf(s[1], sl2, alibi2[1] * pa) : -INFINITY;
      s[2] = (tok_b < ntok) ? fmaf(s[2], sl2, alibi2[0] * pb) : -INFINITY;
      s[3] = (tok_b < ntok) ? fmaf(s[3], sl2, alibi2[1] * pb) : -INFINITY;
    }
    // ---- online softmax: max over the 16 token rows (8 r-lanes x 2) ------------------------------
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
    // P^T as B fragments: transpose the (token-slot x head) 8x8 blocks across the warp
    const uint32_t b0 = movmatrix_trans(M::pack(p0, p1));
    const uint32_t b1 = movmatrix_trans(M::pack(p2, p3));
    // ---- O^T += V^T P^T ------------------------------------------------------------------------
    // zero V of tokens past the end (they may hold NaN/Inf garbage)
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
        const uint32_t wa = t.va[j][0] & vmask0, wb = t.vb[j][0] & vmask0;
        a0 = e4m3x2_to_f16x2((uint16_t)(wa & 0xffffu));
        a2 = e4m3x2_to_f16x2((uint16_t)(wa >> 16));
        a1 = e4m3x2_to_f16x2((uint16_t)(wb & 0xffffu));
        a3 = e4m3x2_to_f16x2((uint16_t)(wb >> 16));
      } else {
        a0 = t.va[j][0] & vmask0;
        a2 = t.va[j][VW - 1] & vmask1;
        a1 = t.vb[j][0] & vmask0;
        a3 = t.vb[j][VW - 1] & vmask1;
      }
      M::mma(o[j], a0, a1, a2, a3, b0, b1);
    }
  };

  // ---- main loop: register double buffer, tiles strided over the warps -------------------------
  {
    Tile ta, tb;
    int tile = warp;
    const int pf = p.prefetch_dist * NWARPS;
    if (pf > 0) {
      for (int t = tile + NWARPS; t < tile + pf && t < ntiles; t += NWARPS) prefetch_tile(t);
    }
    if (tile < ntiles) load_tile(ta, tile);
    while (tile < ntiles) {
      const int n1 = tile + NWARPS;
      if (pf > 0 && tile + pf < ntiles) prefetch_tile(tile + pf);
      if (n1 < ntiles) load_tile(tb, n1);
      compute_tile(ta, tile);
      if (n1 >= ntiles) break;
      const int n2 = n1 + NWARPS;
      if (pf > 0 && n1 + pf < ntiles) prefetch_tile(n1 + pf);
      if (n2 < ntiles) load_tile(ta, n2);
      compute_tile(tb, n1);
      tile = n2;
    }
  }

  // ---- per-warp totals -> shared memory ---------------------------------------------------------
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
  {
    float* so = sm_o + warp * 8 * HEAD_SIZE;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int da = 16 * j + r, db = da + 8;
      if (da < HEAD_SIZE) {
        so[(2 * c) * HEAD_SIZE + da] = o[j][0];
        so[(2 * c + 1) * HEAD_SIZE + da] = o[j][1];
      }
      if (db < HEAD_SIZE) {
        so[(2 * c) * HEAD_SIZE + db] = o[j][2];
        so[(2 * c + 1) * HEAD_SIZE + db] = o[j][3];
      }
    }
  }
  __syncthreads();

  // ---- merge warps, normalise, write --------------------------------------------------------------
  const float v_scale = FP8 ? *p.v_scale : 1.f;
  const bool partitioned = p.max_num_partitions > 1 || p.force_partials;
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
      QT* dst = reinterpret_cast<QT*>(p.tmp_out) +
                (((int64_t)seq_idx * p.num_heads + head) * p.max_num_partitions + part_idx) *
                    HEAD_SIZE;
      dst[d] = from_float<QT>(res);
      if (d == 0) {
        const int64_t o2 = ((int64_t)seq_idx * p.num_heads + head) * p.max_num_partitions + part_idx;
        p.exp_sums[o2] = lsum;
        p.max_logits[o2] = mg * kLn2;  // back to natural-log units, as the reference stores
      }
    } else {
      QT* dst = reinterpret_cast<QT*>(p.out) + ((int64_t)seq_idx * p.num_heads + head) * HEAD_SIZE;
      dst[d] = from_float<QT>(res);
    }
  }
}

// v2 reduce (attention_kernels.cuh:559-664): one warp per (seq, head).
template <typename QT, int HEAD_SIZE>
__global__ void __launch_bounds__(128) paged_attention_reduce_kernel(
    QT* __restrict__ out, const float* __restrict__ exp_sums, const float* __restrict__ max_logits,
    const QT* __restrict__ tmp_out, const int* __restrict__ seq_lens, int num_heads,
    int max_num_partitions, int partition_size, int total_rows, int window_left) {
  pdl_wait();
  pdl_trigger();
  const int row = blockIdx.x * 4 + (threadIdx.x >> 5);  // seq * num_heads + head
  if (row >= total_rows) return;
  const int lane = threadIdx.x & 31;
  const int seq_idx = row / num_heads;
  int seq_len = seq_lens[seq_idx];
  // sliding window (flash-layout kernel): partitions are counted from the block holding the first visible key
  if (window_left >= 0) seq_len -= max(0, seq_len - 1 - window_left) & ~15;
  // seq_len is read on the device and may have grown past the max_seq_len the buffers were sized from (a replayed
  // graph whose seq_lens advance on device): never walk past the row's max_num_partitions entries
  const int np = min(max(1, (seq_len + partition_size - 1) / partition_size), max_num_partitions);
  const float* ml = max_logits + (int64_t)row * max_num_partitions;
  const float* es = exp_sums + (int64_t)row * max_num_partitions;
  const QT* tp = tmp_out + (int64_t)row * max_num_partitions * HEAD_SIZE;
  QT* op = out + (int64_t)row * HEAD_SIZE;
  if (np == 1) {  // copy (attention_kernels.cuh:577-589)
    for (int d = lane; d < HEAD_SIZE; d += 32) op[d] = tp[d];
    return;
  }
  float mg = -INFINITY;
  for (int j = lane; j < np; j += 32) mg = fmaxf(mg, ml[j]);
  mg = warp_max(mg);
  float gsum = 0.f;
  for (int j = lane; j < np; j += 32) gsum += es[j] * __expf(ml[j] - mg);
  gsum = warp_sum(gsum);
  const float inv = __fdividef(1.f, gsum + 1e-6f);
  constexpr int PER = (HEAD_SIZE + 31) / 32;
  float acc[PER];
#pragma unroll
  for (int i = 0; i < PER; ++i) acc[i] = 0.f;
  for (int j = 0; j < np; ++j) {
    const float w = es[j] * __expf(ml[j] - mg) * inv;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      const int d = lane + 32 * i;
      if (d < HEAD_SIZE) acc[i] = fmaf(to_float<QT>(tp[(int64_t)j * HEAD_SIZE + d]), w, acc[i]);
    }
  }
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int d = lane + 32 * i;
    if (d < HEAD_SIZE) op[d] = from_float<QT>(acc[i]);
  }
}

// v2 reduce fused with dynamic_per_token_scaled_fp8_quant of the attention output (the activation quantisation in
// front of o_proj; /root/reference/csrc/quantization/fp8/common.cu:28-83): one CTA per sequence, one warp per head
// (heads beyond 32 loop), the reduced row stays in registers between the reduction, the row amax and the fp8 write.
// Same arithmetic and rounding points as paged_attention_reduce_kernel followed by per_token_quant_kernel: the reduced
// value is rounded to QT first (what the unfused kernel stores), then amax / scale / true division.
// A lane owns PER = HEAD_SIZE / 32 CONSECUTIVE elements of its head (one 8-byte load per partition at head size 128,
// one 4-byte store of the quantised row piece).
template <typename QT, int PER>
__device__ __forceinline__ void load_lane_piece(float (&v)[PER], const QT* p) {
  if constexpr (PER % 4 == 0) {
#pragma unroll
    for (int i = 0; i < PER; i += 4) {
      const uint2 u = *reinterpret_cast<const uint2*>(p + i);
      const QT* e = reinterpret_cast<const QT*>(&u);
#pragma unroll
      for (int j = 0; j < 4; ++j) v[i + j] = to_float<QT>(e[j]);
    }
  } else if constexpr (PER % 2 == 0) {
#pragma unroll
    for (int i = 0; i < PER; i += 2) {
      const uint32_t u = *reinterpret_cast<const uint32_t*>(p + i);
      const QT* e = reinterpret_cast<const QT*>(&u);
      v[i] = to_float<QT>(e[0]);
      v[i + 1] = to_float<QT>(e[1]);
    }
  } else {
#pragma unroll
    for (int i = 0; i < PER; ++i) v[i] = to_float<QT>(p[i]);
  }
}

// HPW: heads per warp.  A warp walks its HPW heads TOGETHER, phase by phase (maxima, sums, partials), so that the
// loads of all its heads are in flight at once; fewer, busier warps keep the CTA small enough for every sequence's
// CTA to be resident in one wave (32 heads: 16 warps, two CTAs per SM).
template <typename QT, int HEAD_SIZE, int HPW>
__global__ void __launch_bounds__(1024) paged_attention_reduce_quant_kernel(
    uint8_t* __restrict__ out_q, float* __restrict__ out_scales, const float* __restrict__ scale_ub,
    const float* __restrict__ exp_sums, const float* __restrict__ max_logits, const QT* __restrict__ tmp_out,
    const int* __restrict__ seq_lens, int num_heads, int max_num_partitions, int partition_size) {
  pdl_wait();
  pdl_trigger();
  constexpr int PER = HEAD_SIZE / 32;
  constexpr int JB = HPW == 4 ? 2 : 4;     // partitions fetched per round
  __shared__ float sm[32];
  const int seq_idx = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int seq_len = seq_lens[seq_idx];
  const int np = min(max(1, (seq_len + partition_size - 1) / partition_size), max_num_partitions);
  float y[HPW][PER];
  const float* ml[HPW];
  const float* es[HPW];
  const QT* tp[HPW];
  bool live[HPW];
#pragma unroll
  for (int hh = 0; hh < HPW; ++hh) {
    const int head = warp + hh * nw;
    live[hh] = head < num_heads;
    const int64_t row = (int64_t)seq_idx * num_heads + (live[hh] ? head : 0);
    ml[hh] = max_logits + row * max_num_partitions;
    es[hh] = exp_sums + row * max_num_partitions;
    tp[hh] = tmp_out + row * max_num_partitions * HEAD_SIZE + lane * PER;
  }
  if (np == 1) {
#pragma unroll
    for (int hh = 0; hh < HPW; ++hh)
      if (live[hh]) load_lane_piece<QT, PER>(y[hh], tp[hh]);
  } else {
    float mg[HPW], gsum[HPW], inv[HPW];
#pragma unroll
    for (int hh = 0; hh < HPW; ++hh) {
      mg[hh] = -INFINITY;
      if (live[hh])
        for (int j = lane; j < np; j += 32) mg[hh] = fmaxf(mg[hh], ml[hh][j]);
    }
#pragma unroll
    for (int hh = 0; hh < HPW; ++hh) mg[hh] = warp_max(mg[hh]);
#pragma unroll
    for (int hh = 0; hh < HPW; ++hh) {
      gsum[hh] = 0.f;
      if (live[hh])
        for (int j = lane; j < np; j += 32) gsum[hh] += es[hh][j] * __expf(ml[hh][j] - mg[hh]);
    }
#pragma unroll
    for (int hh = 0; hh < HPW; ++hh) {
      gsum[hh] = warp_sum(gsum[hh]);
      inv[hh] = __fdividef(1.f, gsum[hh] + 1e-6f);
    }
    float acc[HPW][PER];
#pragma unroll
    for (int hh = 0; hh < HPW; ++hh)
#pragma unroll
      for (int i = 0; i < PER; ++i) acc[hh][i] = 0.f;
    for (int j0 = 0; j0 < np; j0 += JB) {
      float t[HPW][JB][PER];
#pragma unroll
      for (int hh = 0; hh < HPW; ++hh)
#pragma unroll
        for (int jj = 0; jj < JB; ++jj)
          if (live[hh] && j0 + jj < np) load_lane_piece<QT, PER>(t[hh][jj], tp[hh] + (int64_t)(j0 + jj) * HEAD_SIZE);
#pragma unroll
      for (int hh = 0; hh < HPW; ++hh)
#pragma unroll
        for (int jj = 0; jj < JB; ++jj)
          if (live[hh] && j0 + jj < np) {   // same order of accumulation as the unfused reduce kernel: j ascending
            const float w = es[hh][j0 + jj] * __expf(ml[hh][j0 + jj] - mg[hh]) * inv[hh];
#pragma unroll
            for (int i = 0; i < PER; ++i) acc[hh][i] = fmaf(t[hh][jj][i], w, acc[hh][i]);
          }
    }
#pragma unroll
    for (int hh = 0; hh < HPW; ++hh)
#pragma unroll
      for (int i = 0; i < PER; ++i) y[hh][i] = to_float<QT>(from_float<QT>(acc[hh][i]));
  }
  float amax = 0.f;
#pragma unroll
  for (int hh = 0; hh < HPW; ++hh)
    if (live[hh]) {
#pragma unroll
      for (int i = 0; i < PER; ++i) amax = fmaxf(amax, fabsf(y[hh][i]));
    }
  amax = warp_max(amax);
  if (lane == 0) sm[warp] = amax;
  __syncthreads();
  amax = warp_max(lane < nw ? sm[lane] : 0.f);
  float s = scale_ub ? fminf(amax, *scale_ub) : amax;
  s = fmaxf(s / 448.f, 1.0f / (448.f * 512.0f));
  if (threadIdx.x == 0) out_scales[seq_idx] = s;
  const RowDivider div(s, amax);
#pragma unroll
  for (int hh = 0; hh < HPW; ++hh) {
    const int head = warp + hh * nw;
    if (live[hh]) {
      uint8_t* o = out_q + ((int64_t)seq_idx * num_heads + head) * HEAD_SIZE + lane * PER;
      if constexpr (PER % 4 == 0) {
#pragma unroll
        for (int i = 0; i < PER; i += 4)
          *reinterpret_cast<uint32_t*>(o + i) =
              e4m3x4_clamped(div(y[hh][i]), div(y[hh][i + 1]), div(y[hh][i + 2]), div(y[hh][i + 3]));
      } else {
#pragma unroll
        for (int i = 0; i < PER; ++i) o[i] = float_to_e4m3(fmaxf(-448.f, fminf(div(y[hh][i]), 448.f)));
      }
    }
  }
}

}  // namespace vb200
#include "paged_attention_tma.cuh"
#include "paged_attention_flash.cuh"
namespace vb200 {

constexpr int kWarps = 4;
static inline int round_up(int a, int b) { return (a + b - 1) / b * b; }

// bulk-copy load path (block size 16, head size % 16 == 0) unless VB200_PA_TMA=0
static bool pa_use_tma() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("VB200_PA_TMA");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}
// resident CTAs per SM the bulk-copy kernel is compiled for: bandwidth no longer depends on the warp count, so
// more (smaller) CTAs only make the waves finer; 96 registers/thread allow 4 CTAs of 160 threads up to head 128
static inline int pa_tma_ctas_per_sm(int head_size) { return head_size <= 128 ? 4 : 2; }
// ring depth: fill the shared memory of an SM shared by `ctas` CTAs.  MUST be a multiple of the consumer warp
// count: then a stage is always consumed by the same warp, whose program order guarantees it never waits on a
// full barrier that is still one phase behind (parity aliasing).
template <int HEAD_SIZE>
static inline int pa_tma_stages(int cache_elem_bytes, int partition_tokens) {
  const int stage = 2 * HEAD_SIZE * 16 * cache_elem_bytes;
  const int fixed = 2 * kPaMaxStages * 8 + ((partition_tokens + 15) / 16) * 4 + 1024 /* per-CTA reservation */ + 128;
  int budget = (227 * 1024) / pa_tma_ctas_per_sm(HEAD_SIZE) - fixed;
  int st = budget / stage;
  const char* e = getenv("VB200_PA_STAGES");
  if (e && atoi(e) > 0) st = atoi(e);
  if (st > kPaMaxStages) st = kPaMaxStages;
  st = st / kWarps * kWarps;
  if (st < kWarps) st = kWarps;
  return st;
}

template <typename QT, bool FP8, int HEAD_SIZE, int BLOCK_SIZE>
static int launch_pa(const PagedAttnParams& p, int num_seqs, int num_partitions,
                     cudaStream_t stream) {
  const int q_per_kv_ = p.num_heads / p.num_kv_heads;
  if (BLOCK_SIZE == 16 && HEAD_SIZE % 16 == 0 && pa_use_tma()) {
    constexpr int HS = HEAD_SIZE % 16 == 0 ? HEAD_SIZE : 128;  // (only instantiated for multiples of 16)
    auto kt = paged_attention_tma_kernel<QT, FP8, HS, kWarps, (HS <= 128 ? 4 : 2)>;
    const int ce = FP8 ? 1 : 2;
    const int stages = pa_tma_stages<HS>(ce, p.partition_size);
    const size_t smem_t = pa_tma_smem_bytes<HS, kWarps>(stages, ce, p.partition_size);
    VB_CHECK(smem_t <= 227 * 1024, "paged attention: partition of %d tokens needs too much shared memory",
             p.partition_size);
    if (smem_t > 48 * 1024)
      VB_CUDA_CHECK(cudaFuncSetAttribute(kt, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_t));
    dim3 grid_t(p.num_kv_heads * ((q_per_kv_ + 7) / 8), num_seqs, num_partitions);
    VB_CUDA_CHECK(launch_k(kt, dim3(grid_t), dim3((kWarps + 1) * 32), smem_t, stream, p, stages));
    VB_LAUNCH_CHECK();
    if (p.out_q != nullptr) {
      // fused output quantisation: always through the partials (force_partials is set by the entry point)
      if constexpr (HEAD_SIZE % 32 == 0) {
        VB_CHECK(p.num_heads <= 128, "paged_attention_v1_quant: at most 128 query heads (got %d)", p.num_heads);
        const int hpw = p.num_heads <= 8 ? 1 : (p.num_heads <= 64 ? 2 : 4);
        const int warps = (p.num_heads + hpw - 1) / hpw;
#define VB_REDUCE_Q(HPW)                                                                                              \
  VB_CUDA_CHECK(launch_k(paged_attention_reduce_quant_kernel<QT, HEAD_SIZE, HPW>, dim3(num_seqs), dim3(32 * warps), 0, \
                         stream, (uint8_t*)p.out_q, p.out_scales, p.scale_ub, p.exp_sums, p.max_logits,                \
                         (const QT*)p.tmp_out, p.seq_lens, p.num_heads, p.max_num_partitions, p.partition_size))
        if (hpw == 1) VB_REDUCE_Q(1);
        else if (hpw == 2) VB_REDUCE_Q(2);
        else VB_REDUCE_Q(4);
#undef VB_REDUCE_Q
        VB_LAUNCH_CHECK();
        return 0;
      } else {
        VB_CHECK(false, "paged_attention_v1_quant: head size %d must be a multiple of 32", HEAD_SIZE);
      }
    }
    if (p.max_num_partitions > 1 || p.force_partials) {
      const int rows = num_seqs * p.num_heads;
      VB_CUDA_CHECK(launch_k(paged_attention_reduce_kernel<QT, HEAD_SIZE>, dim3((rows + 3) / 4), dim3(128), 0, stream, (QT*)p.out, p.exp_sums, p.max_logits, (const QT*)p.tmp_out, p.seq_lens, p.num_heads,
          p.max_num_partitions, p.partition_size, rows, -1));
      VB_LAUNCH_CHECK();
    }
    return 0;
  }
  VB_CHECK(p.out_q == nullptr, "paged_attention_v1_quant: needs block size 16 and a head size that is a multiple of 32");
  auto kern = paged_attention_kernel<QT, FP8, HEAD_SIZE, BLOCK_SIZE, kWarps>;
  const int nblk = (p.partition_size + BLOCK_SIZE - 1) / BLOCK_SIZE;
  const size_t smem = (size_t)kWarps * 8 * HEAD_SIZE * 4 + 2 * kWarps * 8 * 4 + (size_t)nblk * 4;
  if (smem > 48 * 1024) {
    VB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  const int q_per_kv = p.num_heads / p.num_kv_heads;
  const int groups = (q_per_kv + 7) / 8;
  dim3 grid(p.num_kv_heads * groups, num_seqs, num_partitions);
  VB_CUDA_CHECK(launch_k(kern, dim3(grid), dim3(kWarps * 32), smem, stream, p));
  VB_LAUNCH_CHECK();
  if (p.max_num_partitions > 1 || p.force_partials) {
    const int rows = num_seqs * p.num_heads;
    VB_CUDA_CHECK(launch_k(paged_attention_reduce_kernel<QT, HEAD_SIZE>, dim3((rows + 3) / 4), dim3(128), 0, stream, (QT*)p.out, p.exp_sums, p.max_logits, (const QT*)p.tmp_out, p.seq_lens, p.num_heads,
        p.max_num_partitions, p.partition_size, rows, -1));
    VB_LAUNCH_CHECK();
  }
  return 0;
}

template <typename QT, bool FP8, int HEAD_SIZE>
static int dispatch_block(const PagedAttnParams& p, int block_size, int num_seqs, int np,
                          cudaStream_t s) {
  switch (block_size) {
    case 8: return launch_pa<QT, FP8, HEAD_SIZE, 8>(p, num_seqs, np, s);
    case 16: return launch_pa<QT, FP8, HEAD_SIZE, 16>(p, num_seqs, np, s);
    case 32: return launch_pa<QT, FP8, HEAD_SIZE, 32>(p, num_seqs, np, s);
    default: VB_CHECK(false, "Unsupported block size: %d", block_size);
  }
  return 0;
}

template <typename QT, bool FP8>
static int dispatch_head(const PagedAttnParams& p, int head_size, int block_size, int num_seqs,
                         int np, cudaStream_t s) {
  switch (head_size) {
    case 32: return dispatch_block<QT, FP8, 32>(p, block_size, num_seqs, np, s);
    case 64: return dispatch_block<QT, FP8, 64>(p, block_size, num_seqs, np, s);
    case 80: return dispatch_block<QT, FP8, 80>(p, block_size, num_seqs, np, s);
    case 96: return dispatch_block<QT, FP8, 96>(p, block_size, num_seqs, np, s);
    case 112: return dispatch_block<QT, FP8, 112>(p, block_size, num_seqs, np, s);
    case 128: return dispatch_block<QT, FP8, 128>(p, block_size, num_seqs, np, s);
    case 192: return dispatch_block<QT, FP8, 192>(p, block_size, num_seqs, np, s);
    case 256: return dispatch_block<QT, FP8, 256>(p, block_size, num_seqs, np, s);
    default: break;
  }
  if (!FP8 && head_size == 120) {
    // 120 is not a multiple of the fp8 run (16); the reference's fp8 tests skip it as well.
    return dispatch_block<QT, false, 120>(p, block_size, num_seqs, np, s);
  }
  VB_CHECK(false, "Unsupported head size: %d", head_size);
  return 0;
}

static int run_paged_attention(PagedAttnParams& p, int num_seqs, int head_size, int block_size,
                               int max_seq_len, int q_dtype, int kv_dtype, cudaStream_t stream) {
  VB_CHECK(q_dtype == VB200_F16 || q_dtype == VB200_BF16,
           "paged attention: query dtype must be f16 or bf16 (got %d)", q_dtype);
  VB_CHECK(kv_dtype == VB200_KV_AUTO || kv_dtype == VB200_KV_FP8_E4M3,
           "Unsupported data type of kv cache: %d", kv_dtype);
  VB_CHECK(p.num_kv_heads > 0 && p.num_heads % p.num_kv_heads == 0,
           "num_heads %d must be a multiple of num_kv_heads %d", p.num_heads, p.num_kv_heads);
  VB_CHECK(p.partition_size % 16 == 0 && p.partition_size % block_size == 0,
           "partition size %d must be a multiple of 16 and of the block size", p.partition_size);
  if (num_seqs <= 0) return 0;
  const bool fp8 = kv_dtype == VB200_KV_FP8_E4M3;
  VB_CHECK(!fp8 || (p.k_scale && p.v_scale), "fp8 kv cache needs k_scale and v_scale");
  (void)max_seq_len;
  const int np = p.max_num_partitions;
  if (q_dtype == VB200_F16) {
    return fp8 ? dispatch_head<__half, true>(p, head_size, block_size, num_seqs, np, stream)
               : dispatch_head<__half, false>(p, head_size, block_size, num_seqs, np, stream);
  }
  return fp8 ? dispatch_head<__nv_bfloat16, true>(p, head_size, block_size, num_seqs, np, stream)
             : dispatch_head<__nv_bfloat16, false>(p, head_size, block_size, num_seqs, np, stream);
}

// ---- flash cache layout (NHD / HND views through strides): paged_attention_flash.cuh ---------------------------
struct FlashLayout {
  int64_t block_stride, token_stride, head_stride;  // elements
  int num_blocks;
};

template <typename QT, bool FP8, int HEAD_SIZE>
static int launch_pa_flash(const PagedAttnParams& p, const FlashLayout& L, int num_seqs, int num_partitions,
                           cudaStream_t stream) {
  constexpr int CE = FP8 ? 1 : 2;
  static_assert((HEAD_SIZE * CE) % 128 == 0, "flash-layout kernel: head bytes must be a multiple of 128");
  auto kt = paged_attention_flash_kernel<QT, FP8, HEAD_SIZE, kWarps, (HEAD_SIZE <= 128 ? 4 : 2)>;
  // same ring sizing as the paged-layout kernel, minus the 1 KiB the swizzled boxes need for alignment
  int stages = pa_tma_stages<HEAD_SIZE>(CE, p.partition_size);
  size_t smem = pa_flash_smem_bytes<HEAD_SIZE, kWarps>(stages, CE, p.partition_size);
  const size_t per_cta = (size_t)(227 * 1024) / pa_tma_ctas_per_sm(HEAD_SIZE) - 1024;
  while (smem > per_cta && stages > kWarps) {
    stages -= kWarps;
    smem = pa_flash_smem_bytes<HEAD_SIZE, kWarps>(stages, CE, p.partition_size);
  }
  VB_CHECK(smem <= 227 * 1024, "paged attention (flash layout): partition of %d tokens needs too much shared memory",
           p.partition_size);
  CUtensorMap tk, tv;
  const uint64_t dims[4] = {(uint64_t)HEAD_SIZE * CE, 16, (uint64_t)p.num_kv_heads, (uint64_t)L.num_blocks};
  const uint64_t strides[3] = {(uint64_t)L.token_stride * CE, (uint64_t)L.head_stride * CE,
                               (uint64_t)L.block_stride * CE};
  if (tc::make_tmap_4d_u8(&tk, p.k_cache, dims, strides, 16)) return 1;
  if (tc::make_tmap_4d_u8(&tv, p.v_cache, dims, strides, 16)) return 1;
  if (smem > 48 * 1024)
    VB_CUDA_CHECK(cudaFuncSetAttribute(kt, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int q_per_kv = p.num_heads / p.num_kv_heads;
  dim3 grid(p.num_kv_heads * ((q_per_kv + 7) / 8), num_seqs, num_partitions);
  VB_CUDA_CHECK(launch_k(kt, grid, dim3((kWarps + 1) * 32), smem, stream, tk, tv, p, stages));
  VB_LAUNCH_CHECK();
  if (p.max_num_partitions > 1 || p.force_partials) {
    const int rows = num_seqs * p.num_heads;
    VB_CUDA_CHECK(launch_k(paged_attention_reduce_kernel<QT, HEAD_SIZE>, dim3((rows + 3) / 4), dim3(128), 0, stream,
                           (QT*)p.out, p.exp_sums, p.max_logits, (const QT*)p.tmp_out, p.seq_lens, p.num_heads,
                           p.max_num_partitions, p.partition_size, rows, p.window_left));
    VB_LAUNCH_CHECK();
  }
  return 0;
}

template <typename QT>
static int dispatch_flash(const PagedAttnParams& p, const FlashLayout& L, int head_size, bool fp8, int num_seqs, int np,
                          cudaStream_t s) {
  if (fp8) {
    if (head_size == 128) return launch_pa_flash<QT, true, 128>(p, L, num_seqs, np, s);
    if (head_size == 256) return launch_pa_flash<QT, true, 256>(p, L, num_seqs, np, s);
    VB_CHECK(false, "paged_attention_flash: fp8 KV cache supports head sizes 128 and 256 (got %d)", head_size);
  }
  switch (head_size) {
    case 64: return launch_pa_flash<QT, false, 64>(p, L, num_seqs, np, s);
    case 128: return launch_pa_flash<QT, false, 128>(p, L, num_seqs, np, s);
    case 192: return launch_pa_flash<QT, false, 192>(p, L, num_seqs, np, s);
    case 256: return launch_pa_flash<QT, false, 256>(p, L, num_seqs, np, s);
    default: break;
  }
  VB_CHECK(false, "paged_attention_flash: head size %d is not supported (64, 128, 192, 256)", head_size);
  return 0;
}

}  // namespace vb200

using namespace vb200;

// tuning knob (tiles per warp prefetched into L2 ahead of the register pipeline); VB200_PA_PREFETCH overrides
static int pa_prefetch_dist() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("VB200_PA_PREFETCH");
    v = e ? atoi(e) : 0;
    if (v < 0) v = 0;
    if (v > 64) v = 64;
  }
  return v;
}

extern "C" int vb200_paged_attention_v2(
    void* out, float* exp_sums, float* max_logits, void* tmp_out, const void* query,
    const void* key_cache, const void* value_cache, int num_seqs, int num_heads, int head_size,
    int num_kv_heads, float scale, const int32_t* block_tables, const int32_t* seq_lens,
    int block_size, int max_seq_len, int max_num_blocks_per_seq, const float* alibi_slopes,
    int q_dtype, int kv_dtype, const float* k_scale, const float* v_scale, int64_t q_stride,
    int64_t kv_block_stride, int64_t kv_head_stride, int partition_size, vb200_stream_t stream) {
  VB_CHECK(partition_size > 0, "partition_size must be positive");
  PagedAttnParams p;
  p.out = out;
  p.exp_sums = exp_sums;
  p.max_logits = max_logits;
  p.tmp_out = tmp_out;
  p.q = query;
  p.k_cache = key_cache;
  p.v_cache = value_cache;
  p.block_tables = block_tables;
  p.seq_lens = seq_lens;
  p.alibi_slopes = alibi_slopes;
  p.k_scale = k_scale;
  p.v_scale = v_scale;
  p.num_heads = num_heads;
  p.num_kv_heads = num_kv_heads;
  p.max_num_blocks_per_seq = max_num_blocks_per_seq;
  p.partition_size = partition_size;
  p.max_num_partitions = (max_seq_len + partition_size - 1) / partition_size;
  if (p.max_num_partitions < 1) p.max_num_partitions = 1;
  p.q_stride = q_stride;
  p.kv_block_stride = kv_block_stride;
  p.kv_head_stride = kv_head_stride;
  p.scale = scale;
  p.prefetch_dist = pa_prefetch_dist();
  // exp_sums / max_logits / tmp_out are always written, also with a single partition (the reference's v2 kernel
  // fills them and its reduce kernel copies tmp_out -> out, attention_kernels.cuh:337,577-589,647)
  p.force_partials = 1;
  return run_paged_attention(p, num_seqs, head_size, block_size, max_seq_len, q_dtype, kv_dtype,
                             (cudaStream_t)stream);
}

// resident CTAs per SM of the main kernel (its __launch_bounds__ minimum)
static inline int pa_ctas_per_sm(int head_size, int kv_dtype) {
  return head_size <= 128 ? (kv_dtype == VB200_KV_FP8_E4M3 ? 3 : 2) : 1;
}

extern "C" int vb200_paged_attention_v1_partition_size(int num_seqs, int num_heads,
                                                       int num_kv_heads, int head_size,
                                                       int max_seq_len, int block_size,
                                                       int kv_dtype) {
  // Split-K cost model: time ~ waves(np) * (tokens per partition + fixed per-CTA overhead), where
  // a wave is num_sms x resident CTAs.  Picks the np (not only powers of two) that fills whole
  // waves; no split when the unsplit grid already covers the machine.
  if (num_seqs <= 0 || num_kv_heads <= 0) return 0;
  const int q_per_kv = num_heads / num_kv_heads;
  const int groups = (q_per_kv + 7) / 8;
  const long base_ctas = (long)num_seqs * num_kv_heads * groups;
  const bool tma = block_size == 16 && head_size % 16 == 0 && pa_use_tma();
  const long resident = (long)num_sms() * (tma ? pa_tma_ctas_per_sm(head_size) : pa_ctas_per_sm(head_size, kv_dtype));
  const int align = block_size > 32 ? block_size : 32;
  const int len = max_seq_len > 0 ? max_seq_len : 16;
  const long kOverheadTokens = 256;
  long best_cost = -1;
  int best_part = round_up(len, align);
  for (int np = 1; np <= 64; ++np) {
    const int part = round_up((len + np - 1) / np, align);
    if (np > 1 && part < 128) break;
    const int np_eff = (len + part - 1) / part;
    const long waves = (base_ctas * np_eff + resident - 1) / resident;
    const long cost = waves * (part + kOverheadTokens) + (np_eff > 1 ? 64 : 0);
    if (best_cost < 0 || cost < best_cost) {
      best_cost = cost;
      best_part = part;
    }
  }
  return best_part;
}

extern "C" size_t vb200_paged_attention_v1_workspace_size(int num_seqs, int num_heads,
                                                          int head_size, int max_seq_len,
                                                          int q_dtype, int partition_size) {
  (void)q_dtype;
  if (partition_size <= 0 || max_seq_len <= partition_size) return 0;
  const size_t np = (size_t)(max_seq_len + partition_size - 1) / partition_size;
  const size_t rows = (size_t)num_seqs * num_heads * np;
  return rows * 4 * 2 + rows * head_size * 2 + 64;
}

extern "C" int vb200_paged_attention_v1(
    void* out, const void* query, const void* key_cache, const void* value_cache, int num_seqs,
    int num_heads, int head_size, int num_kv_heads, float scale, const int32_t* block_tables,
    const int32_t* seq_lens, int block_size, int max_seq_len, int max_num_blocks_per_seq,
    const float* alibi_slopes, int q_dtype, int kv_dtype, const float* k_scale,
    const float* v_scale, int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
    int partition_size, void* workspace, size_t workspace_bytes, vb200_stream_t stream) {
  int part = partition_size;
  if (workspace == nullptr) {
    part = 0;
  } else if (part <= 0) {
    part = vb200_paged_attention_v1_partition_size(num_seqs, num_heads, num_kv_heads, head_size,
                                                   max_seq_len, block_size, kv_dtype);
  }
  int np = 1;
  if (part > 0 && max_seq_len > part) np = (max_seq_len + part - 1) / part;
  if (np == 1) {
    part = round_up(max_seq_len > 0 ? max_seq_len : 16, 32);
    if (part % block_size) part = round_up(part, block_size);
  }
  PagedAttnParams p;
  p.out = out;
  p.exp_sums = nullptr;
  p.max_logits = nullptr;
  p.tmp_out = nullptr;
  if (np > 1) {
    const size_t need = vb200_paged_attention_v1_workspace_size(num_seqs, num_heads, head_size,
                                                                max_seq_len, q_dtype, part);
    VB_CHECK(workspace_bytes >= need, "paged_attention_v1 workspace too small: %zu < %zu",
             workspace_bytes, need);
    const size_t rows = (size_t)num_seqs * num_heads * np;
    p.exp_sums = reinterpret_cast<float*>(workspace);
    p.max_logits = p.exp_sums + rows;
    p.tmp_out = reinterpret_cast<void*>(p.max_logits + rows);
  }
  p.q = query;
  p.k_cache = key_cache;
  p.v_cache = value_cache;
  p.block_tables = block_tables;
  p.seq_lens = seq_lens;
  p.alibi_slopes = alibi_slopes;
  p.k_scale = k_scale;
  p.v_scale = v_scale;
  p.num_heads = num_heads;
  p.num_kv_heads = num_kv_heads;
  p.max_num_blocks_per_seq = max_num_blocks_per_seq;
  p.partition_size = part;
  p.max_num_partitions = np;
  p.q_stride = q_stride;
  p.kv_block_stride = kv_block_stride;
  p.kv_head_stride = kv_head_stride;
  p.scale = scale;
  p.prefetch_dist = pa_prefetch_dist();
  return run_paged_attention(p, num_seqs, head_size, block_size, max_seq_len, q_dtype, kv_dtype,
                             (cudaStream_t)stream);
}

// Decode attention over the flash cache layout (see include/vllm_b200.h).
extern "C" size_t vb200_paged_attention_flash_workspace_size(int num_seqs, int num_heads, int head_size, int max_seq_len,
                                                             int q_dtype, int partition_size) {
  return vb200_paged_attention_v1_workspace_size(num_seqs, num_heads, head_size, max_seq_len, q_dtype, partition_size);
}

static int pa_flash_impl(
    void* out, const void* query, const void* key_cache, const void* value_cache, int num_seqs, int num_heads,
    int head_size, int num_kv_heads, float scale, const int32_t* block_tables, const int32_t* seq_lens, int block_size,
    int num_blocks, int max_seq_len, int max_num_blocks_per_seq, const float* alibi_slopes, int q_dtype, int kv_dtype,
    const float* k_scale, const float* v_scale, int64_t q_stride, int64_t block_stride, int64_t token_stride,
    int64_t head_stride, int partition_size, void* workspace, size_t workspace_bytes, int window_left, float softcap,
    vb200_stream_t stream) {
  VB_CHECK(softcap >= 0.f && softcap == softcap, "paged_attention_flash: softcap must be >= 0");
  // with a window a sequence contributes at most window_left + 1 keys plus the 15 that may share their first block
  if (window_left >= 0 && max_seq_len > window_left + 16) max_seq_len = window_left + 16;
  VB_CHECK(block_size == 16, "paged_attention_flash: block size must be 16 (got %d)", block_size);
  VB_CHECK(q_dtype == VB200_F16 || q_dtype == VB200_BF16, "paged_attention_flash: query dtype must be f16 or bf16");
  VB_CHECK(kv_dtype == VB200_KV_AUTO || kv_dtype == VB200_KV_FP8_E4M3, "Unsupported data type of kv cache: %d", kv_dtype);
  VB_CHECK(num_kv_heads > 0 && num_heads % num_kv_heads == 0, "num_heads %d must be a multiple of num_kv_heads %d",
           num_heads, num_kv_heads);
  VB_CHECK(num_blocks > 0, "paged_attention_flash: the cache holds no blocks");
  if (num_seqs <= 0) return 0;
  const bool fp8 = kv_dtype == VB200_KV_FP8_E4M3;
  VB_CHECK(!fp8 || (k_scale && v_scale), "fp8 kv cache needs k_scale and v_scale");
  int part = partition_size;
  if (workspace == nullptr) {
    part = 0;
  } else if (part <= 0) {
    part = vb200_paged_attention_v1_partition_size(num_seqs, num_heads, num_kv_heads, head_size, max_seq_len, block_size,
                                                   kv_dtype);
  }
  int np = 1;
  if (part > 0 && max_seq_len > part) np = (max_seq_len + part - 1) / part;
  if (np == 1) part = round_up(max_seq_len > 0 ? max_seq_len : 16, 32);
  VB_CHECK(part % 16 == 0, "paged_attention_flash: partition size %d must be a multiple of 16", part);
  PagedAttnParams p;
  p.out = out;
  p.exp_sums = nullptr;
  p.max_logits = nullptr;
  p.tmp_out = nullptr;
  if (np > 1) {
    const size_t need = vb200_paged_attention_v1_workspace_size(num_seqs, num_heads, head_size, max_seq_len, q_dtype, part);
    VB_CHECK(workspace_bytes >= need, "paged_attention_flash workspace too small: %zu < %zu", workspace_bytes, need);
    const size_t rows = (size_t)num_seqs * num_heads * np;
    p.exp_sums = reinterpret_cast<float*>(workspace);
    p.max_logits = p.exp_sums + rows;
    p.tmp_out = reinterpret_cast<void*>(p.max_logits + rows);
  }
  p.q = query;
  p.k_cache = key_cache;
  p.v_cache = value_cache;
  p.block_tables = block_tables;
  p.seq_lens = seq_lens;
  p.alibi_slopes = alibi_slopes;
  p.k_scale = k_scale;
  p.v_scale = v_scale;
  p.num_heads = num_heads;
  p.num_kv_heads = num_kv_heads;
  p.max_num_blocks_per_seq = max_num_blocks_per_seq;
  p.partition_size = part;
  p.max_num_partitions = np;
  p.q_stride = q_stride;
  p.kv_block_stride = block_stride;
  p.kv_head_stride = head_stride;
  p.scale = scale;
  p.prefetch_dist = 0;
  p.window_left = window_left < 0 ? -1 : window_left;
  p.softcap = softcap;
  FlashLayout L{block_stride, token_stride, head_stride, num_blocks};
  if (q_dtype == VB200_F16) return dispatch_flash<__half>(p, L, head_size, fp8, num_seqs, np, (cudaStream_t)stream);
  return dispatch_flash<__nv_bfloat16>(p, L, head_size, fp8, num_seqs, np, (cudaStream_t)stream);
}

extern "C" int vb200_paged_attention_flash(
    void* out, const void* query, const void* key_cache, const void* value_cache, int num_seqs, int num_heads,
    int head_size, int num_kv_heads, float scale, const int32_t* block_tables, const int32_t* seq_lens, int block_size,
    int num_blocks, int max_seq_len, int max_num_blocks_per_seq, const float* alibi_slopes, int q_dtype, int kv_dtype,
    const float* k_scale, const float* v_scale, int64_t q_stride, int64_t block_stride, int64_t token_stride,
    int64_t head_stride, int partition_size, void* workspace, size_t workspace_bytes, vb200_stream_t stream) {
  return pa_flash_impl(out, query, key_cache, value_cache, num_seqs, num_heads, head_size, num_kv_heads, scale,
                       block_tables, seq_lens, block_size, num_blocks, max_seq_len, max_num_blocks_per_seq, alibi_slopes,
                       q_dtype, kv_dtype, k_scale, v_scale, q_stride, block_stride, token_stride, head_stride,
                       partition_size, workspace, workspace_bytes, -1, 0.f, stream);
}

// The same decode with a sliding window (flash-attn's window_size = (window_left, 0), flash_attn.py:366-369,507-510): the
// query at position seq_len - 1 attends keys seq_len - 1 - window_left .. seq_len - 1.  Blocks left of the window are
// never loaded: partitions are counted from the block that holds the first visible key, and the split-K plan is made
// for window_left + 16 tokens.  Same workspace contract (sized from max_seq_len is always enough).  softcap > 0: logits
// soft-capping, softcap * tanh(logit / softcap) (flash_attn.py:371-374, applied before ALiBi and the mask).
extern "C" int vb200_paged_attention_flash_window(
    void* out, const void* query, const void* key_cache, const void* value_cache, int num_seqs, int num_heads,
    int head_size, int num_kv_heads, float scale, const int32_t* block_tables, const int32_t* seq_lens, int block_size,
    int num_blocks, int max_seq_len, int max_num_blocks_per_seq, const float* alibi_slopes, int q_dtype, int kv_dtype,
    const float* k_scale, const float* v_scale, int64_t q_stride, int64_t block_stride, int64_t token_stride,
    int64_t head_stride, int partition_size, void* workspace, size_t workspace_bytes, int window_left, float softcap,
    vb200_stream_t stream) {
  return pa_flash_impl(out, query, key_cache, value_cache, num_seqs, num_heads, head_size, num_kv_heads, scale,
                       block_tables, seq_lens, block_size, num_blocks, max_seq_len, max_num_blocks_per_seq, alibi_slopes,
                       q_dtype, kv_dtype, k_scale, v_scale, q_stride, block_stride, token_stride, head_stride,
                       partition_size, workspace, workspace_bytes, window_left, softcap, stream);
}

// paged_attention_v1 + dynamic per-token fp8 quantisation of its output in the split-K reduce kernel (SURVEY.md 8f
// row 1; see include/vllm_b200.h).
extern "C" size_t vb200_paged_attention_v1_quant_workspace_size(int num_seqs, int num_heads, int head_size, int max_seq_len,
                                                                int q_dtype, int partition_size) {
  (void)q_dtype;
  if (partition_size <= 0) return 0;
  size_t np = (size_t)(max_seq_len + partition_size - 1) / partition_size;
  if (np < 1) np = 1;
  const size_t rows = (size_t)num_seqs * num_heads * np;
  return rows * 4 * 2 + rows * head_size * 2 + 64;
}

extern "C" int vb200_paged_attention_v1_quant(
    void* out_q, float* out_scales, const float* scale_ub, const void* query, const void* key_cache,
    const void* value_cache, int num_seqs, int num_heads, int head_size, int num_kv_heads, float scale,
    const int32_t* block_tables, const int32_t* seq_lens, int block_size, int max_seq_len, int max_num_blocks_per_seq,
    const float* alibi_slopes, int q_dtype, int kv_dtype, const float* k_scale, const float* v_scale, int64_t q_stride,
    int64_t kv_block_stride, int64_t kv_head_stride, int partition_size, void* workspace, size_t workspace_bytes,
    vb200_stream_t stream) {
  VB_CHECK(out_q != nullptr && out_scales != nullptr, "paged_attention_v1_quant: out_q and out_scales are required");
  VB_CHECK(block_size == 16 && head_size % 32 == 0, "paged_attention_v1_quant: block size 16 and head size %% 32 == 0");
  VB_CHECK(workspace != nullptr, "paged_attention_v1_quant: a workspace is required");
  int part = partition_size;
  if (part <= 0)
    part = vb200_paged_attention_v1_partition_size(num_seqs, num_heads, num_kv_heads, head_size, max_seq_len, block_size,
                                                   kv_dtype);
  if (part <= 0) return 0;
  int np = (max_seq_len + part - 1) / part;
  if (np < 1) np = 1;
  const size_t need = vb200_paged_attention_v1_quant_workspace_size(num_seqs, num_heads, head_size, max_seq_len, q_dtype, part);
  VB_CHECK(workspace_bytes >= need, "paged_attention_v1_quant workspace too small: %zu < %zu", workspace_bytes, need);
  const size_t rows = (size_t)num_seqs * num_heads * np;
  PagedAttnParams p;
  p.out = nullptr;
  p.exp_sums = reinterpret_cast<float*>(workspace);
  p.max_logits = p.exp_sums + rows;
  p.tmp_out = reinterpret_cast<void*>(p.max_logits + rows);
  p.q = query;
  p.k_cache = key_cache;
  p.v_cache = value_cache;
  p.block_tables = block_tables;
  p.seq_lens = seq_lens;
  p.alibi_slopes = alibi_slopes;
  p.k_scale = k_scale;
  p.v_scale = v_scale;
  p.num_heads = num_heads;
  p.num_kv_heads = num_kv_heads;
  p.max_num_blocks_per_seq = max_num_blocks_per_seq;
  p.partition_size = part;
  p.max_num_partitions = np;
  p.force_partials = 1;
  p.q_stride = q_stride;
  p.kv_block_stride = kv_block_stride;
  p.kv_head_stride = kv_head_stride;
  p.scale = scale;
  p.prefetch_dist = 0;
  p.out_q = out_q;
  p.out_scales = out_scales;
  p.scale_ub = scale_ub;
  return run_paged_attention(p, num_seqs, head_size, block_size, max_seq_len, q_dtype, kv_dtype, (cudaStream_t)stream);
}

// ---- paged_attention_v1 over library-owned scratch: an entry point that allocates nothing ------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The split-K partials
// (exp_sums | max_logits | tmp_out) live in a file-scope device array instead of a caller workspace.  Every GPU that
// loads the library holds the array until exit; every call on a device shares it, so calls that overlap on the GPU
// (two streams, or a graph replay next to eager work) race on it.  Nothing in it has to start at zero: the main
// kernel writes every (sequence, head, partition) slot the reduce kernel reads, so the previous call's contents are
// never observed and no clearing launch is needed.
#define VB200_PA_STATIC_SCRATCH_BYTES (16u << 20)
static __device__ __align__(256) unsigned char g_pa_static_scratch[VB200_PA_STATIC_SCRATCH_BYTES];

// In host code `g_pa_static_scratch` is a host-side stand-in, not a device address: the address a kernel may use
// comes from cudaGetSymbolAddress, asked once per device with that device current.
static int pa_static_scratch_ptr(void** out) {
  constexpr int kMaxDevices = 64;
  static void* ptrs[kMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  VB_CUDA_CHECK(cudaGetDevice(&dev));
  VB_CHECK(dev >= 0 && dev < kMaxDevices, "paged_attention_v1_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    VB_CUDA_CHECK(cudaGetSymbolAddress(&p, g_pa_static_scratch));
    VB_CHECK(p != nullptr, "paged_attention_v1_static: cudaGetSymbolAddress returned a null address");
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return 0;
}

extern "C" size_t vb200_paged_attention_v1_static_max_bytes(void) { return VB200_PA_STATIC_SCRATCH_BYTES; }

extern "C" int vb200_paged_attention_v1_static(
    void* out, const void* query, const void* key_cache, const void* value_cache, int num_seqs, int num_heads,
    int head_size, int num_kv_heads, float scale, const int32_t* block_tables, const int32_t* seq_lens, int block_size,
    int max_seq_len, int max_num_blocks_per_seq, const float* alibi_slopes, int q_dtype, int kv_dtype,
    const float* k_scale, const float* v_scale, int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride,
    int partition_size, vb200_stream_t stream) {
  int part = partition_size;
  if (part <= 0)
    part = vb200_paged_attention_v1_partition_size(num_seqs, num_heads, num_kv_heads, head_size, max_seq_len, block_size,
                                                   kv_dtype);
  const size_t need = vb200_paged_attention_v1_workspace_size(num_seqs, num_heads, head_size, max_seq_len, q_dtype, part);
  VB_CHECK(need <= (size_t)VB200_PA_STATIC_SCRATCH_BYTES,
           "paged_attention_v1_static: the split-K partials of this call need %zu bytes, the library-owned scratch "
           "holds %zu (num_seqs %d x num_heads %d x partitions of %d tokens over max_seq_len %d); use "
           "vb200_paged_attention_v1 with a caller workspace",
           need, (size_t)VB200_PA_STATIC_SCRATCH_BYTES, num_seqs, num_heads, part, max_seq_len);
  void* scratch = nullptr;
  if (int rc = pa_static_scratch_ptr(&scratch)) return rc;
  // same partition plan, kernels and arithmetic as vb200_paged_attention_v1 given a caller workspace of `need` bytes
  return vb200_paged_attention_v1(out, query, key_cache, value_cache, num_seqs, num_heads, head_size, num_kv_heads, scale,
                                  block_tables, seq_lens, block_size, max_seq_len, max_num_blocks_per_seq, alibi_slopes,
                                  q_dtype, kv_dtype, k_scale, v_scale, q_stride, kv_block_stride, kv_head_stride, part,
                                  scratch, (size_t)VB200_PA_STATIC_SCRATCH_BYTES, stream);
}
