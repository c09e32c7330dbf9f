// Variable-length (prefill / chunked-prefill / mixed) causal attention over the paged FLASH-layout KV cache
// ("prefill-with-paged-KV"): the general case of the call the reference's V1 backend makes,
//   flash_attn_varlen_func(q, k=key_cache, v=value_cache, cu_seqlens_q, seqused_k, block_table, causal=True)
//   /root/reference/vllm/v1/attention/backends/flash_attn.py:497-519
// q holds every sequence's NEW tokens back to back (cu_seqlens_q), the cache already holds ALL of a sequence's keys
// and values including the new ones (reshape_and_cache_flash runs first, :448-463), and query token i of a
// sequence with q_len new tokens out of kv_len attends keys 0 .. kv_len - q_len + i (bottom-right aligned causal
// mask, flash-attn's convention).  The decode case (q_len == 1 everywhere) has its own HBM-roofline kernel
// (paged_attention_flash.cuh); this one is the compute-shaped case.
//
// One CTA = 128 (or 64, for short chunks) rows of one sequence x one KV head, a row being a (token, query head of
// that KV head's GQA group) pair; 4 warps x 32 (16) rows.
// FlashAttention-2 structure on the warp-level tensor-core path (mma.sync m16n8k16, fp32 accumulate): K / V tiles
// of 64 tokens are gathered page by page through the block table with cp.async (16-byte chunks, XOR-swizzled
// rows, two stages), every K / V fragment read from shared memory feeds both of a warp's row tiles, S = Q K^T and
// O += P V never leave registers, online softmax in the exp2 domain.  Tiles that lie entirely above the causal
// diagonal are never loaded; a sequence's tiles are issued deepest-context first.
#include <cmath>

#include "common.cuh"

namespace vb200 {

namespace prefill {

constexpr int BN = 64, NWARPS = 4, NTHREADS = NWARPS * 32;

struct Params {
  void* out;
  const void* q;
  const void* k_cache;
  const void* v_cache;
  const int* block_table;
  const int* cu_seqlens_q;
  const int* seqused_k;
  int num_seqs, num_heads, num_kv_heads, max_blocks_per_seq;
  int block_shift;                                  // log2(block_size)
  int group;                                        // query heads per kv head
  int64_t q_stride, o_stride;                       // elements between consecutive tokens
  int64_t block_stride, token_stride, head_stride;  // cache, elements
  float scale_log2;                                 // softmax scale * log2(e)
  int causal;
  const float* k_scale;                             // fp8 cache: per-tensor dequantisation scales (device)
  const float* v_scale;
  // EXTRA kernels only (sliding window / logits soft-cap / log-sum-exp output, flash_attn.py:342-350,507-510,711-760):
  int window_left;                                  // keys p - window_left .. p of a query at position p; < 0: no window
  float cap_pre;                                    // soft-cap: softmax scale / cap (then scale_log2 = cap * log2(e)); 0: none
  float* lse;                                       // optional fp32 [num_heads, lse_stride]: log-sum-exp of every row's logits
  int64_t lse_stride;
  const float* alibi_slopes;                        // optional fp32 [num_heads]: logit -= slope * |key - query position|
};

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int n = valid ? 16 : 0;   // src-size 0: the 16 bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}

template <typename T>
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void mma16816<__half>(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma16816<__nv_bfloat16>(float (&c)[4], const uint32_t (&a)[4], uint32_t b0,
                                                        uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <typename T>
__device__ __forceinline__ uint32_t pack2(float lo, float hi);
template <>
__device__ __forceinline__ uint32_t pack2<__half>(float lo, float hi) {
  const __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float lo, float hi) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// Soft-cap tanh: tanh.approx.f32 (one MUFU operation, 2^-11 relative error), which is what the library kernel the
// reference calls uses (cutlass::fast_tanh).  -DVB200_SOFTCAP_TANH_EXACT builds tanh(x) = 1 - 2 / (e^(2x) + 1) instead
// (two MUFU operations, ~1e-7 absolute): both pass the parity suite; the exact form costs 10 % on a capped prefill
// (profiles/r02_softcap_tanh_s80a.txt).
__device__ __forceinline__ float tanh_2mufu(float x) {
#ifdef VB200_SOFTCAP_TANH_EXACT
  return 1.f - __fdividef(2.f, ex2(x * 2.8853900817779268f) + 1.f);
#else
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
#endif
}

// two e4m3 bytes (low byte first) -> two T values packed in 32 bits; exact (e4m3 fits f16 and bf16)
template <typename T>
__device__ __forceinline__ uint32_t e4m3x2_to(uint16_t v);
template <>
__device__ __forceinline__ uint32_t e4m3x2_to<__half>(uint16_t v) { return e4m3x2_to_f16x2(v); }
template <>
__device__ __forceinline__ uint32_t e4m3x2_to<__nv_bfloat16>(uint16_t v) {
  const uint32_t h = e4m3x2_to_f16x2(v);
  const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&h));
  return pack2<__nv_bfloat16>(f.x, f.y);
}

// byte offset of 16-byte chunk `c` of row `r` in a tile whose rows hold D 16-bit elements
template <int D>
__device__ __forceinline__ uint32_t swz(int r, int c) {
  return (uint32_t)(r * (D * 2) + ((c ^ (r & 7)) << 4));
}

// FP8: the cache holds e4m3 bytes.  Tiles arrive as bytes in a two-stage staging ring and are expanded (exactly) to T
// into ONE 16-bit tile pair in front of the MMAs; the per-tensor scales are folded into the logits (k) and the output
// (v), as the decode kernel does.
// EXTRA: sliding window (tiles left of every row's window are never loaded; the left edge is masked like the diagonal)
// and logits soft-capping (cap * tanh(logit / cap) in front of the mask), and the optional log-sum-exp output; a
// separate instantiation so that the plain causal kernel keeps its registers and instruction stream.
template <typename T, int D, int MT, bool FP8, bool EXTRA>
__global__ void __launch_bounds__(NTHREADS) paged_prefill_flash_kernel(const Params p) {
  pdl_wait();
  pdl_trigger();
  constexpr int BM = 16 * MT * NWARPS;             // query rows per CTA: every warp owns MT 16-row tiles
  constexpr int CPR = D / 8;                       // 16-byte chunks per row of 16-bit elements
  constexpr int CPR_KV = FP8 ? D / 16 : D / 8;     // 16-byte chunks per cache row as it is loaded
  constexpr int KV_PASSES = BN * CPR_KV / NTHREADS;   // chunks of one K (or V) tile per thread
  constexpr int Q_PASSES = BM * CPR / NTHREADS;
  constexpr int TILE16 = BN * D * 2, TILE_LD = FP8 ? BN * D : TILE16;
  static_assert(BN * CPR_KV % NTHREADS == 0 && D % 16 == 0, "tile / head size");
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t sQ = (uint32_t)__cvta_generic_to_shared(smem);
  const uint32_t sK = sQ + BM * D * 2;             // 2 stages (as loaded)
  const uint32_t sV = sK + 2 * TILE_LD;            // 2 stages
  const uint32_t sK16 = sV + 2 * TILE_LD;          // FP8 only: the expanded tiles
  const uint32_t sV16 = sK16 + TILE16;
  const uint8_t* smem_k8 = smem + BM * D * 2;      // generic views of the staging ring (FP8)
  const uint8_t* smem_v8 = smem_k8 + 2 * TILE_LD;

  // The rows of a tile are (token, query head of this kv head's group) pairs, token-major: row r of a sequence is
  // token r / G, head kvh * G + r % G -- a K / V tile is loaded once for the whole GQA group, and a decode row of a
  // mixed batch brings G rows instead of one.
  // which (sequence, BM-row tile) is this CTA?  grid.x covers an upper bound of the tile count
  const int G = p.group;
  int t = blockIdx.x, seq = 0, q0 = 0, q_len = 0;
  for (; seq < p.num_seqs; ++seq) {
    q0 = p.cu_seqlens_q[seq];
    q_len = p.cu_seqlens_q[seq + 1] - q0;
    const int nt = (q_len * G + BM - 1) / BM;
    if (t < nt) {
      t = nt - 1 - t;                              // the tiles with the most keys first
      break;
    }
    t -= nt;
  }
  if (seq >= p.num_seqs) return;
  const int m0 = t * BM;
  const int n_rows = q_len * G;
  const int kv_len = p.seqused_k[seq];
  const int kvh = blockIdx.y;
  const int shift = kv_len - q_len;                // query token i sees keys 0 .. shift + i
  const int last_tok = (min(m0 + BM, n_rows) - 1) / G;
  // keys this tile can see at all
  int n_keys = p.causal ? min(kv_len, shift + last_tok + 1) : kv_len;
  if (n_keys < 0) n_keys = 0;
  const int n_tiles = (n_keys + BN - 1) / BN;
  const int wl = EXTRA && p.causal ? p.window_left : -1;
  int j0 = 0;                                      // first tile any row of this CTA can see
  if (EXTRA && wl >= 0) {
    const int lo = shift + m0 / G - wl;
    if (lo > 0) j0 = min(lo / BN, n_tiles);
  }

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, tq = lane & 3;
  const T* qg = reinterpret_cast<const T*>(p.q) + (int64_t)q0 * p.q_stride + (int64_t)kvh * G * D;
  const T* kc = reinterpret_cast<const T*>(p.k_cache) + (FP8 ? 0 : (int64_t)kvh * p.head_stride);   // 16-bit caches
  const T* vc = reinterpret_cast<const T*>(p.v_cache) + (FP8 ? 0 : (int64_t)kvh * p.head_stride);
  const bool capped = EXTRA && p.cap_pre != 0.f;
  const float scale_log2 = FP8 && !capped ? p.scale_log2 * *p.k_scale : p.scale_log2;
  const float cap_pre = !capped ? 0.f : FP8 ? p.cap_pre * *p.k_scale : p.cap_pre;
  const float out_scale = FP8 ? *p.v_scale : 1.f;
  const int* bt = p.block_table + (int64_t)seq * p.max_blocks_per_seq;
  const int bmask = (1 << p.block_shift) - 1;

  // ---- Q tile -> shared memory
#pragma unroll
  for (int i = 0; i < Q_PASSES; ++i) {
    const int idx = tid + i * NTHREADS;
    const int r = idx / CPR, c = idx % CPR;
    const bool ok = m0 + r < n_rows;
    const int tok = ok ? (m0 + r) / G : 0, hd = ok ? (m0 + r) % G : 0;
    cp_async16(sQ + swz<D>(r, c), qg + (int64_t)tok * p.q_stride + hd * D + c * 8, ok);
  }
  auto load_kv = [&](int j, int stage) {
    const uint32_t dk = sK + stage * TILE_LD, dv = sV + stage * TILE_LD;
#pragma unroll
    for (int i = 0; i < KV_PASSES; ++i) {
      const int idx = tid + i * NTHREADS;
      const int r = idx / CPR_KV, c = idx % CPR_KV;
      const int n = j * BN + r;
      const bool ok = n < n_keys;
      int64_t off = 0;
      if (ok) {
        const int page = bt[n >> p.block_shift];
        off = (int64_t)page * p.block_stride + (int64_t)(n & bmask) * p.token_stride + c * (FP8 ? 16 : 8);
      }
      if (FP8) {   // bytes, rows as they lie (the expansion pass reads them 16 bytes per thread: conflict-free)
        cp_async16(dk + r * D + c * 16, reinterpret_cast<const uint8_t*>(p.k_cache) + (int64_t)kvh * p.head_stride + off, ok);
        cp_async16(dv + r * D + c * 16, reinterpret_cast<const uint8_t*>(p.v_cache) + (int64_t)kvh * p.head_stride + off, ok);
      } else {
        cp_async16(dk + swz<D>(r, c), kc + off, ok);
        cp_async16(dv + swz<D>(r, c), vc + off, ok);
      }
    }
  };
  // FP8: bytes of stage `stage` -> the swizzled 16-bit tiles
  auto expand_kv = [&](int stage) {
#pragma unroll
    for (int i = 0; i < KV_PASSES; ++i) {
      const int idx = tid + i * NTHREADS;
      const int r = idx / CPR_KV, c = idx % CPR_KV;
#pragma unroll
      for (int which = 0; which < 2; ++which) {
        const uint8_t* src = (which ? smem_v8 : smem_k8) + stage * TILE_LD + r * D + c * 16;
        const uint4 b = *reinterpret_cast<const uint4*>(src);
        const uint32_t w[4] = {b.x, b.y, b.z, b.w};
        uint32_t o[8];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          o[2 * k] = e4m3x2_to<T>((uint16_t)(w[k] & 0xffffu));
          o[2 * k + 1] = e4m3x2_to<T>((uint16_t)(w[k] >> 16));
        }
        const uint32_t dst = which ? sV16 : sK16;
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + swz<D>(r, 2 * c)), "r"(o[0]), "r"(o[1]),
                     "r"(o[2]), "r"(o[3]) : "memory");
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + swz<D>(r, 2 * c + 1)), "r"(o[4]), "r"(o[5]),
                     "r"(o[6]), "r"(o[7]) : "memory");
      }
    }
  };
  if (n_tiles > j0) load_kv(j0, 0);
  cp_commit();

  float o[MT][D / 8][4];
  float row_max[MT][2], row_sum[MT][2];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
    for (int i = 0; i < D / 8; ++i) o[mt][i][0] = o[mt][i][1] = o[mt][i][2] = o[mt][i][3] = 0.f;
    row_max[mt][0] = row_max[mt][1] = -INFINITY;
    row_sum[mt][0] = row_sum[mt][1] = 0.f;
  }
  // Q fragments stay in registers when one 16-row tile per warp leaves room; otherwise they are re-read per k-step
  constexpr bool QREG = MT == 1 && D <= 128;
  uint32_t qf[QREG ? D / 16 : 1][4];
  const int wrow = warp * 16 * MT;                 // first row of this warp inside the CTA tile
  int tok_of[MT][2];                               // token (within the sequence) of this thread's rows
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int h = 0; h < 2; ++h) tok_of[mt][h] = (m0 + wrow + mt * 16 + g + h * 8) / G;
  const int warp_first_tok = (m0 + wrow) / G;
  const int warp_last_tok = (m0 + wrow + 16 * MT - 1) / G;
  // ALiBi in the units of the raw logits: (s + b / scale) * scale = s * scale + b
  float alibi_c[MT][2];
  const bool use_alibi = EXTRA && p.alibi_slopes != nullptr;
  if (use_alibi) {
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int h = 0; h < 2; ++h)
        alibi_c[mt][h] = p.alibi_slopes[kvh * G + (m0 + wrow + mt * 16 + g + h * 8) % G] * 1.4426950408889634f / scale_log2;
  }

  for (int j = j0; j < n_tiles; ++j) {
    const int stage = (j - j0) & 1;
    if (j + 1 < n_tiles) load_kv(j + 1, stage ^ 1);
    cp_commit();
    cp_wait<1>();
    __syncthreads();
    if (QREG && j == j0) {
#pragma unroll
      for (int ks = 0; ks < (QREG ? D / 16 : 1); ++ks)
        ldsm_x4(sQ + swz<D>(wrow + (lane & 15), ks * 2 + (lane >> 4)), qf[ks]);
    }
    if (FP8) {
      expand_kv(stage);
      __syncthreads();
    }
    const uint32_t tk = FP8 ? sK16 : sK + stage * TILE16, tv = FP8 ? sV16 : sV + stage * TILE16;
    // ---- S = Q K^T  (16 * MT rows x 64 keys per warp): a K fragment feeds MT row tiles
    float s[MT][BN / 8][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int i = 0; i < BN / 8; ++i) s[mt][i][0] = s[mt][i][1] = s[mt][i][2] = s[mt][i][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) {
      uint32_t qa[MT][4];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        if (QREG) {
#pragma unroll
          for (int e = 0; e < 4; ++e) qa[mt][e] = qf[QREG ? ks : 0][e];
        } else {
          ldsm_x4(sQ + swz<D>(wrow + mt * 16 + (lane & 15), ks * 2 + (lane >> 4)), qa[mt]);
        }
      }
#pragma unroll
      for (int np = 0; np < BN / 16; ++np) {
        uint32_t b[4];
        ldsm_x4(tk + swz<D>(np * 16 + (lane & 7) + ((lane >> 4) << 3), ks * 2 + ((lane >> 3) & 1)), b);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          mma16816<T>(s[mt][2 * np], qa[mt], b[0], b[1]);
          mma16816<T>(s[mt][2 * np + 1], qa[mt], b[2], b[3]);
        }
      }
    }
    if (EXTRA && capped) {
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int i = 0; i < BN / 8; ++i)
#pragma unroll
          for (int e = 0; e < 4; ++e) s[mt][i][e] = tanh_2mufu(s[mt][i][e] * cap_pre);
    }
    const int n_base = j * BN;
    if (use_alibi) {
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int i = 0; i < BN / 8; ++i)
#pragma unroll
          for (int e = 0; e < 4; ++e)
            s[mt][i][e] = fmaf(-alibi_c[mt][e >> 1],
                               fabsf((float)(n_base + i * 8 + 2 * tq + (e & 1) - shift - tok_of[mt][e >> 1])), s[mt][i][e]);
    }
    // ---- mask (only tiles that touch the diagonal, the left edge of the window or the end of the sequence)
    const bool need_mask = n_base + BN > n_keys || (p.causal && n_base + BN - 1 > shift + warp_first_tok) ||
                           (EXTRA && wl >= 0 && n_base < shift + warp_last_tok - wl);
    if (need_mask) {
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int i = 0; i < BN / 8; ++i)
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int n = n_base + i * 8 + 2 * tq + (e & 1);
            const bool vis = n < kv_len && (!p.causal || n <= shift + tok_of[mt][e >> 1]) &&
                             (!(EXTRA && wl >= 0) || n >= shift + tok_of[mt][e >> 1] - wl);
            if (!vis) s[mt][i][e] = -INFINITY;
          }
    }
    // ---- online softmax (exp2 domain)
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      float mx[2] = {row_max[mt][0], row_max[mt][1]};
#pragma unroll
      for (int i = 0; i < BN / 8; ++i) {
        mx[0] = fmaxf(mx[0], fmaxf(s[mt][i][0], s[mt][i][1]));
        mx[1] = fmaxf(mx[1], fmaxf(s[mt][i][2], s[mt][i][3]));
      }
      float alpha[2], mscaled[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
        mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 2));
        mscaled[h] = mx[h] == -INFINITY ? 0.f : mx[h] * scale_log2;     // a fully masked row so far: exp2(-inf) = 0
        alpha[h] = ex2(row_max[mt][h] * scale_log2 - mscaled[h]);     // row_max = -inf -> 0
        row_max[mt][h] = mx[h];
        row_sum[mt][h] *= alpha[h];
      }
#pragma unroll
      for (int i = 0; i < D / 8; ++i) {
        o[mt][i][0] *= alpha[0];
        o[mt][i][1] *= alpha[0];
        o[mt][i][2] *= alpha[1];
        o[mt][i][3] *= alpha[1];
      }
#pragma unroll
      for (int i = 0; i < BN / 8; ++i) {
        s[mt][i][0] = ex2(fmaf(s[mt][i][0], scale_log2, -mscaled[0]));
        s[mt][i][1] = ex2(fmaf(s[mt][i][1], scale_log2, -mscaled[0]));
        s[mt][i][2] = ex2(fmaf(s[mt][i][2], scale_log2, -mscaled[1]));
        s[mt][i][3] = ex2(fmaf(s[mt][i][3], scale_log2, -mscaled[1]));
        row_sum[mt][0] += s[mt][i][0] + s[mt][i][1];
        row_sum[mt][1] += s[mt][i][2] + s[mt][i][3];
      }
    }
    // ---- O += P V: a V fragment feeds MT row tiles
#pragma unroll
    for (int kk = 0; kk < BN / 16; ++kk) {
      uint32_t a[MT][4];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        a[mt][0] = pack2<T>(s[mt][2 * kk][0], s[mt][2 * kk][1]);
        a[mt][1] = pack2<T>(s[mt][2 * kk][2], s[mt][2 * kk][3]);
        a[mt][2] = pack2<T>(s[mt][2 * kk + 1][0], s[mt][2 * kk + 1][1]);
        a[mt][3] = pack2<T>(s[mt][2 * kk + 1][2], s[mt][2 * kk + 1][3]);
      }
#pragma unroll
      for (int dp = 0; dp < D / 16; ++dp) {
        uint32_t b[4];
        ldsm_x4_trans(tv + swz<D>(kk * 16 + (lane & 7) + (((lane >> 3) & 1) << 3), dp * 2 + (lane >> 4)), b);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          mma16816<T>(o[mt][2 * dp], a[mt], b[0], b[1]);
          mma16816<T>(o[mt][2 * dp + 1], a[mt], b[2], b[3]);
        }
      }
    }
    // 16-bit caches: every warp is done with this stage before the next iteration's loads overwrite it.  FP8: the
    // expanded tiles are next written after the following iteration's first barrier, which every warp reaches
    // only after these MMAs.
    if (!FP8) __syncthreads();
  }
  cp_wait<0>();

  // ---- epilogue: O / l  ->  out
  T* og = reinterpret_cast<T*>(p.out) + (int64_t)q0 * p.o_stride + (int64_t)kvh * G * D;
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float l = row_sum[mt][h];
      l += __shfl_xor_sync(0xffffffffu, l, 1);
      l += __shfl_xor_sync(0xffffffffu, l, 2);
      const int row = m0 + wrow + mt * 16 + g + h * 8;
      if (row < n_rows) {
        const float inv = l > 0.f ? out_scale / l : 0.f;
        T* orow = og + (int64_t)tok_of[mt][h] * p.o_stride + (row - tok_of[mt][h] * G) * D;
        if (EXTRA && p.lse != nullptr && tq == 0)   // natural log; a row without a visible key: -inf (merge_attn_states' empty state)
          p.lse[(int64_t)(kvh * G + row - tok_of[mt][h] * G) * p.lse_stride + q0 + tok_of[mt][h]] =
              l > 0.f ? (row_max[mt][h] * scale_log2 + log2f(l)) * 0.6931471805599453f : -INFINITY;
#pragma unroll
        for (int i = 0; i < D / 8; ++i)
          *reinterpret_cast<uint32_t*>(orow + i * 8 + 2 * tq) =
              pack2<T>(o[mt][i][2 * h] * inv, o[mt][i][2 * h + 1] * inv);
      }
    }
  }
}

template <typename T, int D, int MT, bool FP8, bool EXTRA>
static int launch(const Params& p, int total_q, cudaStream_t stream) {
  auto kern = paged_prefill_flash_kernel<T, D, MT, FP8, EXTRA>;
  constexpr int BM = 16 * MT * NWARPS;
  const size_t smem = (size_t)BM * D * 2 + (FP8 ? (size_t)4 * BN * D + (size_t)2 * BN * D * 2 : (size_t)4 * BN * D * 2);
  if (smem > 48 * 1024)
    VB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // every sequence adds at most one partial tile: an upper bound that needs no host read of cu_seqlens_q
  const int max_tiles = (int)(((int64_t)total_q * p.group) / BM) + p.num_seqs;
  VB_CUDA_CHECK(launch_k(kern, dim3(max_tiles, p.num_kv_heads), dim3(NTHREADS), smem, stream, p));
  VB_LAUNCH_CHECK();
  return 0;
}

// 128-row tiles (two 16-row tiles per warp: every K / V fragment read from shared memory feeds two MMAs' worth of
// rows) once some sequence has more than 64 rows; 64-row tiles for short chunks and for head size 256, whose
// accumulators leave no room for the second tile (192 likewise).
template <typename T, bool FP8, bool EXTRA>
static int dispatch(const Params& p, int head_size, int total_q, int max_seqlen_q, cudaStream_t stream) {
  const bool big = (int64_t)max_seqlen_q * p.group > 64;
  switch (head_size) {
    case 64:
      return big ? launch<T, 64, 2, FP8, EXTRA>(p, total_q, stream) : launch<T, 64, 1, FP8, EXTRA>(p, total_q, stream);
    case 128:
      return big ? launch<T, 128, 2, FP8, EXTRA>(p, total_q, stream) : launch<T, 128, 1, FP8, EXTRA>(p, total_q, stream);
    case 192: return launch<T, 192, 1, FP8, EXTRA>(p, total_q, stream);
    case 256: return launch<T, 256, 1, FP8, EXTRA>(p, total_q, stream);
    default: break;
  }
  VB_CHECK(false, "flash_attn_varlen_paged: head size %d is not supported (64, 128, 192, 256)", head_size);
  return 0;
}

}  // namespace prefill
}  // namespace vb200

using namespace vb200;

extern "C" int vb200_flash_attn_varlen_paged_opts(void* out, const void* query, const void* key_cache,
                                                  const void* value_cache, const int32_t* cu_seqlens_q,
                                                  const int32_t* seqused_k, const int32_t* block_table, int num_seqs,
                                                  int total_q, int num_heads, int num_kv_heads, int head_size,
                                                  int block_size, int max_num_blocks_per_seq, int max_seqlen_q,
                                                  float scale, int causal, int dtype, int kv_dtype, const float* k_scale,
                                                  const float* v_scale, int64_t q_stride, int64_t out_stride,
                                                  int64_t block_stride, int64_t token_stride, int64_t head_stride,
                                                  int window_left, float softcap, float* softmax_lse,
                                                  int64_t lse_stride, const float* alibi_slopes,
                                                  vb200_stream_t stream) {
  VB_CHECK(window_left < 0 || causal, "flash_attn_varlen_paged: a sliding window (left, 0) needs causal attention");
  VB_CHECK(softcap >= 0.f && softcap == softcap, "flash_attn_varlen_paged: softcap must be >= 0");
  VB_CHECK(dtype == VB200_F16 || dtype == VB200_BF16, "flash_attn_varlen_paged: dtype must be f16 or bf16");
  VB_CHECK(num_kv_heads > 0 && num_heads % num_kv_heads == 0, "flash_attn_varlen_paged: num_heads %% num_kv_heads != 0");
  VB_CHECK(block_size > 0 && (block_size & (block_size - 1)) == 0,
           "flash_attn_varlen_paged: block_size %d must be a power of two", block_size);
  VB_CHECK(kv_dtype == VB200_KV_AUTO || kv_dtype == VB200_KV_FP8_E4M3, "Unsupported data type of kv cache: %d", kv_dtype);
  const bool fp8 = kv_dtype == VB200_KV_FP8_E4M3;
  VB_CHECK(!fp8 || (k_scale && v_scale), "fp8 kv cache needs k_scale and v_scale");
  VB_CHECK(((q_stride | out_stride) & 7) == 0 && ((block_stride | token_stride | head_stride) & (fp8 ? 15 : 7)) == 0,
           "flash_attn_varlen_paged: strides must be multiples of 16 bytes");
  VB_CHECK(((reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(query) |
             reinterpret_cast<uintptr_t>(key_cache) | reinterpret_cast<uintptr_t>(value_cache)) & 15) == 0,
           "flash_attn_varlen_paged: tensors must be 16-byte aligned");
  if (num_seqs <= 0 || total_q <= 0) return 0;
  prefill::Params p;
  p.out = out;
  p.q = query;
  p.k_cache = key_cache;
  p.v_cache = value_cache;
  p.block_table = block_table;
  p.cu_seqlens_q = cu_seqlens_q;
  p.seqused_k = seqused_k;
  p.num_seqs = num_seqs;
  p.num_heads = num_heads;
  p.num_kv_heads = num_kv_heads;
  p.max_blocks_per_seq = max_num_blocks_per_seq;
  p.block_shift = 0;
  while ((1 << p.block_shift) < block_size) ++p.block_shift;
  p.group = num_heads / num_kv_heads;
  p.q_stride = q_stride;
  p.o_stride = out_stride;
  p.block_stride = block_stride;
  p.token_stride = token_stride;
  p.head_stride = head_stride;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.causal = causal;
  p.k_scale = k_scale;
  p.v_scale = v_scale;
  VB_CHECK(softmax_lse == nullptr || lse_stride >= total_q, "flash_attn_varlen_paged: lse_stride < total_q");
  p.lse = softmax_lse;
  p.lse_stride = lse_stride;
  p.alibi_slopes = alibi_slopes;
  p.window_left = window_left < 0 ? -1 : window_left;
  p.cap_pre = 0.f;
  if (softcap > 0.f) {   // logits = softcap * tanh(q.k * scale / softcap): the tanh takes q.k * cap_pre, the exponent scale is the cap
    p.cap_pre = scale / softcap;
    p.scale_log2 = softcap * 1.4426950408889634f;
  }
  if (max_seqlen_q <= 0) max_seqlen_q = total_q;   // unknown: assume long chunks
  cudaStream_t st = (cudaStream_t)stream;
  if (p.window_left >= 0 || p.cap_pre != 0.f || p.lse != nullptr || p.alibi_slopes != nullptr) {
    if (dtype == VB200_F16)
      return fp8 ? prefill::dispatch<__half, true, true>(p, head_size, total_q, max_seqlen_q, st)
                 : prefill::dispatch<__half, false, true>(p, head_size, total_q, max_seqlen_q, st);
    return fp8 ? prefill::dispatch<__nv_bfloat16, true, true>(p, head_size, total_q, max_seqlen_q, st)
               : prefill::dispatch<__nv_bfloat16, false, true>(p, head_size, total_q, max_seqlen_q, st);
  }
  if (dtype == VB200_F16)
    return fp8 ? prefill::dispatch<__half, true, false>(p, head_size, total_q, max_seqlen_q, st)
               : prefill::dispatch<__half, false, false>(p, head_size, total_q, max_seqlen_q, st);
  return fp8 ? prefill::dispatch<__nv_bfloat16, true, false>(p, head_size, total_q, max_seqlen_q, st)
             : prefill::dispatch<__nv_bfloat16, false, false>(p, head_size, total_q, max_seqlen_q, st);
}

extern "C" int vb200_flash_attn_varlen_paged(void* out, const void* query, const void* key_cache, const void* value_cache,
                                             const int32_t* cu_seqlens_q, const int32_t* seqused_k,
                                             const int32_t* block_table, int num_seqs, int total_q, int num_heads,
                                             int num_kv_heads, int head_size, int block_size,
                                             int max_num_blocks_per_seq, int max_seqlen_q, float scale, int causal,
                                             int dtype, int kv_dtype, const float* k_scale, const float* v_scale,
                                             int64_t q_stride, int64_t out_stride, int64_t block_stride,
                                             int64_t token_stride, int64_t head_stride, vb200_stream_t stream) {
  return vb200_flash_attn_varlen_paged_opts(out, query, key_cache, value_cache, cu_seqlens_q, seqused_k, block_table,
                                            num_seqs, total_q, num_heads, num_kv_heads, head_size, block_size,
                                            max_num_blocks_per_seq, max_seqlen_q, scale, causal, dtype, kv_dtype, k_scale,
                                            v_scale, q_stride, out_stride, block_stride, token_stride, head_stride, -1,
                                            0.f, nullptr, 0, nullptr, stream);
}
