// First-generation fp8 GEMM (one CTA per 128 weight rows, cluster split-K through DSMEM).  Kept only as an A/B
// reference while the CTA-pair kernel in scaled_mm.cu is validated (VB200_GEMM_V1=1 selects it).
#pragma once
#include "tc_common.cuh"

namespace vb200 {
namespace v1 {
using namespace tc;

constexpr int kBlockN = 128;   // weight rows per CTA (UMMA M)
constexpr int kBlockK = 128;   // fp8 elements (= bytes) of K per stage: one 128-byte swizzle row
constexpr int kUmmaK = 32;     // K per tcgen05.mma for 8-bit inputs
constexpr int kMaxTokTile = 256;
constexpr int kMaxStages = 8;
constexpr int kGemmThreads = 192;
constexpr int kWTileBytes = kBlockN * kBlockK;  // 16 KiB
constexpr int kMaxSplits = 8;                   // portable cluster size

struct GemmParams {
  void* out;              // [M, N] out_t, row stride ldc
  const float* a_scales;  // 1 or M
  const float* b_scales;  // 1 or N
  const void* bias;       // nullptr or [N] out_t
  int M, N, K;
  int64_t ldc;
  int a_per_token, b_per_channel;
  int tok_tile;      // tokens per CTA tile (multiple of 16, <= 256)
  int stages;        // pipeline depth
  int splits;        // K splits == cluster size along y
  int kb_per_split;  // k-blocks per split
  int num_kb;        // total k-blocks
  int col_chunk;     // split-K: token columns owned by one CTA of the cluster (multiple of 16)
  int tmem_cols;     // TMEM columns to allocate (power of two >= tok_tile, >= 32)
};

template <typename OutT>
__global__ void __launch_bounds__(kGemmThreads, 1)
    scaled_mm_kernel(const __grid_constant__ CUtensorMap tmap_w,
                     const __grid_constant__ CUtensorMap tmap_a, const GemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [stages][W tile 16 KiB | A tile tok_tile*128]  (re-used as the split-K receive buffer)
  //        | barriers | tmem ptr | a_scale[256]
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const int a_tile_bytes = p.tok_tile * kBlockK;
  const int stage_bytes = kWTileBytes + a_tile_bytes;
  size_t main_bytes = (size_t)p.stages * stage_bytes;
  const size_t recv_bytes = p.splits > 1 ? (size_t)p.splits * p.col_chunk * kBlockN * 4 : 0;
  if (recv_bytes > main_bytes) main_bytes = (recv_bytes + 1023) & ~(size_t)1023;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + main_bytes);
  uint64_t* empty_bar = full_bar + kMaxStages;
  uint64_t* acc_bar = empty_bar + kMaxStages;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(acc_bar + 1);
  float* s_ascale = reinterpret_cast<float*>(tmem_ptr + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_tile = blockIdx.x;
  const int split = blockIdx.y;  // == rank in the (1, splits, 1) cluster
  const int m_tile = blockIdx.z;
  const int n0 = n_tile * kBlockN;
  const int tok0 = m_tile * p.tok_tile;
  const int kb0 = split * p.kb_per_split;
  const int kb1 = min(kb0 + p.kb_per_split, p.num_kb);
  const int nkb = kb1 - kb0;  // >= 1 by construction

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmap_w);
    prefetch_tensormap(&tmap_a);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(acc_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_ptr, (uint32_t)p.tmem_cols);
  if (warp >= 2) {
    // per-token scales of this tile -> smem (scalar scale broadcast)
    for (int i = threadIdx.x - 64; i < p.tok_tile; i += 128) {
      const int t = tok0 + i;
      s_ascale[i] = p.a_per_token ? (t < p.M ? p.a_scales[t] : 0.f) : p.a_scales[0];
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sw = smem + (size_t)stage * stage_bytes;
        uint8_t* sa = sw + kWTileBytes;
        mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)stage_bytes);
        tma_load_2d(sw, &tmap_w, &full_bar[stage], kb * kBlockK, n0, kEvictFirst);
        tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * kBlockK, tok0, kEvictLast);
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      const uint32_t idesc = make_idesc_f8(kBlockN, p.tok_tile);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t sw = smem_u32(smem + (size_t)stage * stage_bytes);
        const uint64_t adesc = make_smem_desc_sw128(sw);
        const uint64_t bdesc = make_smem_desc_sw128(sw + kWTileBytes);
#pragma unroll
        for (int k = 0; k < kBlockK / kUmmaK; ++k) {
          // advance both descriptors by 32 bytes of K inside the swizzle row (address field >> 4)
          umma_f8(tmem_base, adesc + (uint64_t)(k * (kUmmaK >> 4)),
                  bdesc + (uint64_t)(k * (kUmmaK >> 4)), idesc, (kb | k) != 0);
        }
        umma_commit(&empty_bar[stage]);  // frees the stage when these MMAs have read it
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
      umma_commit(acc_bar);  // accumulator complete
    }
    __syncwarp();
  }

  // epilogue warps 2..5: TMEM lane quarter = warp % 4
  const int q = warp & 3;
  const int nl = q * 32 + lane;  // weight row inside the tile
  const int n = n0 + nl;         // output column of this thread
  const bool n_ok = n < p.N;
  const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
  OutT* outp = reinterpret_cast<OutT*>(p.out);

  if (p.splits == 1) {
    if (warp >= 2) {
      const float bsc = p.b_per_channel ? (n_ok ? p.b_scales[n] : 0.f) : p.b_scales[0];
      float biasf = 0.f;
      if (p.bias != nullptr && n_ok) biasf = to_float<OutT>(reinterpret_cast<const OutT*>(p.bias)[n]);
      mbar_wait(acc_bar, 0);
      tc_fence_after();
      for (int c = 0; c < p.tok_tile; c += 16) {
        uint32_t r[16];
        tmem_ld_x16(taddr + c, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int t = tok0 + c + j;
          if (n_ok && t < p.M) {
            // a_scale * (b_scale * acc) [+ bias]  (ScaledEpilogue / ScaledEpilogueBias)
            const float acc = __uint_as_float(r[j]);
            const float v = p.bias != nullptr ? fmaf(s_ascale[c + j], bsc * acc, biasf)
                                              : s_ascale[c + j] * (bsc * acc);
            outp[(int64_t)t * p.ldc + n] = from_float<OutT>(v);
          }
        }
      }
    }
  } else {
    // ===== cluster split-K: exchange partial tiles through distributed shared memory =====
    const int cw = p.col_chunk;
    if (warp >= 2) {
      mbar_wait(acc_bar, 0);  // this CTA's MMAs are done: its stages are free, its TMEM is final
      tc_fence_after();
    }
    cluster_sync_all();  // every CTA of the cluster is past its main loop
    if (warp >= 2) {
      const uint32_t recv_local = smem_u32(smem);
      for (int owner = 0; owner < p.splits; ++owner) {
        const int c_begin = owner * cw;
        const int c_end = min(c_begin + cw, p.tok_tile);
        if (c_begin >= c_end) break;
        const uint32_t remote = map_to_cta(recv_local, (uint32_t)owner) +
                                (uint32_t)(((split * cw) * kBlockN + nl) * 4);
        for (int c = c_begin; c < c_end; c += 16) {
          uint32_t r[16];
          tmem_ld_x16(taddr + c, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j)
            st_cluster_f32(remote + (uint32_t)((c - c_begin + j) * kBlockN * 4), __uint_as_float(r[j]));
        }
      }
    }
    cluster_sync_all();  // all partials have landed (release/acquire at cluster scope)
    if (warp >= 2) {
      const int c_begin = split * cw;
      const int c_end = min(c_begin + cw, p.tok_tile);
      if (c_begin < c_end) {
        const float bsc = p.b_per_channel ? (n_ok ? p.b_scales[n] : 0.f) : p.b_scales[0];
        float biasf = 0.f;
        if (p.bias != nullptr && n_ok) biasf = to_float<OutT>(reinterpret_cast<const OutT*>(p.bias)[n]);
        const float* recv = reinterpret_cast<const float*>(smem);
        for (int c = c_begin; c < c_end; ++c) {
          float acc = 0.f;
          for (int r = 0; r < p.splits; ++r) acc += recv[((r * cw) + (c - c_begin)) * kBlockN + nl];
          const int t = tok0 + c;
          if (n_ok && t < p.M) {
            const float v = p.bias != nullptr ? fmaf(s_ascale[c], bsc * acc, biasf)
                                              : s_ascale[c] * (bsc * acc);
            outp[(int64_t)t * p.ldc + n] = from_float<OutT>(v);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
  }
}

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

struct GemmPlan {
  int tok_tile, m_tiles, n_tiles, stages, splits, kb_per_split, num_kb, col_chunk, tmem_cols;
  size_t smem;
};

static GemmPlan plan_gemm(int M, int N, int K) {
  GemmPlan g;
  g.m_tiles = ceil_div(M, kMaxTokTile);
  g.tok_tile = (ceil_div(M, g.m_tiles) + 15) / 16 * 16;
  g.n_tiles = ceil_div(N, kBlockN);
  g.num_kb = ceil_div(K, kBlockK);
  const int stage_bytes = kWTileBytes + g.tok_tile * kBlockK;
  const int budget = 227 * 1024 - 1024 /*align*/ - 2048 /*barriers + scales*/;
  g.stages = budget / stage_bytes;
  if (g.stages > kMaxStages) g.stages = kMaxStages;
  // cluster split-K: the largest R <= 8 whose grid still fits one wave, keeping >= 4 k-blocks per split
  const int tiles = g.n_tiles * g.m_tiles;
  const int sms = num_sms();
  int splits = 1;
  for (int r = 2; r <= kMaxSplits; ++r)
    if (tiles * r <= sms && g.num_kb / r >= 4) splits = r;
  g.kb_per_split = ceil_div(g.num_kb, splits);
  g.splits = ceil_div(g.num_kb, g.kb_per_split);
  if (g.stages > g.kb_per_split) g.stages = g.kb_per_split < 2 ? 2 : g.kb_per_split;
  g.col_chunk = (ceil_div(g.tok_tile, g.splits) + 15) / 16 * 16;
  size_t main_bytes = (size_t)g.stages * stage_bytes;
  const size_t recv = g.splits > 1 ? (size_t)g.splits * g.col_chunk * kBlockN * 4 : 0;
  if (recv > main_bytes) main_bytes = (recv + 1023) & ~(size_t)1023;
  g.smem = 1024 + main_bytes + 2048;
  g.tmem_cols = 32;
  while (g.tmem_cols < g.tok_tile) g.tmem_cols <<= 1;
  return g;
}

template <typename OutT>
static int launch_gemm(const GemmPlan& g, const CUtensorMap& tw, const CUtensorMap& ta,
                       const GemmParams& p, cudaStream_t stream) {
  auto kern = scaled_mm_kernel<OutT>;
  VB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(g.n_tiles, g.splits, g.m_tiles);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = g.smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1;
  attr[0].val.clusterDim.y = g.splits;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  VB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tw, ta, p));
  return 0;
}

}  // namespace v1
}  // namespace vb200
