// FP8 w8a8 linear (cutlass_scaled_mm semantics,
// /root/reference/csrc/quantization/cutlass_w8a8/scaled_mm_entry.cu:169-240, epilogue
// /root/reference/csrc/cutlass_extensions/epilogue/scaled_mm_epilogues_c3x.hpp:137-216) as a
// hand-written tcgen05 / TMEM / TMA kernel for sm_100a.
//
// Design (weights on the UMMA M side, "swap-AB", on CTA PAIRS):
//   * Two CTAs of one TPC form a cluster and issue ONE tcgen05.mma.cta_group::2 of M = 256 weight
//     rows x N = token tile (16..256) x K = 32: each CTA stages its own 128 weight rows and HALF of
//     the token tile in shared memory, the tensor core reads both halves.  Per 128 bytes of K a CTA
//     therefore pulls 16 KiB of W and <= 16 KiB of A from L2 for 128 x 256 x 128 MACs: at decode
//     batch sizes the L2 -> SM fabric, not HBM, is what a one-CTA 128-row tile saturates (its A tile
//     is re-read by every weight tile), and the pair halves that traffic.
//   * Persistent: grid = min(units, #SM/2) pairs; a unit is (256-row weight tile, token tile, K
//     slice).  The fp32 accumulator (128 lanes x <= 256 columns per CTA) is double-buffered in TMEM
//     (512 columns), so the epilogue of unit i overlaps the main loop of unit i+1.
//   * Warp roles per CTA: warp 0 = TMA producer (128-byte-swizzled [rows x 128 B] boxes into a ring
//     of stages; both CTAs signal the LEADER's full barrier), warp 1 = MMA issuer (leader CTA only,
//     one thread; tcgen05.commit multicasts "stage free" / "accumulator ready" to both CTAs),
//     warps 2-5 = epilogue (tcgen05.ld -> a_scale * (b_scale * acc) (+ bias) -> bf16/f16 -> global).
//   * Split-K when the tile count cannot fill the machine (decode o_proj / down_proj / qkv): the S
//     pairs that share an output tile each accumulate a K slice; partial columns a CTA does not own
//     go to an L2-resident fp32 workspace, a per-tile arrival counter (release/acquire at gpu scope)
//     orders them, and every CTA reduces + writes 1/S of the token columns in split order
//     (deterministic).  Only used when all units are co-resident (units <= #SM/2), so the wait
//     cannot deadlock.
#include "tc_common.cuh"
#include "scaled_mm_1cta.cuh"

#include <cstdlib>

namespace vb200 {
namespace v2 {
using namespace tc;

constexpr int kPairRows = 256;  // weight rows per CTA pair (UMMA M with cta_group::2)
constexpr int kCtaRows = 128;   // weight rows staged (and accumulated) by one CTA
constexpr int kBlockK = 128;    // fp8 elements (= bytes) of K per stage: one 128-byte swizzle row
constexpr int kUmmaK = 32;      // K per tcgen05.mma for 8-bit inputs
constexpr int kMaxTokTile = 256;
constexpr int kMaxStages = 12;
constexpr int kThreads = 192;
constexpr int kWTileBytes = kCtaRows * kBlockK;  // 16 KiB
constexpr int kTmemCols = 512;                   // two accumulator stages of <= 256 columns
constexpr int kAccStride = 256;                  // TMEM column offset of accumulator stage 1
constexpr int kMaxSplits = 8;
constexpr int kCounterBytes = 16 * 1024;         // head of the workspace: arrival / departure counters
constexpr int kEpiBarrier = 1;                   // named barrier id of the 4 epilogue warps

struct Params {
  void* out;              // [M, N] out_t, row stride ldc
  const float* a_scales;  // 1 or M
  const float* b_scales;  // 1 or N
  const void* bias;       // nullptr or [N] out_t
  uint32_t* counters;     // split-K: arrive[2 * tiles] then depart[2 * tiles], zero between launches
  float* partials;        // split-K: [tiles * 2][owner][source slice][col_chunk / 4][128][4] fp32
  int M, N, K;
  int64_t ldc;
  int a_per_token, b_per_channel;
  int tok_tile;      // tokens per tile (multiple of 16, <= 256); each CTA stages tok_tile / 2 rows
  int m_tiles;       // token tiles
  int tiles;         // m_tiles * ceil(N / 256)
  int stages;        // shared-memory pipeline depth
  int stage_bytes;   // per CTA: 16 KiB of W + the half token tile rounded up to 1 KiB
  int splits;        // K slices per tile (1 = none)
  int kb_per_split;  // k-blocks per slice
  int num_kb;        // total k-blocks
  int col_chunk;     // split-K: token columns reduced + written by one slice (multiple of 16)
  int units;         // tiles * splits
  uint32_t main_bytes;  // bytes of the stage ring (barriers follow)
  uint64_t hint_w, hint_a;
};

template <typename OutT>
__device__ __forceinline__ void store_scaled(OutT* outp, int64_t ldc, int t, int n, float acc,
                                             float asc, float bsc, float biasf, bool has_bias) {
  // a_scale * (b_scale * acc) [+ bias]  (ScaledEpilogue / ScaledEpilogueBias)
  const float v = has_bias ? fmaf(asc, bsc * acc, biasf) : asc * (bsc * acc);
  outp[(int64_t)t * ldc + n] = from_float<OutT>(v);
}

template <typename OutT>
__global__ void __launch_bounds__(kThreads, 1)
    scaled_mm_kernel(const __grid_constant__ CUtensorMap tmap_w,
                     const __grid_constant__ CUtensorMap tmap_a, const Params p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [stages][W 16 KiB | A half tile] | barriers | tmem ptr | a_scale[2][256]
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const size_t main_bytes = p.main_bytes;  // >= stages * stage_bytes (and >= the split-K reduce buffer)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + main_bytes);
  uint64_t* empty_bar = full_bar + kMaxStages;
  uint64_t* acc_full = empty_bar + kMaxStages;  // [2] accumulator stage ready (both CTAs)
  uint64_t* acc_empty = acc_full + 2;           // [2] accumulator stage drained (leader's is used)
  uint64_t* red_bar = acc_empty + 2;             // split-K: bulk fetch of the foreign partials
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(red_bar + 1);
  float* s_ascale = reinterpret_cast<float*>(tmem_ptr + 4);  // [2][256]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();  // 0 = leader (issues the MMAs)
  const int pair = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;
  const int tok_half = p.tok_tile >> 1;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmap_w);
    prefetch_tensormap(&tmap_a);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);   // leader's producer arrive (+ transaction bytes of both CTAs)
      mbar_init(&empty_bar[s], 1);  // one multicast tcgen05.commit
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&acc_full[s], 1);   // one multicast tcgen05.commit
      mbar_init(&acc_empty[s], 8);  // 4 epilogue warps x 2 CTAs
    }
    mbar_init(red_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_2cta(tmem_ptr, kTmemCols);
  tc_fence_before();
  cluster_sync_all();  // barrier inits + TMEM allocation visible in both CTAs
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===== TMA producer (both CTAs; one thread) =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int u = pair; u < p.units; u += num_pairs) {
        const int tile = u / p.splits, split = u - tile * p.splits;
        const int n_pt = tile / p.m_tiles, m_t = tile - n_pt * p.m_tiles;
        const int n0 = n_pt * kPairRows + (int)rank * kCtaRows;
        const int tok0 = m_t * p.tok_tile + (int)rank * tok_half;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.num_kb);
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int stage = it % p.stages;
          const uint32_t phase = (it / p.stages) & 1;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sw = smem + (size_t)stage * p.stage_bytes;
          uint8_t* sa = sw + kWTileBytes;
          if (rank == 0)
            mbar_arrive_expect_tx(&full_bar[stage],
                                  2u * (uint32_t)(kWTileBytes + tok_half * kBlockK));
          const uint32_t bar = map_to_cta(smem_u32(&full_bar[stage]), 0);
          tma_load_2d_2cta(sw, &tmap_w, bar, kb * kBlockK, n0, p.hint_w);
          tma_load_2d_2cta(sa, &tmap_a, bar, kb * kBlockK, tok0, p.hint_a);
        }
      }
      // tail: every stage this CTA filled has been released, i.e. all multicast arrivals aimed at
      // this CTA's barriers have landed before it can exit
      const uint32_t n_tail = it < (uint32_t)p.stages ? it : (uint32_t)p.stages;
      for (uint32_t i = 0; i < n_tail; ++i, ++it) {
        const int stage = it % p.stages;
        const uint32_t phase = (it / p.stages) & 1;
        mbar_wait(&empty_bar[stage], phase ^ 1);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA, one thread) =====
    if (rank == 0 && lane == 0) {
      const uint32_t idesc = make_idesc_f8(kPairRows, p.tok_tile);
      uint32_t it = 0, lt = 0;
      for (int u = pair; u < p.units; u += num_pairs, ++lt) {
        const int tile = u / p.splits, split = u - tile * p.splits;
        const int kb0 = split * p.kb_per_split;
        const int nkb = min(kb0 + p.kb_per_split, p.num_kb) - kb0;
        const uint32_t as = lt & 1, aphase = (lt >> 1) & 1;
        mbar_wait(&acc_empty[as], aphase ^ 1);  // both CTAs' epilogues drained this stage
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * kAccStride;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int stage = it % p.stages;
          const uint32_t phase = (it / p.stages) & 1;
          mbar_wait(&full_bar[stage], phase);  // both CTAs' boxes have landed
          tc_fence_after();
          const uint32_t sw = smem_u32(smem + (size_t)stage * p.stage_bytes);
          const uint64_t adesc = make_smem_desc_sw128(sw);
          const uint64_t bdesc = make_smem_desc_sw128(sw + kWTileBytes);
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            // advance both descriptors by 32 bytes of K inside the swizzle row (address field >> 4)
            umma_f8_2cta(d_tmem, adesc + (uint64_t)(k * (kUmmaK >> 4)),
                         bdesc + (uint64_t)(k * (kUmmaK >> 4)), idesc, (kb | k) != 0);
          }
          umma_commit_2cta(&empty_bar[stage], 0x3);  // frees the stage in both CTAs
        }
        umma_commit_2cta(&acc_full[as], 0x3);  // accumulator complete, both CTAs' epilogues may read
      }
    }
    __syncwarp();
  }

  if (warp >= 2) {
    // ===== epilogue (4 warps of each CTA; TMEM lane quarter = warp % 4) =====
    const int q = warp & 3;
    const int row = q * 32 + lane;  // weight row inside this CTA's 128
    const int tid_e = threadIdx.x - 64;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    OutT* outp = reinterpret_cast<OutT*>(p.out);
    const bool has_bias = p.bias != nullptr;
    const uint32_t acc_empty_leader = map_to_cta(smem_u32(&acc_empty[0]), 0);
    uint32_t lt = 0;
    for (int u = pair; u < p.units; u += num_pairs, ++lt) {
      const int tile = u / p.splits, split = u - tile * p.splits;
      const int n_pt = tile / p.m_tiles, m_t = tile - n_pt * p.m_tiles;
      const int n = n_pt * kPairRows + (int)rank * kCtaRows + row;  // output column of this thread
      const bool n_ok = n < p.N;
      const int tok0 = m_t * p.tok_tile;  // accumulator column c <-> token tok0 + c
      const uint32_t as = lt & 1, aphase = (lt >> 1) & 1;
      float* asc = s_ascale + as * kMaxTokTile;
      for (int i = tid_e; i < p.tok_tile; i += 128) {
        const int t = tok0 + i;
        asc[i] = p.a_per_token ? (t < p.M ? p.a_scales[t] : 0.f) : p.a_scales[0];
      }
      const float bsc = p.b_per_channel ? (n_ok ? p.b_scales[n] : 0.f) : p.b_scales[0];
      float biasf = 0.f;
      if (has_bias && n_ok) biasf = to_float<OutT>(reinterpret_cast<const OutT*>(p.bias)[n]);
      named_bar_sync(kEpiBarrier, 128);  // a_scale tile staged (and the previous use of `asc` is over)
      mbar_wait(&acc_full[as], aphase);
      tc_fence_after();
      const uint32_t taddr = lane_addr + as * kAccStride;
      const int ncols = min(p.tok_tile, p.M - tok0);  // valid token columns of this tile

      if (p.splits == 1) {
        for (int c = 0; c < ncols; c += 32) {
          uint32_t r0[16], r1[16];
          tmem_ld_x16(taddr + c, r0);
          const bool second = c + 16 < ncols;
          if (second) tmem_ld_x16(taddr + c + 16, r1);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int t = tok0 + c + j;
            if (n_ok && t < p.M)
              store_scaled(outp, p.ldc, t, n, __uint_as_float(r0[j]), asc[c + j], bsc, biasf, has_bias);
          }
          if (second) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int t = tok0 + c + 16 + j;
              if (n_ok && t < p.M)
                store_scaled(outp, p.ldc, t, n, __uint_as_float(r1[j]), asc[c + 16 + j], bsc, biasf,
                             has_bias);
            }
          }
        }
      } else {
        // ---- split-K: park the columns other slices own, wait for all slices, reduce own columns ----
        // workspace of this (tile, CTA): [owner][source slice][col / 4][128 rows][4] fp32
        const int cw = p.col_chunk, S = p.splits;
        const size_t R = (size_t)cw * kCtaRows;  // floats per (owner, source) region
        float* base = p.partials + (size_t)(tile * 2 + (int)rank) * S * S * R;
        for (int o = 0; o < S; ++o) {
          if (o == split) continue;
          const int c_begin = o * cw, c_end = min(c_begin + cw, p.tok_tile);
          float4* dst = reinterpret_cast<float4*>(base + ((size_t)o * S + split) * R);
          for (int c = c_begin; c < c_end; c += 16) {
            uint32_t r0[16];
            tmem_ld_x16(taddr + c, r0);
            tmem_ld_wait();
#pragma unroll
            for (int jj = 0; jj < 4; ++jj)
              __stcg(dst + (size_t)(((c - c_begin) >> 2) + jj) * kCtaRows + row,
                     make_float4(__uint_as_float(r0[4 * jj]), __uint_as_float(r0[4 * jj + 1]),
                                 __uint_as_float(r0[4 * jj + 2]), __uint_as_float(r0[4 * jj + 3])));
          }
        }
        __threadfence();
        named_bar_sync(kEpiBarrier, 128);
        uint32_t* arrive = p.counters + tile * 2 + (int)rank;
        uint32_t* depart = p.counters + 2 * p.tiles + tile * 2 + (int)rank;
        const int nc = min(cw, p.tok_tile - split * cw);  // columns this slice reduces and writes
        if (tid_e == 0) {
          atomicAdd(arrive, 1u);
          while (ld_acquire_gpu(arrive) < (uint32_t)S) {
          }
          if (nc > 0) {
            // every slice's partials are visible: pull the S-1 foreign regions of my columns into the
            // (now idle) pipeline stages with two bulk copies (the regions left and right of my own)
            fence_proxy_async_all();
            const uint32_t lo = (uint32_t)(split * R * 4), hi = (uint32_t)((S - 1 - split) * R * 4);
            mbar_arrive_expect_tx(red_bar, lo + hi);
            const float* reg = base + (size_t)split * S * R;
            if (lo) bulk_load_1d(smem, reg, lo, red_bar);
            if (hi) bulk_load_1d(smem + lo, reg + (size_t)(split + 1) * R, hi, red_bar);
          }
        }
        if (nc > 0) {
          mbar_wait(red_bar, 0);
          const float4* sp = reinterpret_cast<const float4*>(smem);
          const int cq = cw >> 2;
          for (int c = 0; c < nc; c += 16) {
            uint32_t r0[16];
            tmem_ld_x16(taddr + split * cw + c, r0);
            tmem_ld_wait();
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
              float4 acc = make_float4(__uint_as_float(r0[4 * jj]), __uint_as_float(r0[4 * jj + 1]),
                                       __uint_as_float(r0[4 * jj + 2]), __uint_as_float(r0[4 * jj + 3]));
              for (int s2 = 0; s2 < S - 1; ++s2) {  // own slice first, then the others in slice order
                const float4 v = sp[(size_t)(s2 * cq + (c >> 2) + jj) * kCtaRows + row];
                acc.x += v.x;
                acc.y += v.y;
                acc.z += v.z;
                acc.w += v.w;
              }
              const int cc = split * cw + c + 4 * jj;
              const int t = tok0 + cc;
              if (n_ok) {
                if (t < p.M) store_scaled(outp, p.ldc, t, n, acc.x, asc[cc], bsc, biasf, has_bias);
                if (t + 1 < p.M) store_scaled(outp, p.ldc, t + 1, n, acc.y, asc[cc + 1], bsc, biasf, has_bias);
                if (t + 2 < p.M) store_scaled(outp, p.ldc, t + 2, n, acc.z, asc[cc + 2], bsc, biasf, has_bias);
                if (t + 3 < p.M) store_scaled(outp, p.ldc, t + 3, n, acc.w, asc[cc + 3], bsc, biasf, has_bias);
              }
            }
          }
        }
        named_bar_sync(kEpiBarrier, 128);
        if (tid_e == 0) {
          // the last slice to leave re-arms the counters for the next launch
          if (atomicAdd(depart, 1u) == (uint32_t)S - 1u) {
            atomicExch(arrive, 0u);
            atomicExch(depart, 0u);
          }
        }
      }
      // accumulator stage drained: tell the leader's MMA thread
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(acc_empty_leader + as * 8u);
    }
  }

  // teardown: nobody leaves (or frees TMEM) while the peer may still touch this CTA
  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_2cta(tmem_base, kTmemCols);
  }
}

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

struct Plan {
  int tok_tile, m_tiles, n_ptiles, tiles, stages, stage_bytes, splits, kb_per_split, num_kb, col_chunk,
      units, pairs;
  size_t smem, ws_bytes, main_bytes;
};

static Plan make_plan(int M, int N, int K) {
  Plan g;
  g.m_tiles = ceil_div(M, kMaxTokTile);
  g.tok_tile = (ceil_div(M, g.m_tiles) + 15) / 16 * 16;
  g.n_ptiles = ceil_div(N, kPairRows);
  g.tiles = g.m_tiles * g.n_ptiles;
  g.num_kb = ceil_div(K, kBlockK);
  const int a_bytes = ((g.tok_tile / 2) * kBlockK + 1023) / 1024 * 1024;
  g.stage_bytes = kWTileBytes + a_bytes;
  const int tail = 2 * kMaxStages * 8 + 5 * 8 + 16 + 2 * kMaxTokTile * 4;  // barriers, tmem ptr, a_scale
  const int budget = 227 * 1024 - 1024 /*alignment slack*/ - tail;
  g.stages = budget / g.stage_bytes;
  if (g.stages > kMaxStages) g.stages = kMaxStages;
  const int max_pairs = num_sms() / 2;
  // split-K only when every unit gets its own CTA pair (co-residency makes the in-kernel wait safe)
  int splits = 1;
  for (int s = 2; s <= kMaxSplits; ++s)
    if (g.tiles * s <= max_pairs && g.num_kb / s >= 4) splits = s;
  g.kb_per_split = ceil_div(g.num_kb, splits);
  g.splits = ceil_div(g.num_kb, g.kb_per_split);
  g.col_chunk = (ceil_div(g.tok_tile, g.splits) + 15) / 16 * 16;
  g.units = g.tiles * g.splits;
  g.pairs = g.units < max_pairs ? g.units : max_pairs;
  if (g.stages > g.kb_per_split && g.units <= g.pairs) g.stages = g.kb_per_split < 2 ? 2 : g.kb_per_split;
  size_t main_bytes = (size_t)g.stages * g.stage_bytes;
  const size_t region = (size_t)g.col_chunk * kCtaRows * 4;  // one (owner, source) region
  if (g.splits > 1 && (g.splits - 1) * region > main_bytes)
    main_bytes = ((g.splits - 1) * region + 1023) / 1024 * 1024;
  g.main_bytes = main_bytes;
  g.smem = 1024 + main_bytes + tail;
  g.ws_bytes = g.splits > 1
                   ? (size_t)kCounterBytes + (size_t)g.tiles * 2 * g.splits * g.splits * region
                   : 0;
  return g;
}

template <typename OutT>
static int launch(const Plan& g, const CUtensorMap& tw, const CUtensorMap& ta, const Params& p,
                  cudaStream_t stream) {
  auto kern = scaled_mm_kernel<OutT>;
  VB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * g.pairs, 1, 1);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = g.smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  VB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tw, ta, p));
  return 0;
}

static int run(void* out, const void* a, const void* b, const float* a_scales, const float* b_scales,
               const void* bias, int M, int N, int K, int64_t lda, int64_t ldb, int64_t ldc,
               int a_scales_numel, int b_scales_numel, int out_dtype, void* workspace,
               size_t workspace_bytes, cudaStream_t stream) {
  Plan g = make_plan(M, N, K);
  if (g.splits > 1 && (workspace == nullptr || workspace_bytes < g.ws_bytes)) {
    // no (or too small a) workspace: run unsplit rather than fail
    g.splits = 1;
    g.kb_per_split = g.num_kb;
    g.units = g.tiles;
    g.pairs = g.units < num_sms() / 2 ? g.units : num_sms() / 2;
  }
  CUtensorMap tw, ta;
  if (tc::make_tmap_2d_u8(&tw, b, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, kCtaRows)) return 1;
  if (tc::make_tmap_2d_u8(&ta, a, (uint64_t)M, (uint64_t)K, (uint64_t)lda, (uint32_t)(g.tok_tile / 2)))
    return 1;
  Params p;
  p.out = out;
  p.a_scales = a_scales;
  p.b_scales = b_scales;
  p.bias = bias;
  p.counters = reinterpret_cast<uint32_t*>(workspace);
  p.partials = workspace ? reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + kCounterBytes)
                         : nullptr;
  p.M = M;
  p.N = N;
  p.K = K;
  p.ldc = ldc;
  p.a_per_token = a_scales_numel > 1 ? 1 : 0;
  p.b_per_channel = b_scales_numel > 1 ? 1 : 0;
  p.tok_tile = g.tok_tile;
  p.m_tiles = g.m_tiles;
  p.tiles = g.tiles;
  p.stages = g.stages;
  p.stage_bytes = g.stage_bytes;
  p.splits = g.splits;
  p.kb_per_split = g.kb_per_split;
  p.num_kb = g.num_kb;
  p.col_chunk = g.col_chunk;
  p.units = g.units;
  p.main_bytes = (uint32_t)g.main_bytes;
  // weights are streamed once when there is a single token tile; otherwise they are re-read from L2
  p.hint_w = g.m_tiles == 1 ? kEvictFirst : kEvictNormal;
  p.hint_a = g.m_tiles == 1 ? kEvictLast : kEvictNormal;
  if (out_dtype == VB200_BF16) return launch<__nv_bfloat16>(g, tw, ta, p, stream);
  return launch<__half>(g, tw, ta, p, stream);
}

}  // namespace v2
}  // namespace vb200

using namespace vb200;

extern "C" int vb200_cutlass_scaled_mm_supports_fp8(int cap) { return cap >= 100 ? 1 : 0; }

static bool use_v1() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("VB200_GEMM_V1");
    v = (e && e[0] == '1') ? 1 : 0;
  }
  return v == 1;
}

extern "C" size_t vb200_cutlass_scaled_mm_workspace_size(int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0 || use_v1()) return 0;
  return v2::make_plan(M, N, K).ws_bytes;
}

extern "C" int vb200_cutlass_scaled_mm(void* out, const void* a, const void* b,
                                       const float* a_scales, const float* b_scales,
                                       const void* bias, int M, int N, int K, int64_t lda,
                                       int64_t ldb, int64_t ldc, int a_scales_numel,
                                       int b_scales_numel, int out_dtype, void* workspace,
                                       size_t workspace_bytes, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(out_dtype == VB200_BF16 || out_dtype == VB200_F16,
           "cutlass_scaled_mm: out dtype must be bf16 or f16");
  VB_CHECK(K % 16 == 0 && N % 16 == 0, "cutlass_scaled_mm: K (%d) and N (%d) must be multiples of 16",
           K, N);
  VB_CHECK(a_scales_numel == 1 || a_scales_numel == M, "a_scales must have 1 or M entries");
  VB_CHECK(b_scales_numel == 1 || b_scales_numel == N, "b_scales must have 1 or N entries");
  if (M <= 0 || N <= 0) return 0;
  VB_CHECK(K > 0, "cutlass_scaled_mm: K must be positive");
  if (!use_v1())
    return v2::run(out, a, b, a_scales, b_scales, bias, M, N, K, lda, ldb, ldc, a_scales_numel,
                   b_scales_numel, out_dtype, workspace, workspace_bytes, stream);
  v1::GemmPlan g = v1::plan_gemm(M, N, K);
  VB_CHECK(g.m_tiles <= 65535, "cutlass_scaled_mm: M too large");
  CUtensorMap tw, ta;
  if (tc::make_tmap_2d_u8(&tw, b, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, v1::kBlockN)) return 1;
  if (tc::make_tmap_2d_u8(&ta, a, (uint64_t)M, (uint64_t)K, (uint64_t)lda, (uint32_t)g.tok_tile))
    return 1;
  v1::GemmParams p;
  p.out = out;
  p.a_scales = a_scales;
  p.b_scales = b_scales;
  p.bias = bias;
  p.M = M;
  p.N = N;
  p.K = K;
  p.ldc = ldc;
  p.a_per_token = a_scales_numel > 1 ? 1 : 0;
  p.b_per_channel = b_scales_numel > 1 ? 1 : 0;
  p.tok_tile = g.tok_tile;
  p.stages = g.stages;
  p.splits = g.splits;
  p.kb_per_split = g.kb_per_split;
  p.num_kb = g.num_kb;
  p.col_chunk = g.col_chunk;
  p.tmem_cols = g.tmem_cols;
  if (out_dtype == VB200_BF16) return v1::launch_gemm<__nv_bfloat16>(g, tw, ta, p, stream);
  return v1::launch_gemm<__half>(g, tw, ta, p, stream);
}
