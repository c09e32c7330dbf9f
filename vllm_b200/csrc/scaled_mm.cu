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
  uint32_t* counters;     // split-K: arrive[2 * tiles], zero between launches (self re-arming)
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
  unsigned long long* trace;  // debug: [grid][16] %globaltimer stamps (nullptr = off)
};

__device__ __forceinline__ void stamp(const Params& p, int slot) {
  if (p.trace != nullptr) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.trace[(size_t)blockIdx.x * 16 + slot] = t;
  }
}

__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}

// Epilogue of 16 consecutive token columns of one weight row: out[t0 + j][n] = a_scale[j] * (b_scale * acc[j])
// (+ bias)  (ScaledEpilogue / ScaledEpilogueBias).  The 16 a_scales are read from shared memory with explicit
// ld.shared (a generic load here would be ordered after every preceding global store and serialise the loop).
template <typename OutT>
__device__ __forceinline__ void store_chunk16(OutT* orow, int64_t ldc, const float (&acc)[16], uint32_t asc_smem,
                                              float bsc, float biasf, bool has_bias, int nvalid) {
  float a[16];
#pragma unroll
  for (int q4 = 0; q4 < 4; ++q4) {
    const float4 v = lds_f4(asc_smem + 16 * q4);
    a[4 * q4] = v.x;
    a[4 * q4 + 1] = v.y;
    a[4 * q4 + 2] = v.z;
    a[4 * q4 + 3] = v.w;
  }
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const float sv = bsc * acc[j];
    const float v = has_bias ? fmaf(a[j], sv, biasf) : a[j] * sv;
    if (j < nvalid) *orow = from_float<OutT>(v);
    orow += ldc;
  }
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
  float* s_ascale = reinterpret_cast<float*>(tmem_ptr + 6);  // [2][256], 16-byte aligned (offset 256 past the ring)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();  // 0 = leader (issues the MMAs)
  const int pair = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;
  const int tok_half = p.tok_tile >> 1;
  if (threadIdx.x == 0) stamp(p, 0);

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
  // everything above (barriers, TMEM, tensor-map prefetch) overlapped the previous kernel's tail; from here on the
  // kernel reads what that kernel wrote (activations, scales) and writes buffers it may still be reading
  pdl_wait();
  pdl_trigger();
  if (threadIdx.x == 0) stamp(p, 1);

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
          if (it == 0) stamp(p, 2);
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
          if (it == 0) stamp(p, 3);
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
        stamp(p, 4);
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
      if (tid_e == 0) stamp(p, 5);
      const uint32_t taddr = lane_addr + as * kAccStride;
      const int ncols = min(p.tok_tile, p.M - tok0);  // valid token columns of this tile

      const uint32_t asc_u32 = smem_u32(asc);
      // 16 accumulator columns (registers r[off .. off+15]) -> scaled, converted, stored
      auto emit16 = [&](const uint32_t* r, int c) {
        float f[16];
#pragma unroll
        for (int jx = 0; jx < 16; ++jx) f[jx] = __uint_as_float(r[jx]);
        if (n_ok)
          store_chunk16(outp + (int64_t)(tok0 + c) * p.ldc + n, p.ldc, f, asc_u32 + 4 * c, bsc, biasf, has_bias,
                        ncols - c);
      };
      if (p.splits == 1) {
        // tcgen05.ld is warp-collective (never under a per-lane predicate) and asynchronous: the load of the
        // next 32 columns is in flight while the current 32 are converted and stored
        uint32_t ra[32], rb[32];
        tmem_ld_x32(taddr, ra);
        for (int c = 0; c < ncols; c += 64) {
          tmem_ld_wait();
          const bool more1 = c + 32 < ncols;
          if (more1) tmem_ld_x32(taddr + c + 32, rb);
          emit16(ra, c);
          if (c + 16 < ncols) emit16(ra + 16, c + 16);
          if (more1) {
            tmem_ld_wait();
            if (c + 64 < ncols) tmem_ld_x32(taddr + c + 64, ra);
            emit16(rb, c + 32);
            if (c + 48 < ncols) emit16(rb + 16, c + 48);
          }
        }
      } else {
        // ---- split-K: park the columns other slices own, wait for all slices, reduce own columns ----
        // workspace of this (tile, CTA): [owner][source slice][col / 4][128 rows][4] fp32
        const int cw = p.col_chunk, S = p.splits;
        const size_t R = (size_t)cw * kCtaRows;  // floats per (owner, source) region
        float* base = p.partials + (size_t)(tile * 2 + (int)rank) * S * S * R;
        const int tpc = p.tok_tile >> 4;  // 16-column chunks of the tile
        auto next_foreign = [&](int jc) {
          while (jc < tpc && (jc << 4) / cw == split) ++jc;
          return jc;
        };
        auto park16 = [&](const uint32_t* r, int jc) {
          const int c = jc << 4, o = c / cw;
          float4* dst = reinterpret_cast<float4*>(base + ((size_t)o * S + split) * R) +
                        (size_t)((c - o * cw) >> 2) * kCtaRows + row;
#pragma unroll
          for (int jj = 0; jj < 4; ++jj)
            __stcg(dst + (size_t)jj * kCtaRows,
                   make_float4(__uint_as_float(r[4 * jj]), __uint_as_float(r[4 * jj + 1]),
                               __uint_as_float(r[4 * jj + 2]), __uint_as_float(r[4 * jj + 3])));
        };
        {
          uint32_t ra[16], rb[16];
          int j0 = next_foreign(0);
          if (j0 < tpc) tmem_ld_x16(taddr + (j0 << 4), ra);
          while (j0 < tpc) {
            const int j1 = next_foreign(j0 + 1);
            tmem_ld_wait();
            if (j1 < tpc) tmem_ld_x16(taddr + (j1 << 4), rb);
            park16(ra, j0);
            if (j1 >= tpc) break;
            const int j2 = next_foreign(j1 + 1);
            tmem_ld_wait();
            if (j2 < tpc) tmem_ld_x16(taddr + (j2 << 4), ra);
            park16(rb, j1);
            j0 = j2;
          }
        }
        __threadfence();
        if (tid_e == 0) stamp(p, 6);
        named_bar_sync(kEpiBarrier, 128);
        // arrival counter that re-arms itself: atomicInc wraps to 0 on the S-th arrival, and 0 is what every
        // waiter (who has incremented it itself) waits for; it is 0 again for the next launch
        uint32_t* arrive = p.counters + tile * 2 + (int)rank;
        const int nc = min(cw, p.tok_tile - split * cw);  // columns this slice reduces and writes
        const int c_own = split * cw;
        uint32_t ra[16], rb[16];
        if (nc > 0) tmem_ld_x16(taddr + c_own, ra);  // own columns: in flight during the wait below
        if (tid_e == 0) {
          atomicInc(arrive, (uint32_t)S - 1u);
          while (ld_acquire_gpu(arrive) != 0u) {
          }
          stamp(p, 7);
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
          if (tid_e == 0) stamp(p, 8);
          const uint32_t sp = smem_u32(smem) + (uint32_t)row * 16u;
          const int cq = cw >> 2;
          auto reduce16 = [&](const uint32_t* r, int c) {  // c: column offset inside my chunk
            float f[16];
#pragma unroll
            for (int jx = 0; jx < 16; ++jx) f[jx] = __uint_as_float(r[jx]);
            for (int s2 = 0; s2 < S - 1; ++s2) {  // own slice first, then the others in slice order
#pragma unroll
              for (int jj = 0; jj < 4; ++jj) {
                const float4 v = lds_f4(sp + (uint32_t)((s2 * cq + (c >> 2) + jj) * kCtaRows * 16));
                f[4 * jj] += v.x;
                f[4 * jj + 1] += v.y;
                f[4 * jj + 2] += v.z;
                f[4 * jj + 3] += v.w;
              }
            }
            const int cc = c_own + c;
            if (n_ok)
              store_chunk16(outp + (int64_t)(tok0 + cc) * p.ldc + n, p.ldc, f, asc_u32 + 4 * cc, bsc, biasf, has_bias,
                            p.M - tok0 - cc);
          };
          for (int c = 0; c < nc; c += 32) {
            tmem_ld_wait();
            const bool more1 = c + 16 < nc;
            if (more1) tmem_ld_x16(taddr + c_own + c + 16, rb);
            reduce16(ra, c);
            if (more1) {
              tmem_ld_wait();
              if (c + 32 < nc) tmem_ld_x16(taddr + c_own + c + 32, ra);
              reduce16(rb, c + 16);
            }
          }
        }
      }
      if (tid_e == 0) stamp(p, 9);
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
  if (threadIdx.x == 0) stamp(p, 10);
}

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

static unsigned long long* g_trace = nullptr;  // set through vb200_debug_scaled_mm_trace()

struct Plan {
  int tok_tile, m_tiles, n_ptiles, tiles, stages, stage_bytes, splits, kb_per_split, num_kb, col_chunk,
      units, pairs;
  size_t smem, ws_bytes, main_bytes;
};

// Cost model (microseconds) of one (token tiles, K splits) choice for a problem whose units are all co-resident
// or run in whole rounds.  Constants are read off the per-phase trace (benchmarks/trace_scaled_mm.py,
// profiles/r01_scaled_mm_trace_s18.txt): a CTA pulls ~90 GB/s from L2, a pair retires one 256 x 256 x 128 k-block
// in 0.28 us, the split-K exchange (park partials in L2, wait, fetch, store) costs ~9 us per 256 token columns and
// the plain epilogue ~2.6 us, HBM streams the weights at ~6.4 TB/s.
static float plan_cost(int tok_tile, int tiles, int splits, int num_kb, int max_pairs, double w_bytes) {
  const int kbps = ceil_div(num_kb, splits);
  const int rounds = ceil_div(tiles * splits, max_pairs);
  const float feed = kbps * (kWTileBytes + tok_tile / 2 * kBlockK) / 90e3f;
  const float mma = kbps * tok_tile * 0.0011f;
  const float hbm = (float)(w_bytes / 6.4e6);
  const float main_us = fmaxf(rounds * fmaxf(feed, mma), hbm);
  return main_us + (splits > 1 ? 0.035f : 0.010f) * tok_tile;
}

static Plan make_plan(int M, int N, int K) {
  Plan g;
  g.m_tiles = ceil_div(M, kMaxTokTile);
  g.n_ptiles = ceil_div(N, kPairRows);
  g.num_kb = ceil_div(K, kBlockK);
  const int max_pairs = num_sms() / 2;
  // split-K only when every unit gets its own CTA pair (co-residency makes the in-kernel wait safe)
  int splits = 1;
  for (int s = 2; s <= kMaxSplits; ++s)
    if (g.m_tiles * g.n_ptiles * s <= max_pairs && g.num_kb / s >= 4) splits = s;
  if (M > 128 && M <= kMaxTokTile && 2 * g.n_ptiles <= max_pairs) {
    // one token tile of 129..256 columns: splitting the TOKENS over two pairs (weights re-read from L2, no
    // partial-sum traffic) can beat splitting K, whose exchange moves M x N x 4 bytes per slice through L2
    const double w_bytes = (double)N * K;
    float best = plan_cost((M + 15) / 16 * 16, g.n_ptiles, splits, g.num_kb, max_pairs, w_bytes);
    const int tok2 = (ceil_div(M, 2) + 15) / 16 * 16;
    for (int s = 1; s <= kMaxSplits; ++s) {
      if (s > 1 && !(2 * g.n_ptiles * s <= max_pairs && g.num_kb / s >= 4)) break;
      const float c = plan_cost(tok2, 2 * g.n_ptiles, s, g.num_kb, max_pairs, w_bytes);
      if (c < best * 0.95f) {
        best = c;
        g.m_tiles = 2;
        splits = s;
      }
    }
  }
  g.tok_tile = (ceil_div(M, g.m_tiles) + 15) / 16 * 16;
  g.tiles = g.m_tiles * g.n_ptiles;
  const int a_bytes = ((g.tok_tile / 2) * kBlockK + 1023) / 1024 * 1024;
  g.stage_bytes = kWTileBytes + a_bytes;
  const int tail = 2 * kMaxStages * 8 + 5 * 8 + 24 + 2 * kMaxTokTile * 4;  // barriers, tmem ptr, a_scale
  const int budget = 227 * 1024 - 1024 /*alignment slack*/ - tail;
  g.stages = budget / g.stage_bytes;
  if (g.stages > kMaxStages) g.stages = kMaxStages;
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
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
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
  p.trace = g_trace;
  if (out_dtype == VB200_BF16) return launch<__nv_bfloat16>(g, tw, ta, p, stream);
  return launch<__half>(g, tw, ta, p, stream);
}

}  // namespace v2
}  // namespace vb200

using namespace vb200;

// Debug / profiling hook (not part of the drop-in surface): when `buf` is non-NULL every CTA of the following
// cutlass_scaled_mm launches writes 16 %globaltimer stamps (ns) at buf[blockIdx.x * 16 + i]:
// 0 entry, 1 set-up done, 2 first TMA issued, 3 first stage landed, 4 last MMA committed, 5 accumulator ready,
// 6 split-K partials parked, 7 all slices arrived, 8 foreign partials fetched, 9 epilogue done, 10 exit.
extern "C" void vb200_debug_scaled_mm_trace(void* buf) { v2::g_trace = (unsigned long long*)buf; }

extern "C" int vb200_cutlass_scaled_mm_supports_fp8(int cap) { return cap >= 100 ? 1 : 0; }

extern "C" size_t vb200_cutlass_scaled_mm_workspace_size(int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
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
  return v2::run(out, a, b, a_scales, b_scales, bias, M, N, K, lda, ldb, ldc, a_scales_numel,
                 b_scales_numel, out_dtype, workspace, workspace_bytes, stream);
}
