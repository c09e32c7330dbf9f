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
//   * Warp roles per CTA: warps 0-2 = TMA producers, k-blocks dealt round-robin (128-byte-swizzled [rows x 128 B]
//     boxes into a ring of stages; both CTAs signal the LEADER's full barrier; three issuing threads because one
//     thread sustains only one bulk copy per ~170 ns, see kProdWarps), warp 3 = MMA issuer (leader CTA only,
//     one thread; tcgen05.commit multicasts "stage free" / "accumulator ready" to both CTAs),
//     warps 4-7 = epilogue (tcgen05.ld -> a_scale * (b_scale * acc) (+ bias) -> bf16/f16 -> global).
//   * Split-K when the tile count cannot fill the machine (decode o_proj / down_proj / qkv): the S pairs that share
//     an output tile form ONE thread-block cluster of 2 S CTAs (cluster rank = 2 * slice + half), each pair
//     accumulates a K slice, and the partial accumulators are exchanged through DISTRIBUTED SHARED MEMORY: after a
//     cluster barrier ("my MMAs are done, my stage ring is idle") every CTA pushes the token columns it does not own
//     into the owner's ring with st.shared::cluster.v4 (TMEM -> registers -> peer SM, nothing goes through L2), a
//     second cluster barrier publishes them, and every CTA reduces + writes 1/S of the token columns in slice order
//     (deterministic).  No global workspace, no tickets, no spin-wait: the hardware co-schedules a cluster, so there
//     is nothing to deadlock on and nothing shared between streams.
//   * More tiles than pairs (gate_up: 112 tiles on 74 pairs): whole tiles round-robin, the second round partly empty.
//     A stream-K variant (equal k-block ranges per pair, cut tiles fixed up through an L2 workspace) was built and
//     measured: it paid off only while the issue loops were slow; after the warp-converged issue fix whole tiles win
//     at every batch (gate_up, M = 16 / 64 / 128 / 256: 20.1 / 20.7 / 22.2 / 29.6 us whole tiles vs 22.4 / 24.4 / 26.7 /
//     37.5 us stream-K, profiles/r02_scaled_mm_plan_sweep_s52.jsonl, _s36) and it was removed.
//   * Programmatic dependent launch: barrier init, TMEM allocation, tensor-map prefetch AND the first ring-full of
//     WEIGHT tiles (which no earlier kernel writes) are issued BEFORE griddepcontrol.wait, so they overlap the tail
//     of whatever ran before; only the activation loads, the scale / bias reads and the stores wait.
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
constexpr int kEpiWarps = 4;   // one per TMEM lane quarter.  (8 = two per quarter taking alternate column chunks was
                               // measured SLOWER, profiles/r02_scaled_mm_plan_sweep_s34.jsonl: the epilogue is bound
                               // by the store / DSMEM paths, not by instruction issue)
constexpr int kWpq = kEpiWarps / 4;   // warps per lane quarter
// TMA producer warps.  ONE issuing thread sustains one cp.async.bulk(.tensor) per ~170 ns whatever its size up to
// 16 KiB (benchmarks/l2_feed_probe.cu, profiles/r02_l2_feed_probe_s37.jsonl: 16 / 8 / 4 KiB chunks all land at 5.8 M
// copies/s per SM, 93 / 47 / 23 GB/s) -- and the main loop of the single-producer kernel ran at exactly two copies
// (W box + A box) per 340-390 ns k-block (profiles/r02_scaled_mm_trace_s36.txt), i.e. issue-bound at 50-85 GB/s per
// SM, below both the L2 feed (130 GB/s per SM) and the HBM stream.  Three warps take the k-blocks round-robin
// (3 + MMA + 4 epilogue = 8 warps = two per scheduler, which keeps the 230-register epilogue free of spills; a ninth
// warp would cap the kernel at 168 registers).
constexpr int kProdWarps = 3;
constexpr int kMmaWarp = kProdWarps;          // warp index of the MMA issuer
constexpr int kFirstEpiWarp = kProdWarps + 1; // epilogue warps 4..7 -> TMEM lane quarters 0..3 (= warp % 4)
constexpr int kThreads = 32 * (kFirstEpiWarp + kEpiWarps);
constexpr int kWTileBytes = kCtaRows * kBlockK;  // 16 KiB
constexpr int kTmemCols = 512;                   // two accumulator stages of <= 256 columns
constexpr int kAccStride = 256;                  // TMEM column offset of accumulator stage 1
constexpr int kMaxSplits = 8;                    // cluster of 2 * 8 = 16 CTAs (non-portable size above 4 slices)
constexpr int kEpiBarrier = 1;                   // named barrier id of the epilogue warps

struct Params {
  void* out;              // [M, N] out_t, row stride ldc
  const float* a_scales;  // 1 or M
  const float* b_scales;  // 1 or N
  const void* bias;       // nullptr or [N] out_t
  int M, N, K;
  int64_t ldc;
  int a_per_token, b_per_channel;
  int tok_tile;      // tokens per tile (multiple of 16, <= 256); each CTA stages tok_tile / 2 rows
  int m_tiles;       // token tiles
  int tiles;         // m_tiles * ceil(N / 256)
  int stages;        // shared-memory pipeline depth
  int stage_bytes;   // per CTA: 16 KiB of W + the half token tile rounded up to 1 KiB
  int splits;        // K slices per tile (1 = none) = pairs per cluster; a cluster works on one tile at a time
  int kb_per_split;  // k-blocks per slice
  int num_kb;        // total k-blocks
  int col_chunk;     // split-K: token columns reduced + written by one slice (multiple of 16)
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

// The work list of one pair (cluster): whole tiles cluster_id, cluster_id + num_clusters, ... restricted to the
// pair's K slice [kb0, kb1).
struct SegIter {
  int state, end;
  __device__ __forceinline__ void init(const Params& p, int cluster_id, int num_clusters) {
    (void)num_clusters;
    state = cluster_id;
    end = p.tiles;
  }
  __device__ __forceinline__ bool next(const Params& p, int num_clusters, int kb0, int kb1, int& tile, int& lo, int& hi) {
    (void)p;
    if (state >= end) return false;
    tile = state;
    lo = kb0;
    hi = kb1;
    state += num_clusters;
    return true;
  }
};

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
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(acc_empty + 3);
  float* s_ascale = reinterpret_cast<float*>(tmem_ptr + 6);  // [2][256], 16-byte aligned (offset 256 past the ring)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = cluster_ctarank();  // 2 * K-slice + half
  const uint32_t rank = crank & 1u;          // half of the pair; 0 = leader (issues the MMAs)
  const uint32_t leader = crank & ~1u;       // cluster rank of this pair's leader
  const int split = (int)(crank >> 1);       // K slice of this pair (0 when splits == 1)
  const int cluster_ctas = 2 * p.splits;
  const int cluster_id = blockIdx.x / cluster_ctas;
  const int num_clusters = gridDim.x / cluster_ctas;
  const int tok_half = p.tok_tile >> 1;
  const int kb0 = split * p.kb_per_split;                  // this pair's K slice (the same for every tile)
  const int kb1 = min(kb0 + p.kb_per_split, p.num_kb);
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
      mbar_init(&acc_empty[s], 2 * kEpiWarps);  // epilogue warps x 2 CTAs
    }
    fence_barrier_init();
  }
  if (warp == kMmaWarp) tmem_alloc_2cta(tmem_ptr, kTmemCols);
  tc_fence_before();
  cluster_sync_all();  // barrier inits + TMEM allocation visible in the whole cluster
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_trigger();       // the next kernel on the stream may start its own prologue
  if (threadIdx.x == 0) stamp(p, 1);

  if (warp < kProdWarps) {
    // ===== TMA producers (both CTAs; one thread of each of kProdWarps warps; warp w owns k-block iterations
    // it = w, w + kProdWarps, ... of this CTA's work list: the barrier protocol is per stage, so the warps never
    // have to order themselves against each other) =====
    // (the whole warp runs the loops and waits; one elected lane issues -- see the MMA issuer below for why)
    {
      const uint32_t bar_base = map_to_cta(smem_u32(&full_bar[0]), leader);
      const uint32_t tx_bytes = 2u * (uint32_t)(kWTileBytes + tok_half * kBlockK);
      // ---- before the grid dependency resolves: the first ring-full of WEIGHT tiles of the first tile.  Weights
      // are not written by any kernel, so streaming them may overlap the previous kernel's tail; the barrier of each
      // stage is armed for W + A bytes here, the activation half follows after griddepcontrol.wait.
      uint32_t npre = 0;
      SegIter si;
      si.init(p, cluster_id, num_clusters);
      int tile, lo, hi;
      {
        SegIter s0 = si;
        if (s0.next(p, num_clusters, kb0, kb1, tile, lo, hi)) {
          const int n_pt = tile / p.m_tiles;
          const int n0 = n_pt * kPairRows + (int)rank * kCtaRows;
          npre = (uint32_t)min(hi - lo, p.stages);
          for (uint32_t i = (uint32_t)warp; i < npre; i += kProdWarps) {
            if (elect_one()) {
              if (rank == 0) mbar_arrive_expect_tx(&full_bar[i], tx_bytes);
              tma_load_2d_2cta(smem + (size_t)i * p.stage_bytes, &tmap_w, bar_base + 8u * i, (lo + (int)i) * kBlockK,
                               n0, p.hint_w);
            }
            __syncwarp();
          }
          if (warp == 0 && lane == 0) stamp(p, 2);
        }
      }
      pdl_wait();   // activations (and everything else the previous kernels wrote) are visible from here on
      // (stage, phase, owner) are carried as running counters: this is latency-bound code, and a runtime
      // `it % stages` / `it / stages` pair costs more cycles than the TMA issue itself
      uint32_t it = 0;
      int stage = 0, owner = 0;
      uint32_t phase = 0;
      auto advance = [&]() {
        ++it;
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1u;
        }
        if (++owner == kProdWarps) owner = 0;
      };
      while (si.next(p, num_clusters, kb0, kb1, tile, lo, hi)) {
        const int n_pt = tile / p.m_tiles, m_t = tile - n_pt * p.m_tiles;
        const int n0 = n_pt * kPairRows + (int)rank * kCtaRows;
        const int tok0 = m_t * p.tok_tile + (int)rank * tok_half;
        for (int kb = lo; kb < hi; ++kb, advance()) {
          if (owner != warp) continue;
          uint8_t* sw = smem + (size_t)stage * p.stage_bytes;
          uint8_t* sa = sw + kWTileBytes;
          const uint32_t bar = bar_base + 8u * (uint32_t)stage;
          const bool fresh = it >= npre;   // W of the first `npre` k-blocks is already in flight
          if (fresh) mbar_wait(&empty_bar[stage], phase ^ 1);
          if (elect_one()) {
            if (fresh) {
              if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
              tma_load_2d_2cta(sw, &tmap_w, bar, kb * kBlockK, n0, p.hint_w);
            }
            tma_load_2d_2cta(sa, &tmap_a, bar, kb * kBlockK, tok0, p.hint_a);
          }
          __syncwarp();
        }
      }
      // tail: every stage this CTA filled has been released, i.e. all multicast arrivals aimed at
      // this CTA's barriers have landed before it can exit (each warp drains the iterations it owns)
      const uint32_t n_tail = it < (uint32_t)p.stages ? it : (uint32_t)p.stages;
      for (uint32_t i = 0; i < n_tail; ++i, advance()) {
        if (owner != warp) continue;
        mbar_wait(&empty_bar[stage], phase ^ 1);
      }
    }
    __syncwarp();
  } else if (warp == kMmaWarp) {
    // ===== MMA issuer (leader CTA of each pair).  The WHOLE warp runs the loop and waits on the barriers; one
    // elected lane issues the tcgen05 instructions.  Keeping the warp converged is what lets the compiler hold the
    // descriptors / TMEM addresses / barrier addresses in uniform registers: issued from a `lane == 0` branch, every
    // UTCQMMA and UTCBAR was wrapped in an ELECT + 5 x R2UR.BROADCAST "waterfall" loop (69 cycles per MMA, 440 per
    // k-block measured in the issuer: profiles/r02_scaled_mm_trace_s42.txt) =====
    if (rank == 0) {
      const uint32_t idesc = make_idesc_f8(kPairRows, p.tok_tile);
      const uint16_t pair_mask = (uint16_t)(0x3u << leader);   // the two CTAs of this pair, as cluster ranks
      const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);   // warp-uniform copy of the TMEM base
      const uint32_t smem_base_u = smem_u32(smem);
      uint32_t it = 0, lt = 0;
      int stage = 0;
      uint32_t phase = 0;
      const long long loop_c0 = p.trace != nullptr ? clock64() : 0;
      SegIter si;
      si.init(p, cluster_id, num_clusters);
      int tile, lo, hi;
      for (; si.next(p, num_clusters, kb0, kb1, tile, lo, hi); ++lt) {
        const int nkb = hi - lo;
        const uint32_t as = lt & 1, aphase = (lt >> 1) & 1;
        mbar_wait(&acc_empty[as], aphase ^ 1);  // both CTAs' epilogues drained this stage
        tc_fence_after();
        const uint32_t d_tmem = tmem_u + as * kAccStride;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          mbar_wait(&full_bar[stage], phase);  // both CTAs' boxes have landed
          tc_fence_after();
          if (it == 0 && lane == 0) stamp(p, 3);
          const uint32_t sw = smem_base_u + (uint32_t)stage * (uint32_t)p.stage_bytes;
          const uint64_t adesc = make_smem_desc_sw128(sw);
          const uint64_t bdesc = make_smem_desc_sw128(sw + kWTileBytes);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < kBlockK / kUmmaK; ++k) {
              // advance both descriptors by 32 bytes of K inside the swizzle row (address field >> 4)
              umma_f8_2cta(d_tmem, adesc + (uint64_t)(k * (kUmmaK >> 4)),
                           bdesc + (uint64_t)(k * (kUmmaK >> 4)), idesc, (kb | k) != 0);
            }
            umma_commit_2cta(&empty_bar[stage], pair_mask);  // frees the stage in both CTAs
          }
          __syncwarp();
          // running (stage, phase): no runtime division in this latency-bound loop
          if (++stage == p.stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        if (elect_one()) umma_commit_2cta(&acc_full[as], pair_mask);  // accumulator complete: epilogues may read
        __syncwarp();
        if (lane == 0) stamp(p, 4);
      }
      if (p.trace != nullptr && lane == 0) {   // debug: the issuer's k-block count and its whole-loop cycle count
        p.trace[(size_t)blockIdx.x * 16 + 14] = (unsigned long long)it;
        p.trace[(size_t)blockIdx.x * 16 + 15] = (unsigned long long)(clock64() - loop_c0);
      }
    }
    __syncwarp();
  }
  if (warp < kFirstEpiWarp && p.splits > 1) {
    // split-K: the two cluster barriers of the DSMEM exchange are executed by every thread of the cluster
    cluster_sync_all();
    cluster_sync_all();
  }

  if (warp >= kFirstEpiWarp) {
    // ===== epilogue (kEpiWarps warps of each CTA; TMEM lane quarter = warp % 4, the kWpq warps of a quarter take
    // alternate column chunks) =====
    const int q = warp & 3;
    const int h = (warp - kFirstEpiWarp) >> 2;  // which of the quarter's warps
    const int row = q * 32 + lane;  // weight row inside this CTA's 128
    const int tid_e = threadIdx.x - 32 * kFirstEpiWarp;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    OutT* outp = reinterpret_cast<OutT*>(p.out);
    const bool has_bias = p.bias != nullptr;
    const uint32_t acc_empty_leader = map_to_cta(smem_u32(&acc_empty[0]), leader);
    uint32_t lt = 0;
    pdl_wait();   // scales / bias may come from the previous kernel; `out` may still be read by it
    SegIter si;
    si.init(p, cluster_id, num_clusters);
    int tile, seg_lo, seg_hi;
    for (; si.next(p, num_clusters, kb0, kb1, tile, seg_lo, seg_hi); ++lt) {
      const int n_pt = tile / p.m_tiles, m_t = tile - n_pt * p.m_tiles;
      const int n = n_pt * kPairRows + (int)rank * kCtaRows + row;  // output column of this thread
      const bool n_ok = n < p.N;
      const int tok0 = m_t * p.tok_tile;  // accumulator column c <-> token tok0 + c
      const uint32_t as = lt & 1, aphase = (lt >> 1) & 1;
      float* asc = s_ascale + as * kMaxTokTile;
      for (int i = tid_e; i < p.tok_tile; i += 32 * kEpiWarps) {
        const int t = tok0 + i;
        asc[i] = p.a_per_token ? (t < p.M ? p.a_scales[t] : 0.f) : p.a_scales[0];
      }
      const float bsc = p.b_per_channel ? (n_ok ? p.b_scales[n] : 0.f) : p.b_scales[0];
      float biasf = 0.f;
      if (has_bias && n_ok) biasf = to_float<OutT>(reinterpret_cast<const OutT*>(p.bias)[n]);
      named_bar_sync(kEpiBarrier, 32 * kEpiWarps);  // a_scale tile staged (and the previous use of `asc` is over)
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
        int c = 32 * h;   // this warp: 32-column blocks h, h + kWpq, h + 2 kWpq, ...
        if (c < ncols) tmem_ld_x32(taddr + c, ra);
        for (; c < ncols; c += 64 * kWpq) {
          tmem_ld_wait();
          const int c2 = c + 32 * kWpq;
          if (c2 < ncols) tmem_ld_x32(taddr + c2, rb);
          emit16(ra, c);
          if (c + 16 < ncols) emit16(ra + 16, c + 16);
          if (c2 < ncols) {
            tmem_ld_wait();
            if (c2 + 32 * kWpq < ncols) tmem_ld_x32(taddr + c2 + 32 * kWpq, ra);
            emit16(rb, c2);
            if (c2 + 16 < ncols) emit16(rb + 16, c2 + 16);
          }
        }
      } else {
        // ---- split-K through distributed shared memory (one tile per cluster in this mode) ----
        // receive buffer of a CTA = its (now idle) stage ring: [source slot][col / 4][128 rows][4] fp32, source slot
        // = the sending slice's index among the S - 1 foreign slices, in slice order
        const int cw = p.col_chunk, S = p.splits;
        const int cq = cw >> 2;
        const int tpc = p.tok_tile >> 4;  // 16-column chunks of the tile
        const uint32_t ring_u32 = smem_u32(smem);
        // barrier 1: every CTA of the cluster has seen its accumulator complete => all MMAs that read the rings are
        // done and all TMA loads have landed: the rings may be overwritten by peers
        cluster_sync_all();
        // foreign 16-column chunks (everything but my own slice's columns), dealt alternately to the two warps
        const int own_lo = (split * cw) >> 4, own_n = cw >> 4;
        auto next_foreign = [&](int jc) {
          for (; jc < tpc; ++jc) {
            if (jc >= own_lo && jc < own_lo + own_n) continue;
            const int fi = jc < own_lo ? jc : jc - own_n;
            if (fi % kWpq == h) break;
          }
          return jc;
        };
        auto push16 = [&](const uint32_t* r, int jc) {
          const int c = jc << 4, o = c / cw;
          const int slot = split < o ? split : split - 1;
          const uint32_t dst = map_to_cta(ring_u32, (uint32_t)(2 * o) + rank) +
                               (uint32_t)(((slot * cq + ((c - o * cw) >> 2)) * kCtaRows + row) * 16);
#pragma unroll
          for (int jj = 0; jj < 4; ++jj)
            st_cluster_v4(dst + (uint32_t)(jj * kCtaRows * 16), r[4 * jj], r[4 * jj + 1], r[4 * jj + 2], r[4 * jj + 3]);
        };
        {
          uint32_t ra[16], rb[16];
          int j0 = next_foreign(0);
          if (j0 < tpc) tmem_ld_x16(taddr + (j0 << 4), ra);
          while (j0 < tpc) {
            const int j1 = next_foreign(j0 + 1);
            tmem_ld_wait();
            if (j1 < tpc) tmem_ld_x16(taddr + (j1 << 4), rb);
            push16(ra, j0);
            if (j1 >= tpc) break;
            const int j2 = next_foreign(j1 + 1);
            tmem_ld_wait();
            if (j2 < tpc) tmem_ld_x16(taddr + (j2 << 4), ra);
            push16(rb, j1);
            j0 = j2;
          }
        }
        if (tid_e == 0) stamp(p, 6);
        const int nc = min(cw, p.tok_tile - split * cw);  // columns this slice reduces and writes
        const int c_own = split * cw;
        uint32_t ra[16], rb[16];
        int c = 16 * h;   // this warp: 16-column chunks h, h + kWpq, ... of my slice's columns
        if (c < nc) tmem_ld_x16(taddr + c_own + c, ra);  // in flight across the barrier
        // barrier 2 (release / acquire at cluster scope): the peers' partials of my columns are in my ring
        cluster_sync_all();
        if (tid_e == 0) stamp(p, 7);
        {
          const uint32_t sp = ring_u32 + (uint32_t)row * 16u;
          auto reduce16 = [&](const uint32_t* r, int cc0) {  // cc0: column offset inside my chunk
            float f[16];
#pragma unroll
            for (int jx = 0; jx < 16; ++jx) f[jx] = __uint_as_float(r[jx]);
            for (int s2 = 0; s2 < S - 1; ++s2) {  // own slice first, then the others in slice order
#pragma unroll
              for (int jj = 0; jj < 4; ++jj) {
                const float4 v = lds_f4(sp + (uint32_t)((s2 * cq + (cc0 >> 2) + jj) * kCtaRows * 16));
                f[4 * jj] += v.x;
                f[4 * jj + 1] += v.y;
                f[4 * jj + 2] += v.z;
                f[4 * jj + 3] += v.w;
              }
            }
            const int cc = c_own + cc0;
            if (n_ok)
              store_chunk16(outp + (int64_t)(tok0 + cc) * p.ldc + n, p.ldc, f, asc_u32 + 4 * cc, bsc, biasf, has_bias,
                            p.M - tok0 - cc);
          };
          for (; c < nc; c += 32 * kWpq) {
            tmem_ld_wait();
            const int c2 = c + 16 * kWpq;
            if (c2 < nc) tmem_ld_x16(taddr + c_own + c2, rb);
            reduce16(ra, c);
            if (c2 < nc) {
              tmem_ld_wait();
              if (c2 + 16 * kWpq < nc) tmem_ld_x16(taddr + c_own + c2 + 16 * kWpq, ra);
              reduce16(rb, c2);
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
  cluster_sync_relaxed();   // execution order only: no need to drain the output stores before arriving
  if (warp == kMmaWarp) {
    tc_fence_after();
    tmem_dealloc_2cta(tmem_base, kTmemCols);
  }
  if (threadIdx.x == 0) stamp(p, 10);
}

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

static unsigned long long* g_trace = nullptr;  // set through vb200_debug_scaled_mm_trace()

struct Plan {
  int tok_tile, m_tiles, n_ptiles, tiles, stages, stage_bytes, splits, kb_per_split, num_kb, col_chunk, clusters;
  size_t smem, main_bytes;
};

// Cost model (microseconds) of one (token tiles, K splits) choice.  The constants are FITTED to the exhaustive plan
// sweep (benchmarks/sweep_scaled_mm_plan.py -> profiles/r02_scaled_mm_plan_sweep_s52.jsonl: every placeable plan of the 8
// decode shapes x M in {16, 64, 128, 256}); with them the planner's pick is within 1.3 % of the best measured plan on
// average and within 12 % at worst.  What they stand for: an SM pulls ~100 GB/s through the L2 fabric in this kernel and
// all of them together ~9 TB/s, a pair retires a 256 x 256 x 128 k-block in ~0.15 us, HBM streams the weights at
// ~6.5 TB/s, the plain epilogue stores 256 token columns in ~1.5 us; a split adds two cluster barriers plus the
// teardown (1.6 us) and the partial columns pushed through distributed shared memory at 21 B/clk per SM.
static float plan_cost(int tok_tile, int m_tiles, int n_ptiles, int splits, int num_kb, int max_pairs, double w_bytes,
                       double a_bytes) {
  const int tiles = m_tiles * n_ptiles;
  const int kbps = ceil_div(num_kb, splits);
  const int rounds = ceil_div(tiles * splits, max_pairs);
  const float feed = kbps * (kWTileBytes + tok_tile / 2 * kBlockK) / 100e3f;
  const float mma = kbps * tok_tile * 0.0006f;
  const float hbm = (float)(w_bytes / 6.5e6);
  // bytes every SM pulls through the L2 fabric: the weights once per token tile, the activations once per row tile
  const float fabric = (float)((w_bytes * m_tiles + a_bytes * n_ptiles) / 9e6);
  const float main_us = fmaxf(fmaxf(rounds * fmaxf(feed, mma), hbm), fabric);
  float epi = 0.006f * tok_tile / splits;
  if (splits > 1) epi += 1.6f + 0.018f * tok_tile * (splits - 1) / splits;
  return main_us + epi;
}

static int max_clusters_cached(int cluster_ctas, size_t smem);

static Plan make_plan(int M, int N, int K) {
  Plan g;
  g.n_ptiles = ceil_div(N, kPairRows);
  g.num_kb = ceil_div(K, kBlockK);
  const int max_pairs = num_sms() / 2;
  const int tail = 2 * kMaxStages * 8 + 5 * 8 + 24 + 2 * kMaxTokTile * 4;  // barriers, tmem ptr, a_scale
  const int budget = 227 * 1024 - 1024 /*alignment slack*/ - tail;
  const double w_bytes = (double)N * K, a_bytes = (double)M * K;
  // candidates: token tiles of <= 256 columns (one more split of a single 129..256-token tile in two), K slices
  // 1..8 as long as a slice keeps >= 4 k-blocks and all clusters of the tile grid are resident at once
  const int mt_min = ceil_div(M, kMaxTokTile);
  float best = -1.f;
  g.m_tiles = mt_min;
  g.splits = 1;
  for (int mt = mt_min; mt <= mt_min + ((M > 128 && M <= kMaxTokTile) ? 1 : 0); ++mt) {
    const int tok = (ceil_div(M, mt) + 15) / 16 * 16;
    for (int s = 1; s <= kMaxSplits; ++s) {
      if (s > 1) {
        if (g.num_kb / s < 4) break;
        if (mt * g.n_ptiles * s > max_pairs) break;
        if (ceil_div(g.num_kb, ceil_div(g.num_kb, s)) != s) continue;   // this s does not give s non-empty slices
        if (max_clusters_cached(2 * s, (size_t)227 * 1024) < mt * g.n_ptiles) continue;   // cluster shape not placeable
      }
      const float c = plan_cost(tok, mt, g.n_ptiles, s, g.num_kb, max_pairs, w_bytes, a_bytes);
      if (best < 0.f || c < best - 1e-6f) {
        best = c;
        g.m_tiles = mt;
        g.splits = s;
      }
    }
  }
  g.tok_tile = (ceil_div(M, g.m_tiles) + 15) / 16 * 16;
  g.tiles = g.m_tiles * g.n_ptiles;
  const int a_tile = ((g.tok_tile / 2) * kBlockK + 1023) / 1024 * 1024;
  g.stage_bytes = kWTileBytes + a_tile;
  g.stages = budget / g.stage_bytes;
  if (g.stages > kMaxStages) g.stages = kMaxStages;
  g.kb_per_split = ceil_div(g.num_kb, g.splits);
  g.col_chunk = (ceil_div(g.tok_tile, g.splits) + 15) / 16 * 16;
  g.clusters = g.splits > 1 ? g.tiles : (g.tiles < max_pairs ? g.tiles : max_pairs);
  if (g.stages > g.kb_per_split && g.tiles <= g.clusters) g.stages = g.kb_per_split < 2 ? 2 : g.kb_per_split;
  size_t main_bytes = (size_t)g.stages * g.stage_bytes;
  const size_t region = (size_t)g.col_chunk * kCtaRows * 4;  // one source slot of the DSMEM receive buffer
  if (g.splits > 1 && (g.splits - 1) * region > main_bytes)
    main_bytes = ((g.splits - 1) * region + 1023) / 1024 * 1024;
  g.main_bytes = main_bytes;
  g.smem = 1024 + main_bytes + tail;
  return g;
}

template <typename OutT>
static cudaError_t prepare_kernel(int cluster_ctas) {
  auto kern = scaled_mm_kernel<OutT>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  if (e == cudaSuccess && cluster_ctas > 8)
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  return e;
}

// How many clusters of `cluster_ctas` CTAs (one CTA per SM: the kernel takes the whole shared memory) the device
// can hold at once; 0 when the shape cannot be placed at all.  Asked once per cluster size.
static int max_clusters_cached(int cluster_ctas, size_t smem) {
  static int cache[2 * kMaxSplits + 1];   // 0 = not asked yet, -1 = not placeable
  if (cluster_ctas <= 2) return num_sms() / 2;
  if (cache[cluster_ctas] == 0) {
    int n = 0;
    if (prepare_kernel<__nv_bfloat16>(cluster_ctas) == cudaSuccess) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(cluster_ctas * 64, 1, 1);
      cfg.blockDim = dim3(kThreads);
      cfg.dynamicSmemBytes = smem;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = cluster_ctas;
      attr[0].val.clusterDim.y = 1;
      attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      if (cudaOccupancyMaxActiveClusters(&n, scaled_mm_kernel<__nv_bfloat16>, &cfg) != cudaSuccess) n = 0;
    }
    (void)cudaGetLastError();
    cache[cluster_ctas] = n > 0 ? n : -1;
  }
  return cache[cluster_ctas] > 0 ? cache[cluster_ctas] : 0;
}

template <typename OutT>
static int launch(const Plan& g, const CUtensorMap& tw, const CUtensorMap& ta, const Params& p,
                  cudaStream_t stream) {
  auto kern = scaled_mm_kernel<OutT>;
  const int cluster_ctas = 2 * g.splits;
  VB_CUDA_CHECK(prepare_kernel<OutT>(cluster_ctas));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cluster_ctas * g.clusters, 1, 1);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = g.smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cluster_ctas;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  VB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tw, ta, p));
  return 0;
}

static int g_force_splits = 0;   // debug / tuning hook (vb200_debug_scaled_mm_plan): 0 = planner's choice
static int g_force_mtiles = 0;

static int run(void* out, const void* a, const void* b, const float* a_scales, const float* b_scales,
               const void* bias, int M, int N, int K, int64_t lda, int64_t ldb, int64_t ldc,
               int a_scales_numel, int b_scales_numel, int out_dtype, void* workspace, size_t workspace_bytes,
               cudaStream_t stream) {
  Plan g = make_plan(M, N, K);
  if (g_force_splits > 0 || g_force_mtiles > 0) {
    // re-derive the plan for a forced (token tiles, K slices) pair -- same arithmetic as make_plan's tail
    if (g_force_mtiles > 0) g.m_tiles = g_force_mtiles;
    if (g_force_splits > 0) g.splits = g_force_splits;
    VB_CHECK(ceil_div(M, g.m_tiles) <= kMaxTokTile, "scaled_mm debug plan: token tile too large");
    g.tok_tile = (ceil_div(M, g.m_tiles) + 15) / 16 * 16;
    g.tiles = g.m_tiles * g.n_ptiles;
    g.kb_per_split = ceil_div(g.num_kb, g.splits);
    VB_CHECK(ceil_div(g.num_kb, g.kb_per_split) == g.splits, "scaled_mm debug plan: empty K slice");
    const int tail = 2 * kMaxStages * 8 + 5 * 8 + 24 + 2 * kMaxTokTile * 4;
    const int a_tile = ((g.tok_tile / 2) * kBlockK + 1023) / 1024 * 1024;
    g.stage_bytes = kWTileBytes + a_tile;
    g.stages = (227 * 1024 - 1024 - tail) / g.stage_bytes;
    if (g.stages > kMaxStages) g.stages = kMaxStages;
    g.col_chunk = (ceil_div(g.tok_tile, g.splits) + 15) / 16 * 16;
    const int max_pairs = num_sms() / 2;
    g.clusters = g.splits > 1 ? g.tiles : (g.tiles < max_pairs ? g.tiles : max_pairs);
    size_t main_bytes = (size_t)g.stages * g.stage_bytes;
    const size_t region = (size_t)g.col_chunk * kCtaRows * 4;
    if (g.splits > 1 && (g.splits - 1) * region > main_bytes) main_bytes = ((g.splits - 1) * region + 1023) / 1024 * 1024;
    g.main_bytes = main_bytes;
    g.smem = 1024 + main_bytes + tail;
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
  p.main_bytes = (uint32_t)g.main_bytes;
  (void)workspace;
  (void)workspace_bytes;
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

// Debug / tuning hook (not part of the drop-in surface): force the number of K slices and / or token tiles of the
// following launches (0 = planner's choice); plan[0..5] (if non-NULL) receives the planner's choice for (M, N, K):
// token tile, token tiles, K slices, cluster CTAs, clusters, pipeline stages, stream-K (plan[0..6]).
extern "C" void vb200_debug_scaled_mm_plan(int force_splits, int force_mtiles, int M, int N, int K, int* plan) {
  if (force_splits < 0) force_splits = 0;
  v2::g_force_splits = force_splits;
  v2::g_force_mtiles = force_mtiles;
  if (plan != nullptr && M > 0 && N > 0 && K > 0) {
    const v2::Plan g = v2::make_plan(M, N, K);
    plan[0] = g.tok_tile;
    plan[1] = g.m_tiles;
    plan[2] = g.splits;
    plan[3] = 2 * g.splits;
    plan[4] = g.clusters;
    plan[5] = g.stages;
    plan[6] = 0;
  }
}

// Split-K reduces through distributed shared memory inside a cluster: the kernel needs no workspace (kept in the ABI).
extern "C" size_t vb200_cutlass_scaled_mm_workspace_size(int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  return 0;
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
