// Fused-MoE grouped FP8 GEMM for sm_100a: the work of invoke_fused_moe_kernel
// (/root/reference/vllm/model_executor/layers/fused_moe/fused_moe.py:478-647; Triton kernel :267-475)
// as a tcgen05 / TMEM / TMA kernel driven by moe_align_block_size's outputs.
//
// Decode-shaped design (same "swap-AB" mapping as scaled_mm.cu): an expert's WEIGHT rows are the
// UMMA M side (128 TMEM lanes per CTA), the tokens routed to that expert are the UMMA N side.  The
// N extent of every MMA is trimmed to the number of valid rows of the M-block (rounded up to 16),
// so the padding moe_align_block_size adds costs no tensor work, and a decode batch (tens of tokens
// per expert) reads every expert weight byte exactly once: the kernel is a pure HBM stream of the
// expert weights (1.4 GB / layer for Mixtral-8x7B at TP 1).
//   grid = (ceil(N/128), max M-blocks); CTA (n_tile, p): expert e = expert_ids[p]
//   warp 0: TMA producer  (W tile [128 x 128 B] of expert e, A tile [block_m x 128 B] of the
//           gathered activations, 128-byte swizzle, mbarrier complete_tx)
//   warp 1: tcgen05.mma issuer (one thread), accumulator in TMEM
//   warps 2-5: epilogue: tcgen05.ld -> (acc * routed_weight) * a_scale * b_scale -> out dtype ->
//           C[token_id, n]  (routed weight BEFORE the scales, fused_moe.py:455-468)
// Two CTAs are resident per SM (<= 113 KB of shared memory each), so one CTA's prologue/epilogue
// overlaps the other's weight stream.
// The activation rows of a block are not contiguous in A (row = token_id / top_k): a small gather
// kernel first writes them in sorted order into the workspace, so the GEMM loads A with plain TMA.
#include <cstdlib>

#include "tc_common.cuh"

namespace vb200 {
using namespace tc;

constexpr int kMoeBlockN = 128;   // weight rows per CTA (UMMA M)
constexpr int kMoeBlockK = 128;   // K bytes per stage
constexpr int kMoeUmmaK = 32;
constexpr int kMoeThreads = 192;
constexpr int kMoeMaxStages = 8;
constexpr int kMoeMaxBlockM = 256;
constexpr int kMoeWTile = kMoeBlockN * kMoeBlockK;

struct MoeParams {
  void* c;
  const float* a_scales;
  const float* b_scales;
  const float* topk_weights;  // nullptr: no routed-weight multiply
  const int32_t* sorted_token_ids;
  const int32_t* expert_ids;
  const int32_t* num_tokens_post_pad;
  int num_valid_tokens, top_k, E, N, K, block_m;
  int a_per_token, b_per_channel;
  int stages, num_kb, tmem_cols;
  int tile_cols;                   // TMEM columns of one weight tile's accumulator (tmem_cols = NT * tile_cols)
  // ragged ("CUTLASS-style") mode: experts own contiguous row ranges of A / C given by expert_offsets and
  // problem_sizes (csrc/quantization/cutlass_w8a8/moe/get_group_starts.cuh:11-31); no sorted_token_ids
  const int32_t* expert_offsets;   // [E] first row of expert e
  const int32_t* problem_sizes;    // [E, 3] = (m_e, N, K), or (N, m_e, K) when the producer swapped A/B
  int ragged;
  int in16;                        // 0: fp8 e4m3 operands; 1: f16; 2: bf16 (unquantised experts, test_moe.py:117-233).  K
                                   // and the tile sizes are in BYTES throughout, so only the MMA kind differs
  int m_fastest;                   // CTA launch order: 1 = all M-blocks of a weight-row tile back to back (see the kernel)
};

// A_sorted[r, :] = A[sorted_token_ids[r] / top_k, :] for valid r < num_tokens_post_pad (16-byte
// vectors; padding rows are left untouched: their accumulator columns are never written out).
__global__ void __launch_bounds__(256) moe_gather_rows_kernel(
    uint8_t* __restrict__ a_sorted, const uint8_t* __restrict__ a,
    const int32_t* __restrict__ sorted_token_ids, const int32_t* __restrict__ num_tokens_post_pad,
    int num_valid_tokens, int top_k, int K, int sorted_capacity) {
  pdl_wait();
  pdl_trigger();
  const int npost = min(*num_tokens_post_pad, sorted_capacity);
  const int vec_per_row = K >> 4;
  for (int r = blockIdx.x; r < npost; r += gridDim.x) {
    const int tok = sorted_token_ids[r];
    if (tok >= num_valid_tokens || tok < 0) continue;
    const uint4* src = reinterpret_cast<const uint4*>(a + (int64_t)(tok / top_k) * K);
    uint4* dst = reinterpret_cast<uint4*>(a_sorted + (int64_t)r * K);
    for (int i = threadIdx.x; i < vec_per_row; i += blockDim.x) dst[i] = src[i];
  }
}

// NT = weight-row tiles per CTA.  NT = 2 (larger batches): one CTA multiplies TWO adjacent 128-row weight tiles with ONE
// activation tile per k-block -- two accumulators in TMEM -- so the activation bytes, which every CTA pulls from L2
// for itself, are fetched once per 256 weight rows; one CTA per SM (the stage is 16 KiB larger and TMEM holds 2 x
// tile_cols columns).  NT = 1 (decode batches): the pure weight stream, two CTAs per SM.
template <typename OutT, int NT>
__global__ void __launch_bounds__(kMoeThreads, 2)
    moe_grouped_gemm_kernel(const __grid_constant__ CUtensorMap tmap_w,
                            const __grid_constant__ CUtensorMap tmap_a, const MoeParams p) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const int a_tile_bytes = p.block_m * kMoeBlockK;
  const int stage_bytes = NT * kMoeWTile + a_tile_bytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)p.stages * stage_bytes);
  uint64_t* empty_bar = full_bar + kMoeMaxStages;
  uint64_t* acc_bar = empty_bar + kMoeMaxStages;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(acc_bar + 1);
  int* s_nvalid = reinterpret_cast<int*>(tmem_ptr + 1);
  int* s_tok = reinterpret_cast<int*>(tmem_ptr + 6);          // [block_m] token ids of this block (16-byte aligned)
  float* s_rowscale = reinterpret_cast<float*>(s_tok + kMoeMaxBlockM);  // [block_m] routed_w, a_scale

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // Launch order.  CTAs start in the order x fastest.  With (n_tile, pblk) = (x, y) the M-blocks of one expert -- an
  // expert with more rows than block_m has several -- read the same weight tile a whole sweep of n tiles apart, and
  // the evict-first weight stream has left L2 by then: the expert's weights come from HBM once per M-block.  With
  // m_fastest the M-blocks of a weight-row tile run back to back, so the second read of a tile meets the first in
  // L2 and the weights cross HBM once whatever the routing skew.
  int n_tile = blockIdx.x, pblk = blockIdx.y;
  if (p.m_fastest) {
    const int lin = blockIdx.y * gridDim.x + blockIdx.x;
    n_tile = lin / (int)gridDim.y;
    pblk = lin - n_tile * (int)gridDim.y;
  }
  const int n0 = n_tile * kMoeBlockN * NT;
  const int nt_eff = (NT == 2 && n0 + kMoeBlockN < p.N) ? 2 : 1;   // the last CTA of an odd tile count has one tile
  int row0, expert;
  if (p.ragged) {
    // linear M-block id -> (expert, block inside the expert's row range): experts are walked in order, each owning
    // ceil(m_e / block_m) blocks; the walk is E <= 1024 L2-resident loads by one thread
    if (threadIdx.x == 0) {
      int e = 0, blk = pblk, m_e = 0;
      for (; e < p.E; ++e) {
        // one of the first two entries is N, the other m_e (moe_data.cu:30-44 writes either order)
        const int s0 = p.problem_sizes[e * 3], s1 = p.problem_sizes[e * 3 + 1];
        m_e = s1 == p.N ? s0 : s1;
        const int nb = (m_e + p.block_m - 1) / p.block_m;
        if (blk < nb) break;
        blk -= nb;
      }
      s_nvalid[1] = e < p.E ? e : -1;
      s_nvalid[2] = e < p.E ? p.expert_offsets[e] + blk * p.block_m : 0;
      s_nvalid[0] = e < p.E ? min(p.block_m, m_e - blk * p.block_m) : 0;
    }
    __syncthreads();
    expert = s_nvalid[1];
    row0 = s_nvalid[2];
    const int nv = s_nvalid[0];
    if (nv <= 0) return;  // uniform across the CTA
    for (int i = threadIdx.x; i < p.block_m; i += kMoeThreads) s_tok[i] = i < nv ? row0 + i : -1;
  } else {
    row0 = pblk * p.block_m;
    if (row0 >= *p.num_tokens_post_pad) return;  // uniform across the CTA
    expert = p.expert_ids[pblk];
    // ---- block metadata: token ids, valid count ---------------------------------------------------
    if (threadIdx.x == 0) *s_nvalid = 0;
    __syncthreads();
    int cnt = 0;
    for (int i = threadIdx.x; i < p.block_m; i += kMoeThreads) {
      const int tok = p.sorted_token_ids[row0 + i];
      s_tok[i] = tok;
      if (tok >= 0 && tok < p.num_valid_tokens) cnt = max(cnt, i + 1);
    }
    if (cnt) atomicMax(s_nvalid, cnt);
  }
  __syncthreads();
  const int nvalid = *s_nvalid;                    // valid rows are a prefix of the block
  if (nvalid == 0) return;
  const int umma_n = max(16, (nvalid + 15) & ~15);  // <= block_m (multiple of 16)

  if (expert < 0 || expert >= p.E) {
    // expert not on this rank (expert_map -> -1): write zeros (fused_moe.py:370-377)
    OutT* outp = reinterpret_cast<OutT*>(p.c);
    for (int i = 0; i < nvalid; ++i) {
      const int tok = s_tok[i];
      if (tok < 0 || tok >= p.num_valid_tokens) continue;
      for (int n = n0 + threadIdx.x; n < min(n0 + kMoeBlockN * NT, p.N); n += kMoeThreads)
        outp[(int64_t)tok * p.N + n] = from_float<OutT>(0.f);
    }
    return;
  }

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
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  // Producer and MMA issuer: the WHOLE warp runs the loop and waits on the barriers, one elected lane issues.  With the
  // warp converged the compiler keeps descriptors and barrier addresses in uniform registers; issued from a
  // `lane == 0` branch every UTCQMMA / UTCBAR / UTMALDG was wrapped in an ELECT + R2UR.BROADCAST waterfall loop
  // (measured on the dense GEMM: 69 cycles per MMA issue, profiles/r02_scaled_mm_trace_s42.txt vs _s46.txt).
  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    const int wrow = expert * p.N + n0;  // row of the [E*N, K] weight matrix
    for (int kb = 0; kb < p.num_kb; ++kb) {
      mbar_wait(&empty_bar[stage], phase ^ 1);
      uint8_t* sw = smem + (size_t)stage * stage_bytes;
      if (elect_one()) {
        mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)(nt_eff * kMoeWTile + a_tile_bytes));
        tma_load_2d(sw, &tmap_w, &full_bar[stage], kb * kMoeBlockK, wrow, kEvictFirst);
        if (NT == 2 && nt_eff == 2)
          tma_load_2d(sw + kMoeWTile, &tmap_w, &full_bar[stage], kb * kMoeBlockK, wrow + kMoeBlockN, kEvictFirst);
        tma_load_2d(sw + NT * kMoeWTile, &tmap_a, &full_bar[stage], kb * kMoeBlockK, row0, kEvictLast);
      }
      __syncwarp();
      if (++stage == p.stages) {
        stage = 0;
        phase ^= 1;
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = p.in16 ? make_idesc_16(kMoeBlockN, umma_n, p.in16 == 2) : make_idesc_f8(kMoeBlockN, umma_n);
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t smem_base_u = smem_u32(smem);
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < p.num_kb; ++kb) {
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      const uint32_t sw = smem_base_u + (uint32_t)stage * (uint32_t)stage_bytes;
      const uint64_t bdesc = make_smem_desc_sw128(sw + NT * kMoeWTile);
      if (elect_one()) {
#pragma unroll
        for (int t = 0; t < NT; ++t) {
          if (t < nt_eff) {
            const uint64_t adesc = make_smem_desc_sw128(sw + t * kMoeWTile);
            const uint32_t acc = tmem_u + (uint32_t)(t * p.tile_cols);
            // 32 bytes of K per MMA either way: 32 e4m3 or 16 f16 / bf16 elements
            if (p.in16) {
#pragma unroll
              for (int k = 0; k < kMoeBlockK / kMoeUmmaK; ++k)
                umma_f16(acc, adesc + (uint64_t)(k * (kMoeUmmaK >> 4)),
                         bdesc + (uint64_t)(k * (kMoeUmmaK >> 4)), idesc, (kb | k) != 0);
            } else {
#pragma unroll
              for (int k = 0; k < kMoeBlockK / kMoeUmmaK; ++k)
                umma_f8(acc, adesc + (uint64_t)(k * (kMoeUmmaK >> 4)),
                        bdesc + (uint64_t)(k * (kMoeUmmaK >> 4)), idesc, (kb | k) != 0);
            }
          }
        }
        umma_commit(&empty_bar[stage]);
      }
      __syncwarp();
      if (++stage == p.stages) {
        stage = 0;
        phase ^= 1;
      }
    }
    if (elect_one()) umma_commit(acc_bar);
    __syncwarp();
  } else {
    // ===== epilogue warps 2..5 =====
    const int q = warp & 3;
    const int ep_tid = threadIdx.x - 64;
    // per-row factors: routed weight and per-token activation scale, staged while the GEMM runs
    for (int i = ep_tid; i < umma_n; i += 128) {
      const int tok = s_tok[i];
      const bool ok = tok >= 0 && tok < p.num_valid_tokens;
      s_rowscale[i] = (ok && p.topk_weights) ? p.topk_weights[tok] : 1.f;
      s_rowscale[kMoeMaxBlockM + i] =
          p.a_scales == nullptr ? 1.f : (p.a_per_token ? (ok ? p.a_scales[tok / p.top_k] : 0.f) : p.a_scales[0]);
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    const int nl = q * 32 + lane;
    float bsc_t[NT];
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const int nn = n0 + t * kMoeBlockN + nl;
      bsc_t[t] = p.b_scales == nullptr ? 1.f
                                       : (p.b_per_channel ? (nn < p.N ? p.b_scales[(int64_t)expert * p.N + nn] : 0.f)
                                                          : p.b_scales[expert]);
    }
    const bool mul_w = p.topk_weights != nullptr;
    mbar_wait(acc_bar, 0);
    tc_fence_after();
    OutT* outp = reinterpret_cast<OutT*>(p.c);
#pragma unroll
    for (int t = 0; t < NT; ++t) {
    if (t >= nt_eff) break;
    const int n = n0 + t * kMoeBlockN + nl;
    const bool n_ok = n < p.N;
    const float bsc = bsc_t[t];
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(t * p.tile_cols);
    // token ids and the two per-row factors are read with explicit ld.shared, 16 columns at a time, BEFORE the
    // stores of those columns: a generic load would be ordered behind every earlier global store and serialise
    // the loop on the shared-memory latency
    const uint32_t tok_u32 = smem_u32(s_tok), rs_u32 = smem_u32(s_rowscale);
    for (int c = 0; c < umma_n; c += 16) {
      uint32_t r[16];
      tmem_ld_x16(taddr + c, r);
      int tk[16];
      float rw[16], as_[16];
#pragma unroll
      for (int q4 = 0; q4 < 4; ++q4) {
        uint32_t t0_, t1_, t2_, t3_;
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(t0_), "=r"(t1_), "=r"(t2_), "=r"(t3_)
                     : "r"(tok_u32 + 4 * (c + 4 * q4)));
        tk[4 * q4] = (int)t0_;
        tk[4 * q4 + 1] = (int)t1_;
        tk[4 * q4 + 2] = (int)t2_;
        tk[4 * q4 + 3] = (int)t3_;
        float4 v;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                     : "r"(rs_u32 + 4 * (c + 4 * q4)));
        rw[4 * q4] = v.x;
        rw[4 * q4 + 1] = v.y;
        rw[4 * q4 + 2] = v.z;
        rw[4 * q4 + 3] = v.w;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                     : "r"(rs_u32 + 4 * (kMoeMaxBlockM + c + 4 * q4)));
        as_[4 * q4] = v.x;
        as_[4 * q4 + 1] = v.y;
        as_[4 * q4 + 2] = v.z;
        as_[4 * q4 + 3] = v.w;
      }
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int i = c + j;
        const int tok = tk[j];
        if (i < nvalid && n_ok && tok >= 0 && tok < p.num_valid_tokens) {
          float v = __uint_as_float(r[j]);
          if (mul_w) v = v * rw[j];
          v = (v * as_[j]) * bsc;
          outp[(int64_t)tok * p.N + n] = from_float<OutT>(v);
        }
      }
    }
    }   // weight tile t
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
  }
}

}  // namespace vb200

using namespace vb200;

// stage count / TMEM columns / launch shared by the two entry points
static int launch_moe_gemm(MoeParams p, const CUtensorMap& tw, const CUtensorMap& ta, int m_blocks,
                           int out_dtype, cudaStream_t stream) {
  const int N = p.N, K = p.K, block_m = p.block_m;
  VB_CHECK(block_m >= 16 && block_m <= kMoeMaxBlockM && block_m % 16 == 0,
           "moe grouped GEMM: block_m (%d) must be a multiple of 16 in [16, %d]", block_m, kMoeMaxBlockM);
  p.num_kb = (K + kMoeBlockK - 1) / kMoeBlockK;
  {
    const char* e = getenv("VB200_MOE_ORDER");   // A/B switch for benchmarks: 0 = n tiles fastest (the old order)
    p.m_fastest = (e && e[0] == '0') ? 0 : 1;
  }
  // Two weight tiles per CTA (the kernel's NT comment) once an M-block holds 128 rows or more -- at those batches the
  // L2 -> SM traffic of the activation tiles binds -- provided the halved grid still fills the machine.  Long K (w2, K
  // = intermediate size): always (batch 512: 128 -> 105 us).  Short K (w13, 32 k-blocks): only with two CTAs per SM
  // (block_m <= 128, two-stage rings), whose overlapped prologues / epilogues one CTA per SM lacks (batch 512: 215 ->
  // 205 us, batch 2048: 606 -> 547; one CTA per SM: 223 / 635), and only when ceil(N / 256) x M-blocks >= 2 x #SMs
  // (TP-8 shard at batch 512: 168 CTAs, 39 -> 49 us otherwise).  profiles/r02_moe_gemm_nt2_s74.txt, _s81.txt
  int nt = 1;
  if (block_m >= 128 && N >= 2 * kMoeBlockN) {
    const long ctas2 = (long)((N + 2 * kMoeBlockN - 1) / (2 * kMoeBlockN)) * m_blocks;
    if (K >= 8192 || (block_m <= 128 && ctas2 >= 2L * num_sms())) nt = 2;
  }
  if (const char* e = getenv("VB200_MOE_NT")) {   // A/B switch for benchmarks
    if (e[0] == '1') nt = 1;
    if (e[0] == '2' && N >= 2 * kMoeBlockN) nt = 2;
  }
  const int stage_bytes = nt * kMoeWTile + block_m * kMoeBlockK;
  const int meta = 4096 + kMoeMaxBlockM * 4 * 3;
  // two CTAs per SM when the stages fit in half of the shared memory, else one
  int budget = (227 * 1024) / 2 - 1024 - meta;
  int stages = budget / stage_bytes;
  // NT = 2 with <= 128-row M-blocks: two CTAs per SM still fit (2 x 2 x 128 TMEM columns) with a two-stage ring -- 96 KB
  // in flight per CTA is what the L2 feed needs, the stage COUNT matters less (VB200_MOE_NT2_2CTA=0: one CTA per SM)
  const char* e2 = getenv("VB200_MOE_NT2_2CTA");
  const bool nt2_two_ctas = nt == 2 && block_m <= 128 && stages >= 2 && !(e2 && e2[0] == '0');
  if ((nt == 1 && stages < 3) || (nt == 2 && !nt2_two_ctas)) {
    budget = 227 * 1024 - 1024 - meta;
    stages = budget / stage_bytes;
  }
  if (stages > kMoeMaxStages) stages = kMoeMaxStages;
  if (stages > p.num_kb) stages = p.num_kb < 2 ? 2 : p.num_kb;
  p.stages = stages;
  int cols = 32;
  while (cols < block_m) cols <<= 1;
  p.tile_cols = cols;
  p.tmem_cols = nt * cols;
  size_t smem = 1024 + (size_t)stages * stage_bytes + meta;
  // NT = 2 is compiled for one CTA per SM and may take all 512 TMEM columns: never let two of them share an SM
  if (nt == 2 && !nt2_two_ctas && smem < 120 * 1024) smem = 120 * 1024;
  dim3 grid((N + kMoeBlockN * nt - 1) / (kMoeBlockN * nt), m_blocks);
#define VB_MOE_LAUNCH(T, NT_)                                                                             \
  do {                                                                                                    \
    auto kern = moe_grouped_gemm_kernel<T, NT_>;                                                          \
    VB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));   \
    VB_CUDA_CHECK(launch_k(kern, dim3(grid), dim3(kMoeThreads), smem, stream, tw, ta, p));                \
  } while (0)
  if (out_dtype == VB200_BF16) {
    if (nt == 2) VB_MOE_LAUNCH(__nv_bfloat16, 2);
    else VB_MOE_LAUNCH(__nv_bfloat16, 1);
  } else {
    if (nt == 2) VB_MOE_LAUNCH(__half, 2);
    else VB_MOE_LAUNCH(__half, 1);
  }
#undef VB_MOE_LAUNCH
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" size_t vb200_moe_grouped_gemm_workspace_size(int sorted_capacity, int K) {
  if (sorted_capacity <= 0 || K <= 0) return 0;
  // gathered activations + one extra M-block of slack so a TMA box never leaves the allocation
  return ((size_t)sorted_capacity + kMoeMaxBlockM) * (size_t)K + 256;
}

extern "C" int vb200_moe_grouped_gemm_fp8(
    void* c, const void* a, const void* b, const float* a_scales, const float* b_scales,
    const float* topk_weights, const int32_t* sorted_token_ids, const int32_t* expert_ids,
    const int32_t* num_tokens_post_pad, int num_valid_tokens, int num_a_rows, int top_k, int E,
    int N, int K, int block_m, int sorted_capacity, int a_scales_per_token,
    int b_scales_per_channel, int mul_routed_weight, int out_dtype, void* workspace,
    size_t workspace_bytes, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(out_dtype == VB200_BF16 || out_dtype == VB200_F16,
           "moe_grouped_gemm_fp8: out dtype must be bf16 or f16");
  VB_CHECK(K > 0 && K % 16 == 0, "moe_grouped_gemm_fp8: K (%d) must be a positive multiple of 16", K);
  VB_CHECK(block_m >= 16 && block_m <= kMoeMaxBlockM && block_m % 16 == 0,
           "moe_grouped_gemm_fp8: block_m (%d) must be a multiple of 16 in [16, 256]", block_m);
  VB_CHECK(top_k >= 1 && E >= 1 && N >= 1, "moe_grouped_gemm_fp8: bad top_k / E / N");
  VB_CHECK(!mul_routed_weight || topk_weights != nullptr,
           "moe_grouped_gemm_fp8: mul_routed_weight needs topk_weights");
  (void)num_a_rows;
  if (num_valid_tokens <= 0 || sorted_capacity <= 0) return 0;
  const size_t need = vb200_moe_grouped_gemm_workspace_size(sorted_capacity, K);
  VB_CHECK(workspace != nullptr && workspace_bytes >= need,
           "moe_grouped_gemm_fp8: workspace too small (%zu < %zu)", workspace_bytes, need);
  uint8_t* a_sorted = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(workspace) + 255) & ~static_cast<uintptr_t>(255));
  VB_CHECK((reinterpret_cast<uintptr_t>(a) & 15) == 0, "moe_grouped_gemm_fp8: A must be 16-byte aligned");
  {
    int grid = sorted_capacity < num_sms() * 8 ? sorted_capacity : num_sms() * 8;
    VB_CUDA_CHECK(launch_k(moe_gather_rows_kernel, dim3(grid), dim3(K >= 4096 ? 256 : 128), 0, stream, a_sorted, reinterpret_cast<const uint8_t*>(a), sorted_token_ids, num_tokens_post_pad,
        num_valid_tokens, top_k, K, sorted_capacity));
    VB_LAUNCH_CHECK();
  }
  const int m_blocks = (sorted_capacity + block_m - 1) / block_m;
  CUtensorMap tw, ta;
  if (tc::make_tmap_2d_u8(&tw, b, (uint64_t)E * N, (uint64_t)K, (uint64_t)K, kMoeBlockN)) return 1;
  if (tc::make_tmap_2d_u8(&ta, a_sorted, (uint64_t)m_blocks * block_m, (uint64_t)K, (uint64_t)K,
                          (uint32_t)block_m))
    return 1;
  MoeParams p;
  p.c = c;
  p.a_scales = a_scales;
  p.b_scales = b_scales;
  p.topk_weights = mul_routed_weight ? topk_weights : nullptr;
  p.sorted_token_ids = sorted_token_ids;
  p.expert_ids = expert_ids;
  p.num_tokens_post_pad = num_tokens_post_pad;
  p.num_valid_tokens = num_valid_tokens;
  p.top_k = top_k;
  p.E = E;
  p.N = N;
  p.K = K;
  p.block_m = block_m;
  p.a_per_token = a_scales_per_token ? 1 : 0;
  p.b_per_channel = b_scales_per_channel ? 1 : 0;
  p.expert_offsets = nullptr;
  p.problem_sizes = nullptr;
  p.ragged = 0;
  p.in16 = 0;
  return launch_moe_gemm(p, tw, ta, m_blocks, out_dtype, stream);
}

// The fp8 grouped GEMM fed with UNQUANTISED activations and their per-tensor scale: the rows are quantised while they
// are gathered into expert order (one launch instead of static_scaled_fp8_quant + gather, and no fp8 copy of the
// activations in token order).  Quantising a row is element-wise with one scale, so the bytes the GEMM reads are those
// of "quantise the tensor (fp8/common.cu:14-26), then gather": the result is bit-identical to
// vb200_static_scaled_fp8_quant + vb200_moe_grouped_gemm_fp8 (a_scales_per_token = 0).
namespace vb200 {
int launch_moe_gather_quant_sorted(uint8_t* out, const void* input, const int32_t* sorted_token_ids,
                                   const int32_t* num_tokens_post_pad, const float* static_scale, int sorted_capacity,
                                   int num_src_rows, int top_k, int hidden, int in_dtype, cudaStream_t stream);
}
extern "C" int vb200_moe_grouped_gemm_fp8_q16(
    void* c, const void* a, int a_dtype, const void* b, const float* a_scale, const float* b_scales,
    const float* topk_weights, const int32_t* sorted_token_ids, const int32_t* expert_ids,
    const int32_t* num_tokens_post_pad, int num_valid_tokens, int num_a_rows, int top_k, int E, int N, int K, int block_m,
    int sorted_capacity, int b_scales_per_channel, int mul_routed_weight, int out_dtype, void* workspace,
    size_t workspace_bytes, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(out_dtype == VB200_BF16 || out_dtype == VB200_F16, "moe_grouped_gemm_fp8_q16: out dtype must be bf16 or f16");
  VB_CHECK(a_dtype == VB200_BF16 || a_dtype == VB200_F16, "moe_grouped_gemm_fp8_q16: activations must be bf16 or f16");
  VB_CHECK(K > 0 && K % 16 == 0, "moe_grouped_gemm_fp8_q16: K (%d) must be a positive multiple of 16", K);
  VB_CHECK(block_m >= 16 && block_m <= kMoeMaxBlockM && block_m % 16 == 0,
           "moe_grouped_gemm_fp8_q16: block_m (%d) must be a multiple of 16 in [16, 256]", block_m);
  VB_CHECK(top_k >= 1 && E >= 1 && N >= 1 && num_a_rows >= 1, "moe_grouped_gemm_fp8_q16: bad top_k / E / N / rows");
  VB_CHECK(a_scale != nullptr, "moe_grouped_gemm_fp8_q16: the per-tensor activation scale is required");
  VB_CHECK(!mul_routed_weight || topk_weights != nullptr, "moe_grouped_gemm_fp8_q16: mul_routed_weight needs topk_weights");
  if (num_valid_tokens <= 0 || sorted_capacity <= 0) return 0;
  const size_t need = vb200_moe_grouped_gemm_workspace_size(sorted_capacity, K);
  VB_CHECK(workspace != nullptr && workspace_bytes >= need, "moe_grouped_gemm_fp8_q16: workspace too small (%zu < %zu)",
           workspace_bytes, need);
  uint8_t* a_sorted = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~static_cast<uintptr_t>(255));
  VB_CHECK((reinterpret_cast<uintptr_t>(a) & 15) == 0, "moe_grouped_gemm_fp8_q16: A must be 16-byte aligned");
  if (int rc = launch_moe_gather_quant_sorted(a_sorted, a, sorted_token_ids, num_tokens_post_pad, a_scale, sorted_capacity,
                                              num_a_rows, top_k, K, a_dtype, stream))
    return rc;
  const int m_blocks = (sorted_capacity + block_m - 1) / block_m;
  CUtensorMap tw, ta;
  if (tc::make_tmap_2d_u8(&tw, b, (uint64_t)E * N, (uint64_t)K, (uint64_t)K, kMoeBlockN)) return 1;
  if (tc::make_tmap_2d_u8(&ta, a_sorted, (uint64_t)m_blocks * block_m, (uint64_t)K, (uint64_t)K, (uint32_t)block_m)) return 1;
  MoeParams p;
  p.c = c;
  p.a_scales = a_scale;
  p.b_scales = b_scales;
  p.topk_weights = mul_routed_weight ? topk_weights : nullptr;
  p.sorted_token_ids = sorted_token_ids;
  p.expert_ids = expert_ids;
  p.num_tokens_post_pad = num_tokens_post_pad;
  p.num_valid_tokens = num_valid_tokens;
  p.top_k = top_k;
  p.E = E;
  p.N = N;
  p.K = K;
  p.block_m = block_m;
  p.a_per_token = 0;
  p.b_per_channel = b_scales_per_channel ? 1 : 0;
  p.expert_offsets = nullptr;
  p.problem_sizes = nullptr;
  p.ragged = 0;
  p.in16 = 0;
  return launch_moe_gemm(p, tw, ta, m_blocks, out_dtype, stream);
}

// Unquantised (f16 / bf16) experts: the bf16-weights variant of fused_experts (fused_moe.py:1349-1563 with
// use_fp8_w8a8 = False; reference test tests/kernels/moe/test_moe.py:117-233).  Same kernel -- tiles, TMA boxes and the
// pipeline are sized in bytes -- with kind::f16 MMAs and no scales:
//   C[token_id, n] = OutT( (sum_k A[token_id / top_k, k] * B[e, n, k]) * [topk_weights[token_id]] ),   fp32 accumulate.
// a [num_a_rows, K], b [E, N, K] of `dtype` (F16 / BF16), K % 8 == 0; c [num_valid_tokens, N] of the same dtype.
// workspace: vb200_moe_grouped_gemm_workspace_size(sorted_capacity, 2 * K) bytes.
extern "C" int vb200_moe_grouped_gemm_16(
    void* c, const void* a, const void* b, const float* topk_weights, const int32_t* sorted_token_ids,
    const int32_t* expert_ids, const int32_t* num_tokens_post_pad, int num_valid_tokens, int num_a_rows, int top_k, int E,
    int N, int K, int block_m, int sorted_capacity, int mul_routed_weight, int dtype, void* workspace,
    size_t workspace_bytes, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(dtype == VB200_BF16 || dtype == VB200_F16, "moe_grouped_gemm_16: dtype must be bf16 or f16");
  VB_CHECK(K > 0 && K % 8 == 0, "moe_grouped_gemm_16: K (%d) must be a positive multiple of 8", K);
  VB_CHECK(block_m >= 16 && block_m <= kMoeMaxBlockM && block_m % 16 == 0,
           "moe_grouped_gemm_16: block_m (%d) must be a multiple of 16 in [16, 256]", block_m);
  VB_CHECK(top_k >= 1 && E >= 1 && N >= 1, "moe_grouped_gemm_16: bad top_k / E / N");
  VB_CHECK(!mul_routed_weight || topk_weights != nullptr, "moe_grouped_gemm_16: mul_routed_weight needs topk_weights");
  (void)num_a_rows;
  if (num_valid_tokens <= 0 || sorted_capacity <= 0) return 0;
  const int kb = 2 * K;   // bytes per row
  const size_t need = vb200_moe_grouped_gemm_workspace_size(sorted_capacity, kb);
  VB_CHECK(workspace != nullptr && workspace_bytes >= need, "moe_grouped_gemm_16: workspace too small (%zu < %zu)",
           workspace_bytes, need);
  uint8_t* a_sorted = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~static_cast<uintptr_t>(255));
  VB_CHECK((reinterpret_cast<uintptr_t>(a) & 15) == 0 && (reinterpret_cast<uintptr_t>(b) & 15) == 0,
           "moe_grouped_gemm_16: A and B must be 16-byte aligned");
  {
    int grid = sorted_capacity < num_sms() * 8 ? sorted_capacity : num_sms() * 8;
    VB_CUDA_CHECK(launch_k(moe_gather_rows_kernel, dim3(grid), dim3(kb >= 4096 ? 256 : 128), 0, stream, a_sorted,
                           reinterpret_cast<const uint8_t*>(a), sorted_token_ids, num_tokens_post_pad, num_valid_tokens,
                           top_k, kb, sorted_capacity));
    VB_LAUNCH_CHECK();
  }
  const int m_blocks = (sorted_capacity + block_m - 1) / block_m;
  CUtensorMap tw, ta;
  if (tc::make_tmap_2d_u8(&tw, b, (uint64_t)E * N, (uint64_t)kb, (uint64_t)kb, kMoeBlockN)) return 1;
  if (tc::make_tmap_2d_u8(&ta, a_sorted, (uint64_t)m_blocks * block_m, (uint64_t)kb, (uint64_t)kb, (uint32_t)block_m)) return 1;
  MoeParams p;
  p.c = c;
  p.a_scales = nullptr;
  p.b_scales = nullptr;
  p.topk_weights = mul_routed_weight ? topk_weights : nullptr;
  p.sorted_token_ids = sorted_token_ids;
  p.expert_ids = expert_ids;
  p.num_tokens_post_pad = num_tokens_post_pad;
  p.num_valid_tokens = num_valid_tokens;
  p.top_k = top_k;
  p.E = E;
  p.N = N;
  p.K = kb;
  p.block_m = block_m;
  p.a_per_token = 0;
  p.b_per_channel = 0;
  p.expert_offsets = nullptr;
  p.problem_sizes = nullptr;
  p.ragged = 0;
  p.in16 = dtype == VB200_BF16 ? 2 : 1;
  return launch_moe_gemm(p, tw, ta, m_blocks, dtype, stream);
}


// cutlass_moe_mm: csrc/quantization/cutlass_w8a8/scaled_mm_entry.cu:242-270 (sm100:
// moe/grouped_mm_c3x_sm100.cu:64-140; pointer arrays moe/get_group_starts.cuh:11-31).  Expert e multiplies the
// rows [expert_offsets[e], +m_e) of A with its own weight matrix; here that is the same tcgen05 kernel as
// fused_experts' grouped GEMM in ragged mode: no gather, no pointer arrays, one launch.
extern "C" int vb200_cutlass_moe_mm(void* out, const void* a, const void* b, const float* a_scales,
                                    const float* b_scales, const int32_t* expert_offsets,
                                    const int32_t* problem_sizes, int num_rows, int E, int N, int K,
                                    int per_act_token, int per_out_ch, int out_dtype,
                                    vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(out_dtype == VB200_BF16 || out_dtype == VB200_F16,
           "cutlass_moe_mm: Invalid output type (must be float16 or bfloat16)");
  VB_CHECK(num_rows > 0, "cutlass_moe_mm: No input A tensors provided.");
  VB_CHECK(E > 0 && N > 0, "cutlass_moe_mm: No input B tensors provided.");
  VB_CHECK(K > 0 && K % 16 == 0, "cutlass_moe_mm: K (%d) must be a positive multiple of 16", K);
  VB_CHECK((reinterpret_cast<uintptr_t>(a) & 15) == 0 && (reinterpret_cast<uintptr_t>(b) & 15) == 0,
           "cutlass_moe_mm: A and B must be 16-byte aligned");
  // rows per M-block: about twice the mean rows per expert (skewed routing then still reads an expert's weights
  // once or twice), multiple of 16
  int block_m = (2 * ((num_rows + E - 1) / E) + 15) & ~15;
  if (block_m < 16) block_m = 16;
  // Short K (w13: K = hidden): <= 128 rows per block keeps two CTAs per SM, whose prologue / epilogue overlap is worth
  // more than reading an expert's weights once (batch 512: 272 -> 259 us); long K (w2) prefers the 256-row block
  // (113 vs 134 us) -- profiles/r01_moe_gemm_sweep_s16.jsonl, r01_moe_gemm_sweep_s31.jsonl
  if (block_m > 128 && K <= 8192) block_m = 128;
  // more than 256 rows per expert on average: every expert has several M-blocks whatever the block, and with the
  // M-blocks of a weight tile launched back to back the smaller block (two CTAs per SM) wins (w2, batch 2048: 269 vs
  // 286 us, profiles/r02_moe_gemm_block_m_s73.txt)
  if ((num_rows + E - 1) / E > 256) block_m = 128;
  if (const char* e = getenv("VB200_MOE_RAGGED_BLOCK_M")) {   // benchmark override
    const int v = atoi(e);
    if (v >= 16 && v % 16 == 0) block_m = v;
  }
  if (block_m > kMoeMaxBlockM) block_m = kMoeMaxBlockM;  // s_tok / s_rowscale / TMEM columns / UMMA N all end at 256
  const int m_blocks = E + num_rows / block_m;  // >= sum_e ceil(m_e / block_m)
  CUtensorMap tw, ta;
  if (tc::make_tmap_2d_u8(&tw, b, (uint64_t)E * N, (uint64_t)K, (uint64_t)K, kMoeBlockN)) return 1;
  // rows past num_rows are filled with zeros by the TMA unit (and never stored)
  if (tc::make_tmap_2d_u8(&ta, a, (uint64_t)num_rows, (uint64_t)K, (uint64_t)K, (uint32_t)block_m)) return 1;
  MoeParams p;
  p.c = out;
  p.a_scales = a_scales;
  p.b_scales = b_scales;
  p.topk_weights = nullptr;
  p.sorted_token_ids = nullptr;
  p.expert_ids = nullptr;
  p.num_tokens_post_pad = nullptr;
  p.num_valid_tokens = num_rows;
  p.top_k = 1;
  p.E = E;
  p.N = N;
  p.K = K;
  p.block_m = block_m;
  p.a_per_token = per_act_token ? 1 : 0;
  p.b_per_channel = per_out_ch ? 1 : 0;
  p.expert_offsets = expert_offsets;
  p.problem_sizes = problem_sizes;
  p.ragged = 1;
  p.in16 = 0;
  return launch_moe_gemm(p, tw, ta, m_blocks, out_dtype, stream);
}
