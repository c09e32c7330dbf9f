// MoE routing kernels for sm_100a: topk_softmax, moe_align_block_size (stable counting sort),
// moe_sum.  Semantics: /root/reference/csrc/moe/topk_softmax_kernels.cu:53-391,499-556 and
// /root/reference/csrc/moe/moe_align_sum_kernels.cu:17-123,204-318.  These are latency-bound at
// decode sizes (a few KB), so each is ONE launch (topk: one warp per token row; align: a single
// CTA that keeps histogram, offsets and per-warp ranks in shared memory; sum: one CTA per token).
#include <float.h>

#include "common.cuh"

namespace vb200 {

// ---------------------------------------------------------------------------------------------
// topk_softmax: one warp per row; lane l holds experts l, l+32, ...
// ---------------------------------------------------------------------------------------------
// InT: dtype of the router logits (the reference casts them to fp32 on the host first, fused_moe.py:908; reading
// bf16 / f16 here and widening is the same values).  RENORM: w /= sum(w) over the k winners, summed in k order
// (fused_moe.py:869-870 does it as two torch ops on the [T, k] tensor).
template <int VPT, typename IndT, typename InT = float, bool RENORM = false>
__global__ void __launch_bounds__(128) topk_softmax_kernel(
    float* __restrict__ topk_weights, IndT* __restrict__ topk_indices,
    int* __restrict__ token_expert_indices, const InT* __restrict__ gating, int num_tokens,
    int num_experts, int topk) {
  pdl_wait();
  pdl_trigger();
  const int row = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (row >= num_tokens) return;
  const int lane = threadIdx.x & 31;
  const InT* g = gating + (int64_t)row * num_experts;
  float v[VPT];
  float mx = -FLT_MAX;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int e = lane + 32 * i;
    v[i] = e < num_experts ? to_float<InT>(g[e]) : -FLT_MAX;
    mx = fmaxf(mx, v[i]);
  }
  mx = warp_max(mx);
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int e = lane + 32 * i;
    v[i] = e < num_experts ? expf(v[i] - mx) : 0.f;
    sum += v[i];
  }
  sum = warp_sum(sum);
  const float inv = 1.f / sum;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int e = lane + 32 * i;
    v[i] = e < num_experts ? v[i] * inv : -1.f;  // probabilities are >= 0; -1 never wins
  }
  float wsum = 0.f;   // lane 0: sum of the winners' probabilities
  // k rounds of arg-max, lowest index wins ties (topk_softmax_kernels.cu:336-358)
  for (int k = 0; k < topk; ++k) {
    float best = -2.f;
    int bidx = 0x7fffffff;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int e = lane + 32 * i;
      if (v[i] > best) {  // ascending e inside a lane: strict > keeps the lowest index
        best = v[i];
        bidx = e;
      }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, off);
      const int oi = __shfl_xor_sync(0xffffffffu, bidx, off);
      if (ob > best || (ob == best && oi < bidx)) {
        best = ob;
        bidx = oi;
      }
    }
    if (lane == 0) {
      const int64_t o = (int64_t)row * topk + k;
      topk_weights[o] = best;
      wsum += best;
      topk_indices[o] = (IndT)bidx;
      token_expert_indices[o] = k * num_tokens + row;  // :373
    }
    // mask the winner
    if ((bidx & 31) == lane) {
#pragma unroll
      for (int i = 0; i < VPT; ++i)
        if (lane + 32 * i == bidx) v[i] = -1.f;
    }
  }
  if (RENORM && lane == 0) {
    for (int k = 0; k < topk; ++k) {
      const int64_t o = (int64_t)row * topk + k;
      topk_weights[o] = topk_weights[o] / wsum;   // this lane wrote them: program order makes its own stores visible
    }
  }
}

// ---------------------------------------------------------------------------------------------
// moe_align_block_size: single CTA, stable order inside every expert.
//   smem: hist[E] | start[E+1] | run[E] | warp_cnt[NW][E]
// ---------------------------------------------------------------------------------------------
constexpr int kAlignThreads = 1024;
constexpr int kAlignWarps = kAlignThreads / 32;

__global__ void __launch_bounds__(kAlignThreads) moe_align_kernel(
    const int* __restrict__ topk_ids, int numel, int num_experts, int block_size,
    int* __restrict__ sorted_token_ids, int sorted_capacity, int* __restrict__ expert_ids,
    int expert_ids_capacity, int* __restrict__ num_tokens_post_pad) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ int sm[];
  int* hist = sm;                          // [E]
  int* start = hist + num_experts;         // [E + 1] padded exclusive prefix
  int* run = start + num_experts + 1;      // [E] tokens placed so far
  int* warp_cnt = run + num_experts;       // [NW][E]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  for (int i = tid; i < num_experts; i += kAlignThreads) {
    hist[i] = 0;
    run[i] = 0;
  }
  for (int i = tid; i < kAlignWarps * num_experts; i += kAlignThreads) warp_cnt[i] = 0;
  __syncthreads();
  // pass 1: histogram (ids outside [0, E) are ignored, like the reference's expert_map = -1 rows)
  for (int i = tid; i < numel; i += kAlignThreads) {
    const int e = topk_ids[i];
    if (e >= 0 && e < num_experts) atomicAdd(&hist[e], 1);
  }
  __syncthreads();
  // padded exclusive scan over experts (E <= 1024: one element per thread, warp-shuffle scan)
  {
    __shared__ int wsum[kAlignWarps];
    const int cnt = tid < num_experts ? (hist[tid] + block_size - 1) / block_size * block_size : 0;
    int incl = cnt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += t;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int w = wsum[lane];
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, w, off);
        if (lane >= off) w += t;
      }
      wsum[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const int base = warp > 0 ? wsum[warp - 1] : 0;
    if (tid < num_experts) start[tid] = base + incl - cnt;
    if (tid == num_experts - 1 || (num_experts > kAlignThreads && tid == kAlignThreads - 1))
      start[num_experts] = base + incl;
  }
  __syncthreads();
  const int total = start[num_experts];
  if (tid == 0) *num_tokens_post_pad = total;
  // expert id of every M-block; blocks past the padded total are 0 (reference: zero-initialised)
  for (int b = tid; b < expert_ids_capacity; b += kAlignThreads) {
    const int pos = b * block_size;
    int e = 0;
    if (pos < total) {
      // binary search: last expert with start[e] <= pos
      int lo = 0, hi = num_experts - 1;
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (start[mid] <= pos) lo = mid; else hi = mid - 1;
      }
      e = lo;
    }
    expert_ids[b] = e;
  }
  // padding sentinel everywhere first (entries are then overwritten by real tokens)
  for (int i = tid; i < sorted_capacity; i += kAlignThreads) sorted_token_ids[i] = numel;
  __syncthreads();
  // pass 2: stable scatter, 1024 tokens per tile in flat-index order
  for (int base = 0; base < numel; base += kAlignThreads) {
    const int i = base + tid;
    int e = -1;
    if (i < numel) {
      e = topk_ids[i];
      if (e < 0 || e >= num_experts) e = -1;
    }
    const unsigned peers = __match_any_sync(0xffffffffu, e);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    const bool leader = rank == 0 && e >= 0;
    if (leader) warp_cnt[warp * num_experts + e] = __popc(peers);
    __syncthreads();
    int pos = -1;
    if (e >= 0) {
      int before = 0;
      for (int w = 0; w < warp; ++w) before += warp_cnt[w * num_experts + e];
      pos = start[e] + run[e] + before + rank;
    }
    __syncthreads();
    if (leader) {
      atomicAdd(&run[e], __popc(peers));
      warp_cnt[warp * num_experts + e] = 0;
    }
    if (pos >= 0 && pos < sorted_capacity) sorted_token_ids[pos] = i;
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------
// moe_sum
// ---------------------------------------------------------------------------------------------
template <typename T, int TOPK>
__global__ void __launch_bounds__(256) moe_sum_kernel(T* __restrict__ out,
                                                      const T* __restrict__ in, int topk, int d, bool vec) {
  pdl_wait();
  pdl_trigger();
  const int64_t t = blockIdx.x;
  constexpr int VE = 16 / sizeof(T);
  if (TOPK > 0 && vec) {
    // 16-byte vectors: the kernel is a 3 MB stream at decode sizes and was latency-bound with scalar accesses
    // (11.7 us for [128, 2, 4096] bf16, profiles/r02_launches_cfg4_2steps_s56.txt).  Same arithmetic per element.
    for (int i = threadIdx.x; i < d / VE; i += blockDim.x) {
      union V { uint4 u; T e[VE]; } acc, v;
#pragma unroll
      for (int e = 0; e < VE; ++e) acc.e[e] = from_float<T>(0.f);
#pragma unroll
      for (int k = 0; k < TOPK; ++k) {
        v.u = ldg_stream_v4(in + (t * TOPK + k) * d + (int64_t)i * VE);
#pragma unroll
        for (int e = 0; e < VE; ++e) acc.e[e] = from_float<T>(to_float<T>(acc.e[e]) + to_float<T>(v.e[e]));
      }
      *reinterpret_cast<uint4*>(out + t * d + (int64_t)i * VE) = acc.u;
    }
    return;
  }
  for (int i = threadIdx.x; i < d; i += blockDim.x) {
    if (TOPK > 0) {
      // accumulate in T, in j order, starting from 0 (moe_align_sum_kernels.cu:116-121)
      T x = from_float<T>(0.f);
#pragma unroll
      for (int k = 0; k < TOPK; ++k)
        x = from_float<T>(to_float<T>(x) + to_float<T>(in[(t * TOPK + k) * d + i]));
      out[t * d + i] = x;
    } else {
      // other top-k: at::sum_out (fp32 accumulation, one rounding)
      float x = 0.f;
      for (int k = 0; k < topk; ++k) x += to_float<T>(in[(t * topk + k) * d + i]);
      out[t * d + i] = from_float<T>(x);
    }
  }
}

}  // namespace vb200

using namespace vb200;

template <typename IndT, typename InT = float, bool RENORM = false>
static int launch_topk(float* w, void* idx, int* tei, const InT* g, int T, int E, int k,
                       cudaStream_t s) {
  const int grid = (T + 3) / 4;
  const int vpt = (E + 31) / 32;
#define TOPK_CASE(V)                                                                      \
  if (vpt <= V) {                                                                         \
    VB_CUDA_CHECK(launch_k(topk_softmax_kernel<V, IndT, InT, RENORM>, dim3(grid), dim3(128), 0, s, w, (IndT*)idx, tei, g, T, E, k));    \
    VB_LAUNCH_CHECK();                                                                    \
    return 0;                                                                             \
  }
  TOPK_CASE(1) TOPK_CASE(2) TOPK_CASE(4) TOPK_CASE(8) TOPK_CASE(16) TOPK_CASE(32)
#undef TOPK_CASE
  VB_CHECK(false, "topk_softmax: num_experts %d > 1024 not supported", E);
  return 0;
}

extern "C" int vb200_topk_softmax(float* topk_weights, void* topk_indices,
                                  int32_t* token_expert_indices, const float* gating_output,
                                  int num_tokens, int num_experts, int topk, int indices_dtype,
                                  vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(topk >= 1 && topk <= num_experts, "topk %d must be in [1, num_experts %d]", topk,
           num_experts);
  if (num_tokens <= 0) return 0;
  switch (indices_dtype) {
    case VB200_I32:
      return launch_topk<int32_t>(topk_weights, topk_indices, token_expert_indices, gating_output,
                                  num_tokens, num_experts, topk, stream);
    case VB200_U32:
      return launch_topk<uint32_t>(topk_weights, topk_indices, token_expert_indices,
                                   gating_output, num_tokens, num_experts, topk, stream);
    case VB200_I64:
      return launch_topk<int64_t>(topk_weights, topk_indices, token_expert_indices, gating_output,
                                  num_tokens, num_experts, topk, stream);
    default:
      VB_CHECK(false, "topk_softmax: indices must be int32, uint32 or int64");
  }
  return 0;
}

// fused_topk (fused_moe.py:882-915) as ONE launch: softmax over the router logits in their own dtype, top-k, optional
// renormalisation.  int32 indices.
template <typename InT>
static int fused_topk_in(float* w, int32_t* idx, int32_t* tei, const void* g, int T, int E, int k, int renorm, cudaStream_t s) {
  if (renorm) return launch_topk<int32_t, InT, true>(w, idx, tei, (const InT*)g, T, E, k, s);
  return launch_topk<int32_t, InT, false>(w, idx, tei, (const InT*)g, T, E, k, s);
}

extern "C" int vb200_fused_topk(float* topk_weights, int32_t* topk_indices, int32_t* token_expert_indices,
                                const void* gating_output, int num_tokens, int num_experts, int topk, int renormalize,
                                int in_dtype, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(topk >= 1 && topk <= num_experts, "fused_topk: topk %d must be in [1, num_experts %d]", topk, num_experts);
  if (num_tokens <= 0) return 0;
  switch (in_dtype) {
    case VB200_F32: return fused_topk_in<float>(topk_weights, topk_indices, token_expert_indices, gating_output, num_tokens, num_experts, topk, renormalize, stream);
    case VB200_F16: return fused_topk_in<__half>(topk_weights, topk_indices, token_expert_indices, gating_output, num_tokens, num_experts, topk, renormalize, stream);
    case VB200_BF16: return fused_topk_in<__nv_bfloat16>(topk_weights, topk_indices, token_expert_indices, gating_output, num_tokens, num_experts, topk, renormalize, stream);
    default: VB_CHECK(false, "fused_topk: router logits must be fp32, f16 or bf16");
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Router in one launch: logits = hidden @ w_gate^T (the replicated, unquantised gate Linear: FusedMoE's
// `router_logits, _ = self.gate(hidden_states)`, models/mixtral.py:105-112), rounded to the activation dtype as the
// library GEMM's output would be, then fused_topk (softmax, top-k, renormalisation: fused_moe.py:866-915) -- and, on
// the way, max|hidden| / 448 for the dynamic per-tensor fp8 scale of the experts' first GEMM, since the row is in
// registers anyway.  One CTA of 4 warps per token; E <= 32 (lane e of warp 0 owns expert e); H <= 8 * 8 * 128.
// ---------------------------------------------------------------------------------------------
namespace vb200 {
constexpr int kRouterThreads = 128;
constexpr int kRouterMaxVec = 8;   // 16-byte vectors of the row per thread

template <typename T, bool RENORM>
__global__ void __launch_bounds__(kRouterThreads) moe_router_topk_kernel(
    float* __restrict__ topk_weights, int* __restrict__ topk_indices, int* __restrict__ token_expert_indices,
    T* __restrict__ logits_out, float* __restrict__ absmax_scale, const T* __restrict__ hidden,
    const T* __restrict__ w_gate, int num_tokens, int H, int E, int topk) {
  pdl_wait();
  pdl_trigger();
  __shared__ float part[kRouterThreads / 32][32];
  __shared__ float smax[kRouterThreads / 32];
  const int row = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nvec = H >> 3;
  const T* x = hidden + (int64_t)row * H;
  union V8 { uint4 u; T e[8]; };
  V8 xv[kRouterMaxVec];
  float m = 0.f;
#pragma unroll
  for (int i = 0; i < kRouterMaxVec; ++i) {
    const int v = tid + i * kRouterThreads;
    if (v < nvec) {
      xv[i].u = *reinterpret_cast<const uint4*>(x + (int64_t)v * 8);
#pragma unroll
      for (int j = 0; j < 8; ++j) m = fmaxf(m, fabsf(to_float<T>(xv[i].e[j])));
    }
  }
  for (int e = 0; e < E; ++e) {
    const T* w = w_gate + (int64_t)e * H;
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < kRouterMaxVec; ++i) {
      const int v = tid + i * kRouterThreads;
      if (v < nvec) {
        V8 wv;
        wv.u = *reinterpret_cast<const uint4*>(w + (int64_t)v * 8);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc = fmaf(to_float<T>(xv[i].e[j]), to_float<T>(wv.e[j]), acc);
      }
    }
    acc = warp_sum(acc);
    if (lane == 0) part[warp][e] = acc;
  }
  m = warp_max(m);
  if (lane == 0) smax[warp] = m;
  __syncthreads();
  if (warp != 0) return;
  if (absmax_scale != nullptr && lane == 0) {
    float mm = smax[0];
#pragma unroll
    for (int w = 1; w < kRouterThreads / 32; ++w) mm = fmaxf(mm, smax[w]);
    atomicMax(reinterpret_cast<int*>(absmax_scale), __float_as_int(mm / 448.0f));   // as absmax_scale_kernel
  }
  // lane e: logit of expert e, rounded to T like a GEMM output, then fused_topk on it
  float v = -FLT_MAX;
  if (lane < E) {
    float acc = part[0][lane];
#pragma unroll
    for (int w = 1; w < kRouterThreads / 32; ++w) acc += part[w][lane];
    const T r = from_float<T>(acc);
    if (logits_out != nullptr) logits_out[(int64_t)row * E + lane] = r;
    v = to_float<T>(r);
  }
  const float mx = warp_max(v);
  float p = lane < E ? expf(v - mx) : 0.f;
  const float sum = warp_sum(p);
  const float inv = 1.f / sum;
  p = lane < E ? p * inv : -1.f;
  float wsum = 0.f;
  for (int k = 0; k < topk; ++k) {
    float best = p;
    int bidx = lane;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, off);
      const int oi = __shfl_xor_sync(0xffffffffu, bidx, off);
      if (ob > best || (ob == best && oi < bidx)) {
        best = ob;
        bidx = oi;
      }
    }
    if (lane == 0) {
      const int64_t o = (int64_t)row * topk + k;
      topk_weights[o] = best;
      wsum += best;
      topk_indices[o] = bidx;
      if (token_expert_indices != nullptr) token_expert_indices[o] = k * num_tokens + row;
    }
    if (lane == bidx) p = -1.f;
  }
  if (RENORM && lane == 0) {
    for (int k = 0; k < topk; ++k) {
      const int64_t o = (int64_t)row * topk + k;
      topk_weights[o] = topk_weights[o] / wsum;
    }
  }
}
}  // namespace vb200

extern "C" int vb200_moe_router_topk(float* topk_weights, int32_t* topk_indices, int32_t* token_expert_indices,
                                     void* logits_out, float* absmax_scale, const void* hidden, const void* w_gate,
                                     int num_tokens, int hidden_size, int num_experts, int topk, int renormalize,
                                     int dtype, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(dtype == VB200_BF16 || dtype == VB200_F16, "moe_router_topk: hidden / w_gate must be bf16 or f16");
  VB_CHECK(num_experts >= 1 && num_experts <= 32, "moe_router_topk: num_experts %d must be in [1, 32]", num_experts);
  VB_CHECK(topk >= 1 && topk <= num_experts, "moe_router_topk: topk %d must be in [1, num_experts %d]", topk, num_experts);
  VB_CHECK(hidden_size % 8 == 0 && hidden_size <= 8 * kRouterMaxVec * kRouterThreads,
           "moe_router_topk: hidden size %d must be a multiple of 8 and at most %d", hidden_size,
           8 * kRouterMaxVec * kRouterThreads);
  VB_CHECK(((reinterpret_cast<uintptr_t>(hidden) | reinterpret_cast<uintptr_t>(w_gate)) & 15) == 0,
           "moe_router_topk: hidden and w_gate must be 16-byte aligned");
  if (num_tokens <= 0) return 0;
#define VB_ROUTER(T, R)                                                                                            \
  VB_CUDA_CHECK(launch_k(moe_router_topk_kernel<T, R>, dim3(num_tokens), dim3(kRouterThreads), 0, stream, topk_weights, \
                         topk_indices, token_expert_indices, (T*)logits_out, absmax_scale, (const T*)hidden,       \
                         (const T*)w_gate, num_tokens, hidden_size, num_experts, topk))
  if (dtype == VB200_BF16) {
    if (renormalize) VB_ROUTER(__nv_bfloat16, true);
    else VB_ROUTER(__nv_bfloat16, false);
  } else {
    if (renormalize) VB_ROUTER(__half, true);
    else VB_ROUTER(__half, false);
  }
#undef VB_ROUTER
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" size_t vb200_moe_align_workspace_size(int numel, int num_experts) {
  (void)numel;
  (void)num_experts;
  return 0;  // everything lives in shared memory
}

extern "C" int vb200_moe_align_block_size(const int32_t* topk_ids, int numel, int num_experts,
                                          int block_size, int32_t* sorted_token_ids,
                                          int sorted_capacity, int32_t* expert_ids,
                                          int expert_ids_capacity, int32_t* num_tokens_post_pad,
                                          void* workspace, size_t workspace_bytes,
                                          vb200_stream_t stream_) {
  (void)workspace;
  (void)workspace_bytes;
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(num_experts >= 1 && num_experts <= 1024,
           "moe_align_block_size: num_experts %d must be in [1, 1024]", num_experts);
  VB_CHECK(block_size >= 1, "moe_align_block_size: bad block_size %d", block_size);
  const size_t smem = ((size_t)num_experts * 3 + 1 + (size_t)kAlignWarps * num_experts) * 4;
  if (smem > 48 * 1024) {
    VB_CUDA_CHECK(cudaFuncSetAttribute(moe_align_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  VB_CUDA_CHECK(launch_k(moe_align_kernel, dim3(1), dim3(kAlignThreads), smem, stream, topk_ids, numel, num_experts, block_size,
                                                       sorted_token_ids, sorted_capacity,
                                                       expert_ids, expert_ids_capacity,
                                                       num_tokens_post_pad));
  VB_LAUNCH_CHECK();
  return 0;
}

#define DISPATCH_T(dt, ...)                                                 \
  switch (dt) {                                                             \
    case VB200_F32: { using scalar_t = float; __VA_ARGS__; break; }         \
    case VB200_F16: { using scalar_t = __half; __VA_ARGS__; break; }        \
    case VB200_BF16: { using scalar_t = __nv_bfloat16; __VA_ARGS__; break; } \
    default: VB_CHECK(false, "unsupported dtype %d", dt);                   \
  }

extern "C" int vb200_moe_sum(void* out, const void* input, int num_tokens, int topk, int hidden,
                             int dtype, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (num_tokens <= 0 || hidden <= 0) return 0;
  const int esz = dtype == VB200_F32 ? 4 : 2;
  const bool vec = hidden % (16 / esz) == 0 && (reinterpret_cast<uintptr_t>(input) & 15) == 0 &&
                   (reinterpret_cast<uintptr_t>(out) & 15) == 0;
  const int work = vec ? hidden / (16 / esz) : hidden;
  const int threads = work >= 256 ? 256 : (work >= 128 ? 128 : 64);
  DISPATCH_T(dtype, {
    scalar_t* o = (scalar_t*)out;
    const scalar_t* in = (const scalar_t*)input;
    switch (topk) {
      case 2: VB_CUDA_CHECK(launch_k(moe_sum_kernel<scalar_t, 2>, dim3(num_tokens), dim3(threads), 0, stream, o, in, topk, hidden, vec)); break;
      case 3: VB_CUDA_CHECK(launch_k(moe_sum_kernel<scalar_t, 3>, dim3(num_tokens), dim3(threads), 0, stream, o, in, topk, hidden, vec)); break;
      case 4: VB_CUDA_CHECK(launch_k(moe_sum_kernel<scalar_t, 4>, dim3(num_tokens), dim3(threads), 0, stream, o, in, topk, hidden, vec)); break;
      default: VB_CUDA_CHECK(launch_k(moe_sum_kernel<scalar_t, 0>, dim3(num_tokens), dim3(threads), 0, stream, o, in, topk, hidden, vec)); break;
    }
  });
  VB_LAUNCH_CHECK();
  return 0;
}
