// Routing metadata and row permutation of the CUTLASS-style fused-MoE API:
//   get_cutlass_moe_mm_data      /root/reference/csrc/quantization/cutlass_w8a8/moe/moe_data.cu:11-159
//   get_cutlass_pplx_moe_mm_data /root/reference/csrc/quantization/cutlass_w8a8/moe/moe_data.cu:161-195
//   shuffle_rows                 /root/reference/vllm/_custom_ops.py:912-922 (torch.ops._moe_C.shuffle_rows)
// The reference runs five small kernels (zero-fill, per-expert count with atomics, serial offsets, per-expert
// arg-sort with atomics -> order inside an expert depends on thread scheduling).  Here ONE CTA does a stable
// counting sort in shared memory: same expert_offsets / problem sizes, and the permutation is the deterministic
// member (ascending flat index inside every expert) of the set of permutations the reference may return.
#include "common.cuh"

namespace vb200 {

constexpr int kDataThreads = 1024;
constexpr int kDataWarps = kDataThreads / 32;

// smem: hist[E] | start[E+1] | run[E] | warp_cnt[NW][E]
__global__ void __launch_bounds__(kDataThreads) moe_mm_data_kernel(
    const int* __restrict__ topk_ids, int numel, int topk, int num_experts, int n, int k, int swap_ab,
    int* __restrict__ expert_offsets, int* __restrict__ problem_sizes1, int* __restrict__ problem_sizes2,
    int* __restrict__ input_permutation, int* __restrict__ output_permutation,
    int* __restrict__ blockscale_offsets) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ int sm[];
  int* hist = sm;
  int* start = hist + num_experts;
  int* run = start + num_experts + 1;
  int* warp_cnt = run + num_experts;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < num_experts; i += kDataThreads) {
    hist[i] = 0;
    run[i] = 0;
  }
  for (int i = tid; i < kDataWarps * num_experts; i += kDataThreads) warp_cnt[i] = 0;
  __syncthreads();
  for (int i = tid; i < numel; i += kDataThreads) {
    const int e = topk_ids[i];
    if (e >= 0 && e < num_experts) atomicAdd(&hist[e], 1);
  }
  __syncthreads();
  // problem sizes (moe_data.cu:28-44): (m_e, 2n, k) / (m_e, k, n), first two entries swapped for swap_ab
  for (int e = tid; e < num_experts; e += kDataThreads) {
    const int m = hist[e];
    problem_sizes1[e * 3 + (swap_ab ? 1 : 0)] = m;
    problem_sizes1[e * 3 + (swap_ab ? 0 : 1)] = 2 * n;
    problem_sizes1[e * 3 + 2] = k;
    problem_sizes2[e * 3 + (swap_ab ? 1 : 0)] = m;
    problem_sizes2[e * 3 + (swap_ab ? 0 : 1)] = k;
    problem_sizes2[e * 3 + 2] = n;
  }
  // exclusive prefix over experts (serial over <= 1024 shared-memory words: moe_data.cu:48-59)
  if (tid == 0) {
    int tot = 0, tot_round = 0;
    if (blockscale_offsets) blockscale_offsets[0] = 0;
    for (int e = 0; e < num_experts; ++e) {
      start[e] = tot;
      expert_offsets[e] = tot;
      tot += hist[e];
      if (blockscale_offsets) {
        tot_round += (hist[e] + 127) / 128 * 128;   // moe_data.cu:75
        blockscale_offsets[e + 1] = tot_round;
      }
    }
    start[num_experts] = tot;
    expert_offsets[num_experts] = tot;
  }
  __syncthreads();
  const int total = start[num_experts];
  // stable scatter, 1024 flat indices per pass
  for (int base = 0; base < numel; base += kDataThreads) {
    const int i = base + tid;
    int e = -1;
    bool invalid = false;
    if (i < numel) {
      e = topk_ids[i];
      invalid = e == -1;
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
    if (pos >= 0) {
      input_permutation[pos] = i / topk;
      output_permutation[i] = pos;
    } else if (invalid) {
      output_permutation[i] = total;  // moe_data.cu:92-99: row `total` of the zero-padded output
    }
    __syncthreads();
  }
}

__global__ void moe_pplx_data_kernel(int* __restrict__ expert_offsets, int* __restrict__ problem_sizes1,
                                     int* __restrict__ problem_sizes2,
                                     const int* __restrict__ expert_num_tokens, int num_local_experts,
                                     int padded_m, int n, int k) {
  pdl_wait();
  pdl_trigger();
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= num_local_experts) return;
  expert_offsets[e] = e * padded_m;
  const int m = expert_num_tokens[e];
  problem_sizes1[e * 3] = m;
  problem_sizes1[e * 3 + 1] = 2 * n;
  problem_sizes1[e * 3 + 2] = k;
  problem_sizes2[e * 3] = m;
  problem_sizes2[e * 3 + 1] = k;
  problem_sizes2[e * 3 + 2] = n;
}

// out[r, :] = in[dst2src[r], :], rows of row_bytes bytes (16-byte vectors when aligned, bytes otherwise)
template <typename V>
__global__ void __launch_bounds__(256) shuffle_rows_kernel(uint8_t* __restrict__ out,
                                                           const uint8_t* __restrict__ in,
                                                           const int* __restrict__ dst2src, int num_dst,
                                                           int num_src, int64_t row_bytes) {
  pdl_wait();
  pdl_trigger();
  const int64_t nvec = row_bytes / (int64_t)sizeof(V);
  for (int r = blockIdx.x; r < num_dst; r += gridDim.x) {
    const int s = dst2src[r];
    if (s < 0 || s >= num_src) continue;
    const V* src = reinterpret_cast<const V*>(in + (int64_t)s * row_bytes);
    V* dst = reinterpret_cast<V*>(out + (int64_t)r * row_bytes);
    for (int64_t i = threadIdx.x; i < nvec; i += blockDim.x) dst[i] = src[i];
  }
}

}  // namespace vb200

using namespace vb200;

extern "C" int vb200_get_cutlass_moe_mm_data(const int32_t* topk_ids, int num_tokens, int topk,
                                             int32_t* expert_offsets, int32_t* problem_sizes1,
                                             int32_t* problem_sizes2, int32_t* input_permutation,
                                             int32_t* output_permutation, int num_experts, int n, int k,
                                             int32_t* blockscale_offsets, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(num_experts >= 1 && num_experts <= 1024,
           "get_cutlass_moe_mm_data: num_experts %d must be in [1, 1024]", num_experts);
  VB_CHECK(topk >= 1 && num_tokens >= 0, "get_cutlass_moe_mm_data: bad topk_ids shape");
  const int numel = num_tokens * topk;
  // moe_data.cu:9,121-122: problem sizes are written M/N-swapped for <= 64 routed rows (never for the fp4 path)
  const int swap_ab = (blockscale_offsets == nullptr && numel <= 64) ? 1 : 0;
  const size_t smem = ((size_t)num_experts * 3 + 1 + (size_t)kDataWarps * num_experts) * 4;
  if (smem > 48 * 1024)
    VB_CUDA_CHECK(cudaFuncSetAttribute(moe_mm_data_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
  VB_CUDA_CHECK(launch_k(moe_mm_data_kernel, dim3(1), dim3(kDataThreads), smem, stream, topk_ids, numel, topk, num_experts, n, k, swap_ab,
                                                        expert_offsets, problem_sizes1, problem_sizes2,
                                                        input_permutation, output_permutation,
                                                        blockscale_offsets));
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_get_cutlass_pplx_moe_mm_data(int32_t* expert_offsets, int32_t* problem_sizes1,
                                                  int32_t* problem_sizes2,
                                                  const int32_t* expert_num_tokens, int num_local_experts,
                                                  int padded_m, int n, int k, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(num_local_experts >= 1, "get_cutlass_pplx_moe_mm_data: no experts");
  VB_CUDA_CHECK(launch_k(moe_pplx_data_kernel, dim3((num_local_experts + 127) / 128), dim3(128), 0, stream, expert_offsets, problem_sizes1, problem_sizes2, expert_num_tokens, num_local_experts, padded_m, n, k));
  VB_LAUNCH_CHECK();
  return 0;
}

extern "C" int vb200_shuffle_rows(void* out, const void* in, const int32_t* dst2src_map, int num_dst_rows,
                                  int num_src_rows, int64_t row_bytes, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(row_bytes > 0, "shuffle_rows: empty rows");
  if (num_dst_rows <= 0) return 0;
  const int grid = num_dst_rows < num_sms() * 8 ? num_dst_rows : num_sms() * 8;
  const bool vec = row_bytes % 16 == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
                   (reinterpret_cast<uintptr_t>(in) & 15) == 0;
  uint8_t* o = reinterpret_cast<uint8_t*>(out);
  const uint8_t* i = reinterpret_cast<const uint8_t*>(in);
  if (vec)
    VB_CUDA_CHECK(launch_k(shuffle_rows_kernel<uint4>, dim3(grid), dim3(row_bytes >= 4096 ? 256 : 128), 0, stream, o, i, dst2src_map, num_dst_rows, num_src_rows, row_bytes));
  else
    VB_CUDA_CHECK(launch_k(shuffle_rows_kernel<uint8_t>, dim3(grid), dim3(256), 0, stream, o, i, dst2src_map, num_dst_rows, num_src_rows,
                                                           row_bytes));
  VB_LAUNCH_CHECK();
  return 0;
}
