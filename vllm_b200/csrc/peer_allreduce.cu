// Row-parallel all-reduce over NVLink peer memory, fused with its consumer.
//
// The reference's analogue is the P2P custom all-reduce (/root/reference/csrc/custom_all_reduce.cuh:297-314,
// 560-580: one-shot `cross_device_reduce_1stage`, flag barriers :184-260) plus the compilation pass that fuses
// all-reduce + residual-add + RMSNorm (/root/reference/vllm/compilation/collective_fusion.py).  Here that is ONE
// kernel: every rank maps every peer's partial-sum buffer (CUDA IPC) and, after a flag barrier at system scope,
// each CTA pulls its token rows from all peers with 16-byte loads over NVLink, sums them in rank order in fp32
// (the same order on every rank => bitwise identical results everywhere), rounds to the activation dtype (the value
// NCCL's all-reduce would have delivered for 2 ranks), and goes straight on with the residual add, the RMSNorm and
// -- for the FP8 linear that follows -- the dynamic per-token quantisation, using the same arithmetic (and rounding
// points) as rms_norm_kernel in layer_glue.cu.  One launch replaces all-reduce + fused_add_rms_norm (+ quant).
//   MODE 0: out = T(sum_r x_r)                                                    (flat, any numel % 8 == 0)
//   MODE 1: x = T(sum); z = T(x + res); res = z; out = T(T(z * inv) * w)          (fused_add_rms_norm)
//   MODE 2: x = T(sum); z = x + res (fp32); res = T(z); y = T(T(z * inv) * w); s = max(amax|y| / 448, 1/(448*512));
//           out = fp8(clamp(y / s)); scales[row] = s                              (rms_norm_dynamic_per_token_quant)
// Barrier: per CTA index, a monotonically increasing counter kept in device memory (graph-replay safe); CTA b of
// rank r stores the counter into slot [b][r] of every peer's signal block (st.release.sys after a system fence)
// and spins (ld.acquire.sys) until all peers' counters show up in its own block.  A second barrier at the end keeps
// a rank from overwriting its buffer (next GEMM) while a peer is still reading it.
#include <cstring>

#include "common.cuh"

namespace vb200 {

constexpr int kArMaxRanks = 8;
constexpr int kArMaxBlocks = 320;
constexpr int kArThreads = 256;
constexpr int kArMaxVec = 8;   // rows up to 256 * 8 * 8 = 16384 elements stay in registers
constexpr float kFp8MaxA = 448.f;

struct ArSignal {
  uint32_t self_counter[kArMaxBlocks][kArMaxRanks];
  uint32_t peer_counter[2][kArMaxBlocks][kArMaxRanks];
};

struct ArParams {
  const void* data[kArMaxRanks];   // peer partial sums (this process's mapping), [rows, hidden] contiguous
  ArSignal* signals[kArMaxRanks];  // peer signal blocks; signals[rank] is ours
  int rank;
  void* out;                       // MODE 0/1: T [rows, hidden]; MODE 2: fp8 [rows, hidden]
  void* residual;                  // MODE 1/2: T [rows, hidden], updated in place
  const void* weight;              // MODE 1/2: T [hidden]
  float* scales;                   // MODE 2: [rows]
  const float* scale_ub;           // MODE 2: optional
  float eps;
  int rows, hidden;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <int N>
__device__ __forceinline__ void ar_barrier(const ArParams& p) {
  __syncthreads();  // every thread of the CTA is done with what the barrier protects
  if (threadIdx.x < N) {
    ArSignal* self = p.signals[p.rank];
    const uint32_t flag = self->self_counter[blockIdx.x][threadIdx.x] + 1;
    self->self_counter[blockIdx.x][threadIdx.x] = flag;
    __threadfence_system();  // this rank's earlier writes (the GEMM output, via the kernel boundary) before the flag
    st_release_sys(&p.signals[threadIdx.x]->peer_counter[flag & 1][blockIdx.x][p.rank], flag);
    const uint32_t* mine = &self->peer_counter[flag & 1][blockIdx.x][threadIdx.x];
    while (ld_acquire_sys(mine) != flag) {
    }
  }
  __syncthreads();
}

__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
  uint4 v;  // peer memory is written by another GPU: no non-coherent (.nc) path, no L1 allocation
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

template <typename T>
union ArVec8 {
  uint4 u;
  T e[8];
};

template <bool IS_MAX>
__device__ __forceinline__ float ar_block_reduce(float v, float* sm) {
  v = IS_MAX ? warp_max(v) : warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = ((threadIdx.x & 31) < (kArThreads >> 5)) ? sm[threadIdx.x & 31] : 0.f;
  return IS_MAX ? warp_max(r) : warp_sum(r);
}

// sum of the N peers' vectors at element offset `off`, fp32, rank order; rounded to T
template <typename T, int N>
__device__ __forceinline__ void ar_sum8(const ArParams& p, int64_t off, float (&acc)[8]) {
  ArVec8<T> v[N];
#pragma unroll
  for (int r = 0; r < N; ++r) v[r].u = ld_peer_v4(reinterpret_cast<const T*>(p.data[r]) + off);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float s = to_float<T>(v[0].e[j]);
#pragma unroll
    for (int r = 1; r < N; ++r) s += to_float<T>(v[r].e[j]);
    acc[j] = to_float<T>(from_float<T>(s));
  }
}

template <typename T, int N, int MODE>
__global__ void __launch_bounds__(kArThreads, 2) peer_allreduce_kernel(const ArParams p) {
  __shared__ float sm[kArThreads / 32];
  ar_barrier<N>(p);
  if (MODE == 0) {
    const int64_t nvec = (int64_t)p.rows * p.hidden / 8;
    for (int64_t i = (int64_t)blockIdx.x * kArThreads + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * kArThreads) {
      float acc[8];
      ar_sum8<T, N>(p, i * 8, acc);
      ArVec8<T> o;
#pragma unroll
      for (int j = 0; j < 8; ++j) o.e[j] = from_float<T>(acc[j]);
      *reinterpret_cast<uint4*>(reinterpret_cast<T*>(p.out) + i * 8) = o.u;
    }
  } else {
    const int nvec = p.hidden >> 3;
    const T* weight = reinterpret_cast<const T*>(p.weight);
    for (int row = blockIdx.x; row < p.rows; row += gridDim.x) {
      T* res = reinterpret_cast<T*>(p.residual) + (int64_t)row * p.hidden;
      float zf[kArMaxVec][8];
      float ss = 0.f;
#pragma unroll
      for (int k = 0; k < kArMaxVec; ++k) {
        const int i = threadIdx.x + k * kArThreads;
        if (i < nvec) {
          float x[8];
          ar_sum8<T, N>(p, (int64_t)row * p.hidden + i * 8, x);
          ArVec8<T> b, z;
          b.u = *reinterpret_cast<const uint4*>(res + i * 8);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float s = x[j] + to_float<T>(b.e[j]);
            z.e[j] = from_float<T>(s);
            zf[k][j] = (MODE == 1) ? to_float<T>(z.e[j]) : s;
            ss = fmaf(zf[k][j], zf[k][j], ss);
          }
          *reinterpret_cast<uint4*>(res + i * 8) = z.u;
        }
      }
      ss = ar_block_reduce<false>(ss, sm);
      const float inv = rsqrtf(ss / (float)p.hidden + p.eps);
      float amax = 0.f;
#pragma unroll
      for (int k = 0; k < kArMaxVec; ++k) {
        const int i = threadIdx.x + k * kArThreads;
        if (i < nvec) {
          ArVec8<T> w, y;
          w.u = *reinterpret_cast<const uint4*>(weight + i * 8);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const T n = from_float<T>(zf[k][j] * inv);
            y.e[j] = from_float<T>(to_float<T>(n) * to_float<T>(w.e[j]));
            if (MODE == 2) {
              zf[k][j] = to_float<T>(y.e[j]);
              amax = fmaxf(amax, fabsf(zf[k][j]));
            }
          }
          if (MODE == 1)
            *reinterpret_cast<uint4*>(reinterpret_cast<T*>(p.out) + (int64_t)row * p.hidden + i * 8) = y.u;
        }
      }
      if (MODE == 2) {
        amax = ar_block_reduce<true>(amax, sm);
        float s = p.scale_ub ? fminf(amax, *p.scale_ub) : amax;
        s = fmaxf(s / kFp8MaxA, 1.0f / (kFp8MaxA * 512.0f));
        if (threadIdx.x == 0) p.scales[row] = s;
        uint8_t* o = reinterpret_cast<uint8_t*>(p.out) + (int64_t)row * p.hidden;
#pragma unroll
        for (int k = 0; k < kArMaxVec; ++k) {
          const int i = threadIdx.x + k * kArThreads;
          if (i < nvec) {
            union { uint2 u; uint8_t b[8]; } q;
#pragma unroll
            for (int j = 0; j < 8; ++j)
              q.b[j] = float_to_e4m3(fmaxf(-kFp8MaxA, fminf(zf[k][j] / s, kFp8MaxA)));
            *reinterpret_cast<uint2*>(o + i * 8) = q.u;
          }
        }
      }
    }
  }
  ar_barrier<N>(p);
}

template <typename T, int N>
static int launch_mode(int mode, int grid, const ArParams& p, cudaStream_t stream) {
  if (mode == 0) peer_allreduce_kernel<T, N, 0><<<grid, kArThreads, 0, stream>>>(p);
  else if (mode == 1) peer_allreduce_kernel<T, N, 1><<<grid, kArThreads, 0, stream>>>(p);
  else peer_allreduce_kernel<T, N, 2><<<grid, kArThreads, 0, stream>>>(p);
  VB_LAUNCH_CHECK();
  return 0;
}

template <typename T>
static int launch_ranks(int n, int mode, int grid, const ArParams& p, cudaStream_t stream) {
  switch (n) {
    case 2: return launch_mode<T, 2>(mode, grid, p, stream);
    case 4: return launch_mode<T, 4>(mode, grid, p, stream);
    case 8: return launch_mode<T, 8>(mode, grid, p, stream);
    default: VB_CHECK(false, "peer_allreduce: world size %d not in {2, 4, 8}", n);
  }
  return 1;
}

}  // namespace vb200

using namespace vb200;

extern "C" size_t vb200_peer_allreduce_signal_bytes(void) { return sizeof(ArSignal); }

// mode 0: plain sum; 1: + residual add + RMSNorm; 2: + residual add + RMSNorm + dynamic per-token fp8 quant.
// peer_data / peer_signals: HOST arrays of `world` device pointers (this process's mappings of every rank's
// partial-sum buffer and zero-initialised signal block; entry `rank` is this rank's own).  Every rank must call
// with the same rows / hidden / mode, in the same order.
extern "C" int vb200_peer_allreduce(int mode, void* out, const void* const* peer_data,
                                    void* const* peer_signals, int rank, int world, int rows, int hidden,
                                    int dtype, void* residual, const void* weight, float epsilon,
                                    float* scales, const float* scale_ub, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(mode >= 0 && mode <= 2, "peer_allreduce: bad mode %d", mode);
  VB_CHECK(dtype == VB200_BF16 || dtype == VB200_F16, "peer_allreduce: dtype must be bf16 or f16");
  VB_CHECK(rank >= 0 && rank < world && world <= kArMaxRanks, "peer_allreduce: bad rank %d / world %d", rank, world);
  VB_CHECK(rows > 0 && hidden > 0 && hidden % 8 == 0, "peer_allreduce: hidden (%d) must be a positive multiple of 8", hidden);
  VB_CHECK(mode == 0 || hidden <= kArThreads * kArMaxVec * 8, "peer_allreduce: hidden %d too large for the fused modes", hidden);
  VB_CHECK(mode == 0 || (residual != nullptr && weight != nullptr), "peer_allreduce: fused modes need residual and weight");
  VB_CHECK(mode != 2 || scales != nullptr, "peer_allreduce: mode 2 needs scales");
  ArParams p;
  for (int r = 0; r < kArMaxRanks; ++r) {
    p.data[r] = r < world ? peer_data[r] : nullptr;
    p.signals[r] = r < world ? reinterpret_cast<ArSignal*>(peer_signals[r]) : nullptr;
    VB_CHECK(r >= world || (p.data[r] != nullptr && p.signals[r] != nullptr), "peer_allreduce: null peer pointer");
    VB_CHECK(r >= world || (reinterpret_cast<uintptr_t>(p.data[r]) & 15) == 0, "peer_allreduce: peer buffers must be 16-byte aligned");
  }
  p.rank = rank;
  p.out = out;
  p.residual = residual;
  p.weight = weight;
  p.scales = scales;
  p.scale_ub = scale_ub;
  p.eps = epsilon;
  p.rows = rows;
  p.hidden = hidden;
  int grid;
  if (mode == 0) {
    const int64_t nvec = (int64_t)rows * hidden / 8;
    grid = (int)((nvec + kArThreads - 1) / kArThreads);
  } else {
    grid = rows;
  }
  // every CTA must be co-resident with its same-index CTA on the peers: two 256-thread CTAs per SM at most
  const int cap = 2 * num_sms() < kArMaxBlocks ? 2 * num_sms() : kArMaxBlocks;
  if (grid > cap) grid = cap;
  if (dtype == VB200_BF16) return launch_ranks<__nv_bfloat16>(world, mode, grid, p, stream);
  return launch_ranks<__half>(world, mode, grid, p, stream);
}

// ---- peer-visible buffers (CUDA IPC) ----------------------------------------------------------------------------
// The reference does the same in csrc/custom_all_reduce.cu (allocate_shared_buffer_and_handle / open_mem_handle):
// cudaMalloc + cudaIpcGetMemHandle on the owner, cudaIpcOpenMemHandle(cudaIpcMemLazyEnablePeerAccess) on the peers,
// each on ITS OWN current device, so the NVLink peer mapping is created for the device whose kernels will use it.
extern "C" int vb200_peer_buffer_alloc(size_t bytes, void** ptr, void* handle64) {
  VB_CHECK(bytes > 0 && ptr != nullptr && handle64 != nullptr, "peer_buffer_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  VB_CUDA_CHECK(cudaMalloc(&p, bytes));
  VB_CUDA_CHECK(cudaMemset(p, 0, bytes));
  VB_CUDA_CHECK(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  VB_CUDA_CHECK(cudaIpcGetMemHandle(&h, p));
  memcpy(handle64, &h, sizeof(h));
  *ptr = p;
  return 0;
}

extern "C" int vb200_peer_buffer_open(const void* handle64, void** ptr) {
  VB_CHECK(handle64 != nullptr && ptr != nullptr, "peer_buffer_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  void* p = nullptr;
  VB_CUDA_CHECK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr = p;
  return 0;
}

extern "C" int vb200_peer_buffer_close(void* ptr) {
  if (ptr) VB_CUDA_CHECK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

extern "C" int vb200_peer_buffer_free(void* ptr) {
  if (ptr) VB_CUDA_CHECK(cudaFree(ptr));
  return 0;
}
