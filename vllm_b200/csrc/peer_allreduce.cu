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
// rank r stores the counter into slot [b][r] of every peer's signal block (volatile store over NVLink) and spins
// (volatile loads of its own block) until all peers' counters show up.  A second barrier at the end keeps
// a rank from overwriting its buffer (next GEMM) while a peer is still reading it.
#include <cstring>

#include "common.cuh"

namespace vb200 {

constexpr int kArMaxRanks = 8;
constexpr int kArMaxBlocks = 160;
constexpr int kArThreads = 512;
constexpr int kArMaxVec = 4;   // rows up to 512 * 4 * 8 = 16384 elements stay in registers
constexpr int kArStages = 3;   // staged kernel: rows in flight per CTA (cp.async ring in shared memory)
constexpr int kArMaxStageBytes = 64 * 1024;
constexpr float kFp8MaxA = 448.f;

struct ArSignal {
  uint32_t self_counter[kArMaxBlocks][kArMaxRanks];
  uint32_t peer_counter[2][kArMaxBlocks][kArMaxRanks];
  // two-shot kernel: grid-wide "my shard is reduced" hand-shake
  uint32_t grid_arrive;                  // CTAs of this rank that finished the reduce-scatter (atomicInc, wraps to 0)
  uint32_t grid_epoch;                   // two-shot launches completed by this rank (every rank agrees on it)
  uint32_t done_flag[kArMaxRanks];       // [r]: last epoch whose shard rank r has published (written by rank r)
};

struct ArParams {
  const void* data[kArMaxRanks];   // peer partial sums (this process's mapping), [rows, hidden] contiguous
  void* gather[kArMaxRanks];       // two-shot: peer buffers holding each rank's REDUCED shard (same shape as data)
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

// Flag traffic is volatile (relaxed, system-visible) stores and loads WITHOUT fences, as in the reference's first /
// final barriers (custom_all_reduce.cuh:196-240): the data a flag announces was written by the PREVIOUS kernel on
// the stream (complete at this kernel's start), and the end barrier only has to order "my loads of your buffer have
// returned" (their values were consumed before the bar.sync) before "you may overwrite it".  A fenced version
// (fence.sys + st.release.sys / ld.acquire.sys in the spin) cost ~9 us per barrier.
__device__ __forceinline__ void st_flag(uint32_t* p, uint32_t v) {
  asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_flag(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <int N>
__device__ __forceinline__ void ar_barrier(const ArParams& p) {
  __syncthreads();  // every thread of the CTA is done with what the barrier protects
  if (threadIdx.x < N) {
    ArSignal* self = p.signals[p.rank];
    const uint32_t flag = self->self_counter[blockIdx.x][threadIdx.x] + 1;
    self->self_counter[blockIdx.x][threadIdx.x] = flag;
    st_flag(&p.signals[threadIdx.x]->peer_counter[flag & 1][blockIdx.x][p.rank], flag);
    const uint32_t* mine = &self->peer_counter[flag & 1][blockIdx.x][threadIdx.x];
    while (ld_flag(mine) != flag) {
    }
  }
  __syncthreads();
}

// Peer data is read with plain (weak) 16-byte loads that do not allocate in L1: every address is read once per
// kernel, after the barrier (bar.sync after the acquiring threads orders the CTA's loads behind it), and L1 does not
// survive a kernel boundary, so no stale line can be hit.  (Strong .sys loads here cost 2x in the first version.)
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* g) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
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

// residual add + RMSNorm (+ per-token fp8 quant) of one row whose all-reduced values `fetch(i, x)` delivers as 8 fp32
// numbers (already rounded to T) for 16-byte vector i of the row.  Arithmetic = rms_norm_kernel in layer_glue.cu.
template <typename T, int MODE, typename Fetch>
__device__ __forceinline__ void ar_norm_row(const ArParams& p, int row, float* sm, Fetch fetch) {
  const int nvec = p.hidden >> 3;
  const T* weight = reinterpret_cast<const T*>(p.weight);
  T* res = reinterpret_cast<T*>(p.residual) + (int64_t)row * p.hidden;
  float zf[kArMaxVec][8];
  float ss = 0.f;
#pragma unroll
  for (int k = 0; k < kArMaxVec; ++k) {
    const int i = threadIdx.x + k * kArThreads;
    if (i < nvec) {
      float x[8];
      fetch(i, x);
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

template <typename T, int N, int MODE>
__global__ void __launch_bounds__(kArThreads, 1) peer_allreduce_kernel(const ArParams p) {
  __shared__ float sm[kArThreads / 32];
  pdl_wait();   // the partial sums come from the previous kernel on the stream: flags only after it has completed
  pdl_trigger();
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
    for (int row = blockIdx.x; row < p.rows; row += gridDim.x)
      ar_norm_row<T, MODE>(p, row, sm,
                           [&](int i, float (&x)[8]) { ar_sum8<T, N>(p, (int64_t)row * p.hidden + i * 8, x); });
  }
  ar_barrier<N>(p);
}

// ---- two-shot (reduce-scatter + all-gather) variant ---------------------------------------------------------------
// The one-shot kernel pulls (N-1) copies of the whole tensor into every GPU: 14 MiB per GPU for [256 x 4096] bf16 at 8
// ranks, ~19 us of NVLink time alone.  Here rank r first reduces only ITS shard (rows [r R/N, (r+1) R/N), a contiguous
// range of 16-byte vectors dealt out to all CTAs in warp-sized chunks: one NVLink round trip, (N-1)/N of the tensor
// pulled) into its peer-visible gather buffer, then every rank reads each row ONCE from its owner and goes on with the
// residual add / norm / quant: 2 (N-1)/N of the tensor per GPU instead of (N-1).  The reference makes the same
// switch above 256 KB at 8 ranks (/root/reference/csrc/custom_all_reduce.cuh:320-360 cross_device_reduce_2stage,
// :560-572 dispatch).  Synchronisation:
//   start   per-CTA-index flag barrier, unfenced (partials come from the previous kernel; also: every peer has left
//           the previous all-reduce, so my gather buffer is free to overwrite)
//   middle  grid-wide: each CTA fences (sys) and bumps a local counter; the last one pushes epoch e into every peer's
//           done_flag[rank]; everybody spins locally until all N ranks show e, then fences (acquire)
//   end     none: peers read my partials only before `middle`, and my gather buffer only until they enter the next
//           all-reduce's `start`.
template <typename T, int N, int MODE>
__global__ void __launch_bounds__(kArThreads, 1) peer_allreduce_2shot_kernel(const ArParams p) {
  __shared__ float sm[kArThreads / 32];
  pdl_wait();
  pdl_trigger();
  ArSignal* self = p.signals[p.rank];
  ar_barrier<N>(p);
  const int nvec = p.hidden >> 3;
  // ---- reduce-scatter: my shard as a contiguous vector range [v0, v1) ----
  int64_t v0, v1;
  int rows_per = 0;
  if (MODE == 0) {
    const int64_t total = (int64_t)p.rows * nvec;
    const int64_t per = ((total + N - 1) / N + 31) & ~(int64_t)31;
    v0 = min((int64_t)p.rank * per, total);
    v1 = min(v0 + per, total);
  } else {
    rows_per = (p.rows + N - 1) / N;
    v0 = (int64_t)min(p.rank * rows_per, p.rows) * nvec;
    v1 = (int64_t)min((p.rank + 1) * rows_per, p.rows) * nvec;
  }
  {
    T* mine = reinterpret_cast<T*>(p.gather[p.rank]);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t nchunk = (v1 - v0 + 31) >> 5;
    for (int64_t c = blockIdx.x + (int64_t)gridDim.x * warp; c < nchunk; c += (int64_t)gridDim.x * (kArThreads / 32)) {
      const int64_t i = v0 + (c << 5) + lane;
      if (i < v1) {
        float acc[8];
        ar_sum8<T, N>(p, i * 8, acc);
        ArVec8<T> o;
#pragma unroll
        for (int j = 0; j < 8; ++j) o.e[j] = from_float<T>(acc[j]);
        *reinterpret_cast<uint4*>(mine + i * 8) = o.u;
      }
    }
  }
  // ---- middle barrier ----
  __syncthreads();
  if (threadIdx.x == 0) {
    // every CTA reads the epoch BEFORE it arrives and only the last arrival advances it: all CTAs see the same e,
    // whatever grids earlier launches used
    const uint32_t e = ld_flag(&self->grid_epoch) + 1;
    __threadfence_system();   // release: this CTA's gather stores (ordered before by the bar.sync) before the count
    if (atomicInc(&self->grid_arrive, gridDim.x - 1) == gridDim.x - 1) {
      __threadfence_system();   // acquire the other CTAs' counts (and with them their stores), then publish
      st_flag(&self->grid_epoch, e);
#pragma unroll
      for (int r = 0; r < N; ++r) st_flag(&p.signals[r]->done_flag[p.rank], e);
    }
    sm[0] = __uint_as_float(e);
  }
  __syncthreads();
  if (threadIdx.x < N) {
    const uint32_t e = __float_as_uint(sm[0]);
    while ((int32_t)(ld_flag(&self->done_flag[threadIdx.x]) - e) < 0) {
    }
    __threadfence_system();
  }
  __syncthreads();
  // ---- all-gather fused with the consumer ----
  if (MODE == 0) {
    const int64_t total = (int64_t)p.rows * nvec;
    const int64_t per = ((total + N - 1) / N + 31) & ~(int64_t)31;
    for (int64_t i = (int64_t)blockIdx.x * kArThreads + threadIdx.x; i < total; i += (int64_t)gridDim.x * kArThreads) {
      const int owner = (int)(i / per);
      const uint4 v = ld_peer_v4(reinterpret_cast<const T*>(p.gather[owner]) + i * 8);
      *reinterpret_cast<uint4*>(reinterpret_cast<T*>(p.out) + i * 8) = v;
    }
  } else {
    for (int row = blockIdx.x; row < p.rows; row += gridDim.x) {
      const T* src = reinterpret_cast<const T*>(p.gather[row / rows_per]) + (int64_t)row * p.hidden;
      ar_norm_row<T, MODE>(p, row, sm, [&](int i, float (&x)[8]) {
        ArVec8<T> v;
        v.u = ld_peer_v4(src + i * 8);
#pragma unroll
        for (int j = 0; j < 8; ++j) x[j] = to_float<T>(v.e[j]);
      });
    }
  }
}

// Staged variant of MODE 1 / 2 (kept as a tuning option, off by default -- see g_ar_staged): a CTA owns rows blockIdx.x, blockIdx.x + grid, ... and
// keeps kArStages rows in flight -- all N peers' copies of a row are fetched with cp.async (16 bytes, L2-only) into a
// shared-memory ring, so the NVLink round trip of the next rows overlaps the reduction / norm / quant of this one and
// far fewer CTAs (each one pays two system-scope barriers) cover the batch.
template <typename T, int N, int MODE>
__global__ void __launch_bounds__(kArThreads, 1) peer_allreduce_staged_kernel(const ArParams p) {
  extern __shared__ __align__(16) uint8_t ring[];
  __shared__ float sm[kArThreads / 32];
  const int nvec = p.hidden >> 3;
  const int row_bytes = p.hidden * (int)sizeof(T);
  const int stage_bytes = N * row_bytes;
  const T* weight = reinterpret_cast<const T*>(p.weight);
  const uint32_t ring_u32 = (uint32_t)__cvta_generic_to_shared(ring);
  pdl_wait();
  pdl_trigger();
  ar_barrier<N>(p);
  auto issue = [&](int k) {
    const int row = blockIdx.x + k * gridDim.x;
    if (row < p.rows) {
      const uint32_t dst = ring_u32 + (uint32_t)((k % kArStages) * stage_bytes);
#pragma unroll
      for (int r = 0; r < N; ++r) {
        const uint8_t* src = reinterpret_cast<const uint8_t*>(p.data[r]) + (int64_t)row * row_bytes;
        for (int i = threadIdx.x; i < nvec; i += kArThreads) cp_async16(dst + r * row_bytes + i * 16, src + i * 16);
      }
    }
    cp_async_commit();   // always: keeps the group count in step with k
  };
#pragma unroll
  for (int k = 0; k < kArStages - 1; ++k) issue(k);
  for (int k = 0;; ++k) {
    const int row = blockIdx.x + k * gridDim.x;
    if (row >= p.rows) break;
    issue(k + kArStages - 1);
    cp_async_wait<kArStages - 1>();
    __syncthreads();
    const uint8_t* st = ring + (size_t)(k % kArStages) * stage_bytes;
    T* res = reinterpret_cast<T*>(p.residual) + (int64_t)row * p.hidden;
    float zf[kArMaxVec][8];
    float ss = 0.f;
#pragma unroll
    for (int kk = 0; kk < kArMaxVec; ++kk) {
      const int i = threadIdx.x + kk * kArThreads;
      if (i < nvec) {
        ArVec8<T> v[N], b, z;
#pragma unroll
        for (int r = 0; r < N; ++r) v[r].u = *reinterpret_cast<const uint4*>(st + r * row_bytes + i * 16);
        b.u = *reinterpret_cast<const uint4*>(res + i * 8);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float a = to_float<T>(v[0].e[j]);
#pragma unroll
          for (int r = 1; r < N; ++r) a += to_float<T>(v[r].e[j]);
          const float s = to_float<T>(from_float<T>(a)) + to_float<T>(b.e[j]);
          z.e[j] = from_float<T>(s);
          zf[kk][j] = (MODE == 1) ? to_float<T>(z.e[j]) : s;
          ss = fmaf(zf[kk][j], zf[kk][j], ss);
        }
        *reinterpret_cast<uint4*>(res + i * 8) = z.u;
      }
    }
    ss = ar_block_reduce<false>(ss, sm);   // its barriers also fence this stage's reads from the refill two rows on
    const float inv = rsqrtf(ss / (float)p.hidden + p.eps);
    float amax = 0.f;
#pragma unroll
    for (int kk = 0; kk < kArMaxVec; ++kk) {
      const int i = threadIdx.x + kk * kArThreads;
      if (i < nvec) {
        ArVec8<T> w, y;
        w.u = *reinterpret_cast<const uint4*>(weight + i * 8);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const T n = from_float<T>(zf[kk][j] * inv);
          y.e[j] = from_float<T>(to_float<T>(n) * to_float<T>(w.e[j]));
          if (MODE == 2) {
            zf[kk][j] = to_float<T>(y.e[j]);
            amax = fmaxf(amax, fabsf(zf[kk][j]));
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
      for (int kk = 0; kk < kArMaxVec; ++kk) {
        const int i = threadIdx.x + kk * kArThreads;
        if (i < nvec) {
          union { uint2 u; uint8_t b[8]; } q;
#pragma unroll
          for (int j = 0; j < 8; ++j) q.b[j] = float_to_e4m3(fmaxf(-kFp8MaxA, fminf(zf[kk][j] / s, kFp8MaxA)));
          *reinterpret_cast<uint2*>(o + i * 8) = q.u;
        }
      }
    }
  }
  cp_async_wait<0>();
  ar_barrier<N>(p);
}

// ---- split variant: reduce-scatter + consumer | all-gather, as TWO kernels -------------------------------------------
// What the measurements said about the single-kernel two-shot above (profiles/r02_multi_gpu_tp8_s44.txt: 34.8 us at
// 8 ranks vs 33.2 us one-shot; 27.8 vs 14.3 us at 2 ranks): its grid-wide + cross-GPU MIDDLE barrier, which has to be
// fenced at system scope because it publishes data written inside the same kernel, costs more than the NVLink time
// the shard-wise reduction saves.  Splitting at that point turns the middle barrier into a kernel boundary: kernel 1
// (start barrier, unfenced as before) reduces ONLY this rank's rows, runs the residual add / RMSNorm (/ quant) on
// them -- the consumer's work is sharded N ways too -- and leaves normed rows, updated residual rows and scales in its
// peer-visible gather buffer; kernel 2 starts with the same unfenced flag barrier (what it announces was written by
// the peer's PREVIOUS kernel) and copies every row from its owner.  Per GPU (N-1)/N of the partial sums are pulled
// once, and the gathered bytes are the fp8 rows + residual rows instead of N-1 copies of the bf16 tensor.
//   gather buffer of a rank: [out rows | residual rows | scales], each region sized for ALL rows (only the owner's
//   rows are valid): offsets rows*hidden*sizeof(T), 2*rows*hidden*sizeof(T).
template <typename T, int N, int MODE>
__global__ void __launch_bounds__(kArThreads, 1) peer_reduce_scatter_kernel(const ArParams p) {
  __shared__ float sm[kArThreads / 32];
  pdl_wait();
  pdl_trigger();
  ar_barrier<N>(p);
  const int rows_per = (p.rows + N - 1) / N;
  const int r0 = min(p.rank * rows_per, p.rows), r1 = min(r0 + rows_per, p.rows);
  uint8_t* g = reinterpret_cast<uint8_t*>(p.gather[p.rank]);
  const size_t region = (size_t)p.rows * p.hidden * sizeof(T);
  if (MODE == 0) {
    const int nvec = p.hidden >> 3;
    T* dst = reinterpret_cast<T*>(g);
    for (int64_t i = (int64_t)r0 * nvec + (int64_t)blockIdx.x * kArThreads + threadIdx.x; i < (int64_t)r1 * nvec;
         i += (int64_t)gridDim.x * kArThreads) {
      float acc[8];
      ar_sum8<T, N>(p, i * 8, acc);
      ArVec8<T> o;
#pragma unroll
      for (int j = 0; j < 8; ++j) o.e[j] = from_float<T>(acc[j]);
      *reinterpret_cast<uint4*>(dst + i * 8) = o.u;
    }
    return;
  }
  ArParams q = p;
  q.out = g;                                                  // normed (/ quantised) rows, indexed by global row
  q.scales = reinterpret_cast<float*>(g + 2 * region);
  T* res_copy = reinterpret_cast<T*>(g + region);
  const int nvec = p.hidden >> 3;
  for (int row = r0 + blockIdx.x; row < r1; row += gridDim.x) {
    ar_norm_row<T, MODE>(q, row, sm,
                         [&](int i, float (&x)[8]) { ar_sum8<T, N>(p, (int64_t)row * p.hidden + i * 8, x); });
    // the updated residual row travels too: every rank ends with the whole residual, as after the reference's op
    __syncthreads();
    const T* res = reinterpret_cast<const T*>(p.residual) + (int64_t)row * p.hidden;
    for (int i = threadIdx.x; i < nvec; i += kArThreads)
      *reinterpret_cast<uint4*>(res_copy + (int64_t)row * p.hidden + i * 8) = *reinterpret_cast<const uint4*>(res + i * 8);
  }
}

template <typename T, int N, int MODE>
__global__ void __launch_bounds__(kArThreads, 1) peer_all_gather_kernel(const ArParams p) {
  pdl_wait();     // this rank's kernel 1 has completed (and with it everything it wrote)
  pdl_trigger();
  ar_barrier<N>(p);   // every peer is past ITS kernel 1 as well
  const int rows_per = (p.rows + N - 1) / N;
  const size_t region = (size_t)p.rows * p.hidden * sizeof(T);
  const int out_row_bytes = MODE == 2 ? p.hidden : p.hidden * (int)sizeof(T);
  const int ovec = out_row_bytes >> 4, rvec = (p.hidden * (int)sizeof(T)) >> 4;
  const int per_row = ovec + (MODE == 0 ? 0 : rvec);
  const int64_t total = (int64_t)p.rows * per_row;
  for (int64_t i = (int64_t)blockIdx.x * kArThreads + threadIdx.x; i < total; i += (int64_t)gridDim.x * kArThreads) {
    const int row = (int)(i / per_row);
    const int v = (int)(i - (int64_t)row * per_row);
    const uint8_t* g = reinterpret_cast<const uint8_t*>(p.gather[row / rows_per]);
    if (v < ovec) {
      const uint4 x = ld_peer_v4(g + (int64_t)row * out_row_bytes + v * 16);
      *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(p.out) + (int64_t)row * out_row_bytes + v * 16) = x;
    } else if (row / rows_per != p.rank) {   // my own residual rows are already in place
      const int vv = v - ovec;
      const uint4 x = ld_peer_v4(g + region + (int64_t)row * p.hidden * sizeof(T) + vv * 16);
      *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(p.residual) + (int64_t)row * p.hidden * sizeof(T) + vv * 16) = x;
    }
  }
  if (MODE == 2) {
    for (int row = blockIdx.x * kArThreads + threadIdx.x; row < p.rows; row += gridDim.x * kArThreads) {
      const float* sc = reinterpret_cast<const float*>(reinterpret_cast<const uint8_t*>(p.gather[row / rows_per]) + 2 * region);
      float v;
      asm volatile("ld.global.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(sc + row) : "memory");
      p.scales[row] = v;
    }
  }
}

static int g_ar_blocks = 0;       // 0 = default; debug / tuning hook below
static int g_ar_staged = 0;       // measured: direct loads win at 8 ranks (33 vs 52 us) and tie at 2 (14.6 vs 15.1 us)

template <typename T, int N>
static int launch_mode(int mode, int grid, const ArParams& p, bool two_shot, cudaStream_t stream) {
  const int stage_bytes = N * p.hidden * (int)sizeof(T);
  if (two_shot) {
    if (mode == 0) VB_CUDA_CHECK(launch_k(peer_allreduce_2shot_kernel<T, N, 0>, dim3(grid), dim3(kArThreads), 0, stream, p));
    else if (mode == 1) VB_CUDA_CHECK(launch_k(peer_allreduce_2shot_kernel<T, N, 1>, dim3(grid), dim3(kArThreads), 0, stream, p));
    else VB_CUDA_CHECK(launch_k(peer_allreduce_2shot_kernel<T, N, 2>, dim3(grid), dim3(kArThreads), 0, stream, p));
    return 0;
  }
  if (mode != 0 && g_ar_staged && stage_bytes <= kArMaxStageBytes) {
    const size_t smem = (size_t)kArStages * stage_bytes;
    if (mode == 1) {
      auto kern = peer_allreduce_staged_kernel<T, N, 1>;
      VB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kArStages * kArMaxStageBytes));
      VB_CUDA_CHECK(launch_k(kern, dim3(grid), dim3(kArThreads), smem, stream, p));
    } else {
      auto kern = peer_allreduce_staged_kernel<T, N, 2>;
      VB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kArStages * kArMaxStageBytes));
      VB_CUDA_CHECK(launch_k(kern, dim3(grid), dim3(kArThreads), smem, stream, p));
    }
    return 0;
  }
  if (mode == 0) VB_CUDA_CHECK(launch_k(peer_allreduce_kernel<T, N, 0>, dim3(grid), dim3(kArThreads), 0, stream, p));
  else if (mode == 1) VB_CUDA_CHECK(launch_k(peer_allreduce_kernel<T, N, 1>, dim3(grid), dim3(kArThreads), 0, stream, p));
  else VB_CUDA_CHECK(launch_k(peer_allreduce_kernel<T, N, 2>, dim3(grid), dim3(kArThreads), 0, stream, p));
  return 0;
}

template <typename T, int N>
static int launch_split(int mode, const ArParams& p, cudaStream_t stream) {
  const int rows_per = (p.rows + N - 1) / N;
  int cap = g_ar_blocks > 0 ? g_ar_blocks : num_sms();
  if (cap > num_sms()) cap = num_sms();
  if (cap > kArMaxBlocks) cap = kArMaxBlocks;
  int g1 = mode == 0 ? (int)(((int64_t)rows_per * p.hidden / 8 + kArThreads - 1) / kArThreads) : rows_per;
  if (g1 > cap) g1 = cap;
  if (g1 < 1) g1 = 1;
  const int64_t vecs = (int64_t)p.rows * p.hidden * (mode == 2 ? 3 : (mode == 1 ? 4 : 2)) / 16;
  int g2 = (int)((vecs + kArThreads - 1) / kArThreads);
  if (g2 > cap) g2 = cap;
  if (g2 < 1) g2 = 1;
  if (mode == 0) {
    VB_CUDA_CHECK(launch_k(peer_reduce_scatter_kernel<T, N, 0>, dim3(g1), dim3(kArThreads), 0, stream, p));
    VB_CUDA_CHECK(launch_k(peer_all_gather_kernel<T, N, 0>, dim3(g2), dim3(kArThreads), 0, stream, p));
  } else if (mode == 1) {
    VB_CUDA_CHECK(launch_k(peer_reduce_scatter_kernel<T, N, 1>, dim3(g1), dim3(kArThreads), 0, stream, p));
    VB_CUDA_CHECK(launch_k(peer_all_gather_kernel<T, N, 1>, dim3(g2), dim3(kArThreads), 0, stream, p));
  } else {
    VB_CUDA_CHECK(launch_k(peer_reduce_scatter_kernel<T, N, 2>, dim3(g1), dim3(kArThreads), 0, stream, p));
    VB_CUDA_CHECK(launch_k(peer_all_gather_kernel<T, N, 2>, dim3(g2), dim3(kArThreads), 0, stream, p));
  }
  return 0;
}

template <typename T>
static int launch_ranks(int n, int mode, int grid, const ArParams& p, bool two_shot, cudaStream_t stream) {
  switch (n) {
    case 2: return launch_mode<T, 2>(mode, grid, p, two_shot, stream);
    case 4: return launch_mode<T, 4>(mode, grid, p, two_shot, stream);
    case 8: return launch_mode<T, 8>(mode, grid, p, two_shot, stream);
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
// with the same rows / hidden / mode, in the same order.  peer_gather (two-shot only): the same for every rank's
// gather buffer (at least rows * hidden elements, 16-byte aligned).
static int run_peer_allreduce(int mode, void* out, const void* const* peer_data, void* const* peer_gather,
                              void* const* peer_signals, int rank, int world, int rows, int hidden, int dtype,
                              void* residual, const void* weight, float epsilon, float* scales,
                              const float* scale_ub, cudaStream_t stream) {
  VB_CHECK(mode >= 0 && mode <= 2, "peer_allreduce: bad mode %d", mode);
  VB_CHECK(dtype == VB200_BF16 || dtype == VB200_F16, "peer_allreduce: dtype must be bf16 or f16");
  VB_CHECK(rank >= 0 && rank < world && world <= kArMaxRanks, "peer_allreduce: bad rank %d / world %d", rank, world);
  VB_CHECK(rows > 0 && hidden > 0 && hidden % 8 == 0, "peer_allreduce: hidden (%d) must be a positive multiple of 8", hidden);
  VB_CHECK(mode == 0 || hidden <= kArThreads * kArMaxVec * 8, "peer_allreduce: hidden %d too large for the fused modes", hidden);
  VB_CHECK(mode == 0 || (residual != nullptr && weight != nullptr), "peer_allreduce: fused modes need residual and weight");
  VB_CHECK(mode != 2 || scales != nullptr, "peer_allreduce: mode 2 needs scales");
  const bool two_shot = peer_gather != nullptr;
  ArParams p;
  for (int r = 0; r < kArMaxRanks; ++r) {
    p.data[r] = r < world ? peer_data[r] : nullptr;
    p.gather[r] = (two_shot && r < world) ? peer_gather[r] : nullptr;
    p.signals[r] = r < world ? reinterpret_cast<ArSignal*>(peer_signals[r]) : nullptr;
    VB_CHECK(r >= world || (p.data[r] != nullptr && p.signals[r] != nullptr), "peer_allreduce: null peer pointer");
    VB_CHECK(r >= world || (reinterpret_cast<uintptr_t>(p.data[r]) & 15) == 0, "peer_allreduce: peer buffers must be 16-byte aligned");
    VB_CHECK(!two_shot || r >= world || (p.gather[r] != nullptr && (reinterpret_cast<uintptr_t>(p.gather[r]) & 15) == 0),
             "peer_allreduce: gather buffers must be non-null and 16-byte aligned");
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
  // Grid: every CTA pays the system-scope barriers, and all CTAs of all ranks must be co-resident
  // (1 CTA/SM, grid <= #SMs).  More CTAs = fewer rows per CTA = fewer serialised NVLink round trips (measured).
  int cap = g_ar_blocks > 0 ? g_ar_blocks : num_sms();
  if (cap > num_sms()) cap = num_sms();
  if (cap > kArMaxBlocks) cap = kArMaxBlocks;
  int grid;
  if (mode == 0) {
    const int64_t nvec = (int64_t)rows * hidden / 8;
    grid = (int)((nvec + kArThreads - 1) / kArThreads);
  } else {
    grid = rows;
  }
  if (grid > cap) grid = cap;
  if (dtype == VB200_BF16) return launch_ranks<__nv_bfloat16>(world, mode, grid, p, two_shot, stream);
  return launch_ranks<__half>(world, mode, grid, p, two_shot, stream);
}

extern "C" int vb200_peer_allreduce(int mode, void* out, const void* const* peer_data,
                                    void* const* peer_signals, int rank, int world, int rows, int hidden,
                                    int dtype, void* residual, const void* weight, float epsilon,
                                    float* scales, const float* scale_ub, vb200_stream_t stream_) {
  return run_peer_allreduce(mode, out, peer_data, nullptr, peer_signals, rank, world, rows, hidden, dtype, residual,
                            weight, epsilon, scales, scale_ub, (cudaStream_t)stream_);
}

extern "C" int vb200_peer_allreduce_2shot(int mode, void* out, const void* const* peer_data,
                                          void* const* peer_gather, void* const* peer_signals, int rank, int world,
                                          int rows, int hidden, int dtype, void* residual, const void* weight,
                                          float epsilon, float* scales, const float* scale_ub,
                                          vb200_stream_t stream_) {
  VB_CHECK(peer_gather != nullptr, "peer_allreduce_2shot: peer_gather is required");
  return run_peer_allreduce(mode, out, peer_data, peer_gather, peer_signals, rank, world, rows, hidden, dtype,
                            residual, weight, epsilon, scales, scale_ub, (cudaStream_t)stream_);
}

// Split variant (reduce-scatter + consumer | all-gather, two launches; see the kernels).  peer_gather: every rank's
// gather buffer, at least 2 * rows * hidden * sizeof(T) + 4 * rows bytes.  Same results as vb200_peer_allreduce.
extern "C" int vb200_peer_allreduce_split(int mode, void* out, const void* const* peer_data, void* const* peer_gather,
                                          void* const* peer_signals, int rank, int world, int rows, int hidden,
                                          int dtype, void* residual, const void* weight, float epsilon, float* scales,
                                          const float* scale_ub, vb200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  VB_CHECK(mode >= 0 && mode <= 2, "peer_allreduce_split: bad mode %d", mode);
  VB_CHECK(dtype == VB200_BF16 || dtype == VB200_F16, "peer_allreduce_split: dtype must be bf16 or f16");
  VB_CHECK(rank >= 0 && rank < world && world <= kArMaxRanks, "peer_allreduce_split: bad rank %d / world %d", rank, world);
  VB_CHECK(rows > 0 && hidden > 0 && hidden % 16 == 0, "peer_allreduce_split: hidden (%d) must be a positive multiple of 16", hidden);
  VB_CHECK(mode == 0 || hidden <= kArThreads * kArMaxVec * 8, "peer_allreduce_split: hidden %d too large for the fused modes", hidden);
  VB_CHECK(mode == 0 || (residual != nullptr && weight != nullptr), "peer_allreduce_split: fused modes need residual and weight");
  VB_CHECK(mode != 2 || scales != nullptr, "peer_allreduce_split: mode 2 needs scales");
  VB_CHECK(peer_gather != nullptr, "peer_allreduce_split: peer_gather is required");
  ArParams p;
  for (int r = 0; r < kArMaxRanks; ++r) {
    p.data[r] = r < world ? peer_data[r] : nullptr;
    p.gather[r] = r < world ? peer_gather[r] : nullptr;
    p.signals[r] = r < world ? reinterpret_cast<ArSignal*>(peer_signals[r]) : nullptr;
    VB_CHECK(r >= world || (p.data[r] && p.gather[r] && p.signals[r]), "peer_allreduce_split: null peer pointer");
    VB_CHECK(r >= world || ((reinterpret_cast<uintptr_t>(p.data[r]) | reinterpret_cast<uintptr_t>(p.gather[r])) & 15) == 0,
             "peer_allreduce_split: peer buffers must be 16-byte aligned");
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
#define VB_SPLIT(T_)                                                     \
  switch (world) {                                                       \
    case 2: return launch_split<T_, 2>(mode, p, stream);                 \
    case 4: return launch_split<T_, 4>(mode, p, stream);                 \
    case 8: return launch_split<T_, 8>(mode, p, stream);                 \
    default: VB_CHECK(false, "peer_allreduce_split: world size %d not in {2, 4, 8}", world); \
  }
  if (dtype == VB200_BF16) { VB_SPLIT(__nv_bfloat16) } else { VB_SPLIT(__half) }
#undef VB_SPLIT
  return 1;
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


// Debug / tuning hook (not part of the drop-in surface): CTAs per launch (0 = default 64) and whether the fused modes
// use the cp.async-staged kernel or the direct-load kernel (default); CTAs 0 = one per SM.  Must be set identically on every rank.
extern "C" void vb200_debug_peer_allreduce_tune(int blocks, int staged) {
  g_ar_blocks = blocks;
  g_ar_staged = staged;
}
