// Probe of the L2/HBM -> shared-memory feed rate of the bulk-copy (TMA) engine on B200, per SM and chip-wide, with and
// without cluster multicast.  Question it answers for the GEMM design: when every SM of a
// cluster needs the SAME activation tile, does a multicast load (each CTA fetches 1/C of the tile and the fabric
// replicates it into all C shared memories) relieve the per-SM limit, or is the limit the SM's own inbound port?
//
//   l2_feed_probe <grid> <cluster> <multicast 0|1> <source 0=L2-resident 1=HBM stream> [chunk_kb=16] [stages=8] [producers=1] [consumers=1]
//
// `producers` = number of warps (one issuing thread each) that take the chunks round-robin: one thread sustains only
// one bulk copy per ~170 ns whatever its size, several threads overlap their issue latencies.
//
// Every CTA keeps a ring of `stages` chunks in flight; a chunk counts as delivered when all its bytes have landed in
// THIS CTA's shared memory (mbarrier complete_tx).  Prints delivered GB/s per SM and in total, and the bytes actually
// requested from L2 per SM.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o bin/l2_feed_probe l2_feed_probe.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x)                                                                                   \
  do {                                                                                          \
    cudaError_t e = (x);                                                                        \
    if (e != cudaSuccess) {                                                                     \
      printf("CUDA error %s at line %d: %s\n", cudaGetErrorName(e), __LINE__, cudaGetErrorString(e)); \
      exit(1);                                                                                  \
    }                                                                                           \
  } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t n) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n));
}
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(b)), "r"(parity)
        : "memory");
  }
}
__device__ __forceinline__ uint32_t mapa(uint32_t a, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
  return r;
}
__device__ __forceinline__ void arrive_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_load_mc(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::
          "r"(smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar)), "h"(mask)
      : "memory");
}

struct P {
  const uint8_t* src;
  size_t region;      // bytes addressable (power of two)
  size_t stride;      // start offset of cluster i = i * stride
  int iters, stages, chunk, csize, multicast, producers, consumers;
};

__global__ void __launch_bounds__(288, 1) feed_kernel(P p, unsigned long long* sink) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty = full + 16;
  uint8_t* ring = smem + 256;
  uint32_t crank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
  const int cluster_id = blockIdx.x / p.csize;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], p.multicast ? p.csize : 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (p.csize > 1) {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
  const size_t mask = p.region - 1;
  const size_t base = (size_t)cluster_id * p.stride;
  if ((threadIdx.x & 31) == 0 && (int)(threadIdx.x >> 5) < p.producers) {
    // producer warp w: chunks w, w + producers, ...
    for (int it = threadIdx.x >> 5; it < p.iters; it += p.producers) {
      const int s = it % p.stages;
      if (it >= p.stages) mbar_wait(&empty[s], ((it / p.stages) & 1) ^ 1);
      mbar_expect(&full[s], (uint32_t)p.chunk);
      // without multicast every CTA of the cluster fetches the SAME chunk on its own (what the GEMM does today)
      const uint8_t* g = p.src + ((base + (size_t)it * p.chunk) & mask);
      uint8_t* d = ring + (size_t)s * p.chunk;
      if (!p.multicast) {
        bulk_load(d, g, (uint32_t)p.chunk, &full[s]);
      } else {
        const uint32_t part = (uint32_t)p.chunk / p.csize;
        bulk_load_mc(d + crank * part, g + crank * part, part, &full[s], (uint16_t)((1u << p.csize) - 1u));
      }
    }
  } else if ((threadIdx.x & 31) == 0 && (int)(threadIdx.x >> 5) < p.producers + p.consumers) {
    // consumer warp c: chunks c, c + consumers, ...: the chunk has landed -> release the stage (in every CTA that may
    // write into it)
    unsigned long long acc = 0;
    for (int it = (threadIdx.x >> 5) - p.producers; it < p.iters; it += p.consumers) {
      const int s = it % p.stages;
      mbar_wait(&full[s], (it / p.stages) & 1);
      acc += ring[(size_t)s * p.chunk + (it & 63)];
      if (!p.multicast) {
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&empty[s])) : "memory");
      } else {
        for (int r = 0; r < p.csize; ++r) arrive_remote(mapa(smem_u32(&empty[s]), (uint32_t)r));
      }
    }
    if (acc == 0xdeadbeefULL) *sink = acc;
  }
  __syncthreads();
  if (p.csize > 1) {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
}

int main(int argc, char** argv) {
  if (argc < 5) {
    printf("usage: %s grid cluster multicast source [chunk_kb] [stages]\n", argv[0]);
    return 1;
  }
  const int grid = atoi(argv[1]), csize = atoi(argv[2]), mc = atoi(argv[3]), source = atoi(argv[4]);
  const int chunk = (argc > 5 ? atoi(argv[5]) : 16) * 1024;
  const int stages = argc > 6 ? atoi(argv[6]) : 8;
  const int producers = argc > 7 ? atoi(argv[7]) : 1;
  const int consumers = argc > 8 ? atoi(argv[8]) : 1;
  if (grid % csize || stages > 16 || chunk % (16 * csize)) {
    printf("bad arguments\n");
    return 1;
  }
  const size_t region = source ? ((size_t)4 << 30) : ((size_t)32 << 20);
  uint8_t* buf;
  CK(cudaMalloc(&buf, region));
  CK(cudaMemset(buf, 1, region));
  unsigned long long* sink;
  CK(cudaMalloc(&sink, 8));
  P p;
  p.src = buf;
  p.region = region;
  p.csize = csize;
  p.multicast = mc;
  p.chunk = chunk;
  p.stages = stages;
  p.producers = producers;
  p.consumers = consumers;
  const int clusters = grid / csize;
  // L2-resident: clusters start 1 MiB + 4 KiB apart inside 32 MiB and walk 8 MiB each; HBM: private slices
  p.stride = source ? (region / clusters) / 4096 * 4096 : ((1u << 20) + 4096);
  const size_t per_cta = source ? (region / clusters) / 4096 * 4096 : ((size_t)64 << 20);
  p.iters = (int)(per_cta / chunk);
  const size_t smem = 256 + (size_t)stages * chunk;
  CK(cudaFuncSetAttribute(feed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (csize > 8) CK(cudaFuncSetAttribute(feed_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(32 * (producers + consumers));
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = csize;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int maxc = 0;
  cudaOccupancyMaxActiveClusters(&maxc, feed_kernel, &cfg);
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CK(cudaEventRecord(e0));
    CK(cudaLaunchKernelEx(&cfg, feed_kernel, p, sink));
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep > 0 && ms < best) best = ms;
  }
  const double delivered = (double)p.iters * chunk;            // bytes landed in each CTA's shared memory
  const double requested = mc ? delivered / csize : delivered;  // bytes each CTA asked L2 for
  printf("{\"grid\": %d, \"cluster\": %d, \"multicast\": %d, \"source\": \"%s\", \"chunk_kb\": %d, \"stages\": %d, \"producers\": %d, \"consumers\": %d, "
         "\"max_active_clusters\": %d, \"ms\": %.3f, \"delivered_GBs_per_sm\": %.1f, \"delivered_TBs_total\": %.2f, "
         "\"requested_TBs_total\": %.2f}\n",
         grid, csize, mc, source ? "hbm" : "l2", chunk / 1024, stages, producers, consumers, maxc, best, delivered / best / 1e6,
         delivered * grid / best / 1e9, requested * grid / best / 1e9);
  return 0;
}
