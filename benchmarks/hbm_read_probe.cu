// Read-only HBM bandwidth probe (bench infrastructure, not product): streams a buffer much larger than L2
// with 16-byte ld.global.nc.L1::no_allocate loads, UNROLL independent loads per thread in flight, and reports
// GB/s per (CTAs/SM, threads, unroll).  Gives the read ceiling the paged-attention kernel can approach
// (MEASURED_PEAKS.json's hbm_gbs is a copy: reads + writes).
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o benchmarks/bin/hbm_read_probe benchmarks/hbm_read_probe.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

template <int UNROLL>
__global__ void read_kernel(const uint4* __restrict__ p, size_t n_vec, unsigned* sink) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned acc = 0;
  for (; i + (UNROLL - 1) * stride < n_vec; i += UNROLL * stride) {
    uint4 v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const uint4* a = p + i + u * stride;
      asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                   : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w)
                   : "l"(a));
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x12345678u) *sink = acc;
}

template <int UNROLL>
static void run(const uint4* buf, size_t n_vec, unsigned* sink, int ctas_per_sm, int threads, int sms) {
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  const int grid = sms * ctas_per_sm;
  read_kernel<UNROLL><<<grid, threads>>>(buf, n_vec, sink);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int it = 0; it < 5; ++it) {
    cudaEventRecord(a);
    read_kernel<UNROLL><<<grid, threads>>>(buf, n_vec, sink);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    if (ms < best) best = ms;
  }
  printf("{\"probe\": \"hbm_read\", \"ctas_per_sm\": %d, \"threads\": %d, \"unroll\": %d, \"bytes_in_flight_per_sm\": %d, "
         "\"GBs\": %.1f}\n",
         ctas_per_sm, threads, UNROLL, ctas_per_sm * threads * UNROLL * 16, n_vec * 16.0 / best / 1e6);
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const size_t bytes = (size_t)4 << 30;
  uint4* buf;
  unsigned* sink;
  if (cudaMalloc(&buf, bytes) != cudaSuccess) return 1;
  cudaMalloc(&sink, 4);
  cudaMemset(buf, 1, bytes);
  const size_t n_vec = bytes / 16;
  for (int ctas = 1; ctas <= 4; ctas *= 2) {
    run<2>(buf, n_vec, sink, ctas, 256, sms);
    run<4>(buf, n_vec, sink, ctas, 256, sms);
    run<8>(buf, n_vec, sink, ctas, 256, sms);
    run<16>(buf, n_vec, sink, ctas, 256, sms);
  }
  run<8>(buf, n_vec, sink, 3, 128, sms);
  run<16>(buf, n_vec, sink, 3, 128, sms);
  run<8>(buf, n_vec, sink, 8, 128, sms);
  return 0;
}
