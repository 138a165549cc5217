// Micro-benchmark (bring-up tool, not product code): does cp.async.bulk.prefetch.L2 make a cold buffer L2-resident?
// Times the same 128-bit ld.global.cg streaming pass over a 32 MiB buffer (a) cold (after 512 MiB of other traffic),
// (b) warm (second pass), (c) cold + bulk prefetches of `pf_bytes` each issued by a separate kernel first.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2_prefetch l2_prefetch.cu && ./l2_prefetch
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256, 1) read_kernel(const uint4* __restrict__ src, size_t cta_u4, uint32_t* sink) {
  const uint4* base = src + (size_t)blockIdx.x * cta_u4;
  uint32_t acc = 0;
  for (size_t i = threadIdx.x; i < cta_u4; i += 256 * 4) {
    uint4 v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = (i + j * 256 < cta_u4) ? __ldcg(base + i + j * 256) : make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int j = 0; j < 4; ++j) acc ^= v[j].x ^ v[j].y ^ v[j].z ^ v[j].w;
  }
  if (acc == 0x12345678u) sink[0] = acc;
}
__global__ void flush_kernel(uint4* buf, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) buf[i] = make_uint4(i, 1, 2, 3);
}
__global__ void prefetch_kernel(const uint8_t* src, size_t total, int pf_bytes) {
  const size_t n = total / pf_bytes;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src + i * pf_bytes), "r"(pf_bytes) : "memory");
}
__global__ void prefetch_line_kernel(const uint8_t* src, size_t total) {
  const size_t n = total / 128;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    asm volatile("prefetch.global.L2 [%0];" ::"l"(src + i * 128) : "memory");
}

int main() {
  const int ctas = 128;
  const size_t total = 32u << 20, flush_bytes = 512u << 20;
  uint8_t *src, *fl; uint32_t* sink;
  cudaMalloc(&src, total); cudaMemset(src, 1, total); cudaMalloc(&fl, flush_bytes); cudaMalloc(&sink, 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto timed_read = [&]() {
    cudaEventRecord(e0); read_kernel<<<ctas, 256>>>((const uint4*)src, total / 16 / ctas, sink); cudaEventRecord(e1);
    cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
  };
  auto flush = [&]() { flush_kernel<<<1024, 256>>>((uint4*)fl, flush_bytes / 16); cudaDeviceSynchronize(); };
  flush(); float cold = timed_read(); float warm = timed_read(); float warm2 = timed_read();
  printf("32 MiB read pass: cold %.1f us (%.2f TB/s), warm %.1f us (%.2f TB/s), warm again %.1f us\n", cold * 1e3, total / (cold * 1e-3) / 1e12,
         warm * 1e3, total / (warm * 1e-3) / 1e12, warm2 * 1e3);
  for (int pf : {512, 4096, 16384, 65536}) {
    flush();
    cudaEventRecord(e0); prefetch_kernel<<<128, 64>>>(src, total, pf); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float pms; cudaEventElapsedTime(&pms, e0, e1);
    cudaDeviceSynchronize();
    float t = timed_read();
    printf("bulk prefetch %6d B each (kernel %.1f us), then read pass: %.1f us (%.2f TB/s)\n", pf, pms * 1e3, t * 1e3, total / (t * 1e-3) / 1e12);
  }
  {
    flush();
    prefetch_line_kernel<<<256, 256>>>(src, total); cudaDeviceSynchronize();
    float t = timed_read();
    printf("prefetch.global.L2 per 128-B line, then read pass: %.1f us (%.2f TB/s)\n", t * 1e3, total / (t * 1e-3) / 1e12);
  }
  {  // prefetch with a delay before the read (is the prefetch still in flight when the kernel ends?)
    flush();
    prefetch_kernel<<<128, 64>>>(src, total, 4096); cudaDeviceSynchronize();
    flush_kernel<<<8, 64>>>((uint4*)fl, 4096); cudaDeviceSynchronize();   // ~10 us of nothing much
    float t = timed_read();
    printf("bulk prefetch 4096 B + delay, then read pass: %.1f us (%.2f TB/s)\n", t * 1e3, total / (t * 1e-3) / 1e12);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
