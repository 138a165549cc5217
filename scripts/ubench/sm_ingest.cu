// Micro-benchmark (bring-up tool, not product code): how many bytes per second ONE SM pulls from L2 into registers with
// 128-bit ld.global.cg loads (the decode attention's granule pattern, D granules in flight per warp) and into shared memory
// with cp.async.bulk (N copies of `bytes` in flight per CTA), 128 CTAs at once, source resident in L2.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o sm_ingest sm_ingest.cu && ./sm_ingest
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int D>
__global__ void __launch_bounds__(256, 1) ldg_kernel(const uint4* __restrict__ src, size_t cta_u4, int iters, uint32_t* sink) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint4* base = src + (size_t)blockIdx.x * cta_u4 + (size_t)w * (cta_u4 / 8);
  const int n_gran = (int)(cta_u4 / 8 / 256);           // 4 KiB granules per warp
  uint4 G[D][8];
  uint32_t acc = 0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int d = 0; d < D - 1; ++d)
#pragma unroll
      for (int j = 0; j < 8; ++j) G[d][j] = __ldcg(base + (size_t)d * 256 + j * 32 + lane);
    for (int g = 0; g < n_gran; g += D) {
#pragma unroll
      for (int k = 0; k < D; ++k) {
        if (g + k < n_gran) {
          if (g + k + D - 1 < n_gran)
#pragma unroll
            for (int j = 0; j < 8; ++j) G[(k + D - 1) % D][j] = __ldcg(base + (size_t)(g + k + D - 1) * 256 + j * 32 + lane);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc ^= G[k][j].x ^ G[k][j].y ^ G[k][j].z ^ G[k][j].w;
        }
      }
    }
  }
  if (acc == 0x12345678u) sink[0] = acc;
}

__global__ void __launch_bounds__(256, 1) tma_kernel(const uint8_t* __restrict__ src, size_t cta_bytes, int bytes, int nslots, int iters, uint32_t* sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint8_t* ring = smem + 1024;
  const uint8_t* base = src + (size_t)blockIdx.x * cta_bytes;
  const int n = (int)(cta_bytes / bytes);
  if (threadIdx.x == 0) {
    for (int i = 0; i < nslots; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[i])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  uint32_t acc = 0;
  if (threadIdx.x == 0) {
    // one thread: issue, wait in order, re-issue into the freed slot (the consumer side is a single word read)
    int issued = 0, done = 0;
    const int total = n * iters;
    while (done < total) {
      while (issued < total && issued - done < nslots) {
        const int s = issued % nslots;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[s])), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         smem_u32(ring + (size_t)s * bytes)),
                     "l"(base + (size_t)(issued % n) * bytes), "r"(bytes), "r"(smem_u32(&bars[s]))
                     : "memory");
        ++issued;
      }
      const int s = done % nslots;
      const uint32_t parity = (done / nslots) & 1;
      uint32_t ok = 0;
      while (!ok)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok)
                     : "r"(smem_u32(&bars[s])), "r"(parity)
                     : "memory");
      acc ^= *reinterpret_cast<volatile uint32_t*>(ring + (size_t)s * bytes);
      ++done;
    }
  }
  if (acc == 0x12345678u) sink[0] = acc;
}

// `lanes` lanes of warp 0 issue one copy each at the same time (one barrier per lane), wait, repeat
__global__ void __launch_bounds__(256, 1) tma_par_kernel(const uint8_t* __restrict__ src, size_t cta_bytes, int bytes, int lanes, int iters, uint32_t* sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint8_t* ring = smem + 1024;
  const uint8_t* base = src + (size_t)blockIdx.x * cta_bytes;
  const int n = (int)(cta_bytes / bytes);
  if (threadIdx.x == 0) {
    for (int i = 0; i < 32; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[i])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  uint32_t acc = 0;
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    const int rounds = n * iters / lanes;
    for (int r = 0; r < rounds; ++r) {
      if (lane < lanes) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[lane])), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         smem_u32(ring + (size_t)lane * bytes)),
                     "l"(base + (size_t)((r * lanes + lane) % n) * bytes), "r"(bytes), "r"(smem_u32(&bars[lane]))
                     : "memory");
        uint32_t ok = 0;
        while (!ok)
          asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                       : "=r"(ok)
                       : "r"(smem_u32(&bars[lane])), "r"((uint32_t)(r & 1))
                       : "memory");
        acc ^= *reinterpret_cast<volatile uint32_t*>(ring + (size_t)lane * bytes);
      }
      __syncwarp();
    }
  }
  if (acc == 0x12345678u) sink[0] = acc;
}

int main() {
  const int ctas = 128;
  const size_t cta_bytes = 768 * 1024;
  uint8_t* src; uint32_t* sink;
  cudaMalloc(&src, ctas * cta_bytes); cudaMemset(src, 1, ctas * cta_bytes); cudaMalloc(&sink, 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 50;
  auto report = [&](const char* name, float ms) {
    const double bytes = (double)ctas * cta_bytes * iters;
    printf("%-40s %8.3f ms  %7.1f GB/s per SM  %6.2f TB/s chip\n", name, ms, bytes / ctas / (ms * 1e-3) / 1e9, bytes / (ms * 1e-3) / 1e12);
  };
#define RUN_LDG(D)                                                                                      \
  {                                                                                                     \
    ldg_kernel<D><<<ctas, 256>>>((const uint4*)src, cta_bytes / 16, 2, sink);                           \
    cudaEventRecord(e0); ldg_kernel<D><<<ctas, 256>>>((const uint4*)src, cta_bytes / 16, iters, sink);  \
    cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1);         \
    char nm[64]; snprintf(nm, 64, "ld.global.cg 128-bit, %d granules/warp", D); report(nm, ms);         \
  }
  RUN_LDG(2) RUN_LDG(3) RUN_LDG(4) RUN_LDG(6)
  cudaFuncSetAttribute(tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(tma_par_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int sizes[] = {4096, 16384, 32768, 65536};
  for (int bytes : sizes)
    for (int nslots : {2, 4, 8}) {
      if ((size_t)bytes * nslots > 192 * 1024) continue;
      const size_t sm = 1024 + (size_t)bytes * nslots;
      tma_kernel<<<ctas, 256, sm>>>(src, cta_bytes, bytes, nslots, 2, sink);
      cudaEventRecord(e0); tma_kernel<<<ctas, 256, sm>>>(src, cta_bytes, bytes, nslots, iters, sink);
      cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1);
      char nm[64]; snprintf(nm, 64, "cp.async.bulk %5d B x %2d in flight", bytes, nslots); report(nm, ms);
    }
  for (int bytes : {2048, 4096, 16384})
    for (int lanes : {1, 4, 8, 16, 32}) {
      if ((size_t)bytes * lanes > 192 * 1024) continue;
      const size_t sm = 1024 + (size_t)bytes * lanes;
      tma_par_kernel<<<ctas, 256, sm>>>(src, cta_bytes, bytes, lanes, 2, sink);
      cudaEventRecord(e0); tma_par_kernel<<<ctas, 256, sm>>>(src, cta_bytes, bytes, lanes, iters, sink);
      cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1);
      char nm[64]; snprintf(nm, 64, "bulk %5d B, %2d lanes at once, wait all", bytes, lanes); report(nm, ms);
    }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
