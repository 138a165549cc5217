// Micro-benchmark (bring-up tool, not product code): the LayerNorm row exchange of the head-cluster decode kernels in isolation.
// 32 clusters of 4 CTAs; per round every CTA delivers `rows` rows of 2 KiB (bf16 row of 1024) to the same place in all 4 CTAs:
//   (a) st.shared::cluster 8-byte stores by one warp per row + barrier.cluster (release / acquire)      -- what ln_cluster does
//   (b) local st.shared + fence.proxy.async + one cp.async.bulk.shared::cluster per (row, peer), completion on the receiver's mbarrier
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dsmem_exchange dsmem_exchange.cu && ./dsmem_exchange
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t cta_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t map_to(uint32_t local, uint32_t rank) {
  uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank)); return r;
}

constexpr int kRowBytes = 2064;

template <int MODE>
__global__ void __cluster_dims__(4, 1, 1) __launch_bounds__(256, 1) exch_kernel(int rows, int iters, unsigned long long* out, uint32_t* sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
  uint8_t* a_rm = smem + 1024;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t rank = cta_rank();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  uint32_t acc = 0;
  unsigned long long t0 = 0;
  for (int it = 0; it < iters + 10; ++it) {
    if (it == 10) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    const int r = rows * rank + w;                    // the row this warp delivers (w < rows)
    if (MODE == 0) {
      if (w < rows) {
        uint8_t* dst = a_rm + (size_t)r * kRowBytes;
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
          for (int d = 0; d < 4; ++d) {
            const uint32_t ra = map_to(smem_u32(dst + (128 * j + 4 * lane) * 2), d);
            asm volatile("st.shared::cluster.v2.u32 [%0], {%1, %2};" ::"r"(ra), "r"(it + lane), "r"(it + j) : "memory");
          }
      }
      asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    } else if (MODE == 1) {
      if (threadIdx.x == 0)
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(3 * rows * 2048) : "memory");
      if (w < rows) {
        uint8_t* dst = a_rm + (size_t)r * kRowBytes;
#pragma unroll
        for (int j = 0; j < 8; ++j) *reinterpret_cast<uint2*>(dst + (128 * j + 4 * lane) * 2) = make_uint2(it + lane, it + j);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane < 4 && lane != (int)rank) {
          const uint32_t ra = map_to(smem_u32(dst), lane), rb = map_to(smem_u32(bar), lane);
          asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(ra),
                       "r"(smem_u32(dst)), "r"(2048), "r"(rb)
                       : "memory");
        }
      }
      uint32_t ok = 0;
      while (!ok)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"((uint32_t)(it & 1)) : "memory");
    }
    // consume: every warp reads a bit of every row (the GEMM's ldmatrix)
    acc ^= *reinterpret_cast<volatile uint32_t*>(a_rm + (size_t)((w * 2 + it) % (4 * rows)) * kRowBytes + lane * 4);
    __syncthreads();     // everybody is done reading before the next round's rows may land
    if (MODE >= 1) asm volatile("barrier.cluster.arrive.relaxed.aligned;\n\tbarrier.cluster.wait.aligned;" ::: "memory");   // WAR guard (the kernel gets this from data dependencies)
  }
  unsigned long long t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  if (acc == 0x12345678u) sink[0] = acc;
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

int main() {
  unsigned long long* out; uint32_t* sink;
  cudaMallocManaged(&out, 8); cudaMalloc(&sink, 4);
  const int iters = 2000, smem = 1024 + 16 * kRowBytes;
  cudaFuncSetAttribute(exch_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(exch_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(exch_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  exch_kernel<2><<<128, 256, smem>>>(2, iters, out, sink); cudaDeviceSynchronize();
  printf("skeleton only (CTA barrier + relaxed cluster barrier, no data): %.3f us per round\n", out[0] / 1e3 / iters);
  for (int rows : {2, 4}) {
    exch_kernel<0><<<128, 256, smem>>>(rows, iters, out, sink); cudaDeviceSynchronize();
    printf("rows/CTA %d: st.shared::cluster + cluster barrier        %.3f us per exchange (incl. a CTA barrier)\n", rows, out[0] / 1e3 / iters);
    exch_kernel<1><<<128, 256, smem>>>(rows, iters, out, sink); cudaDeviceSynchronize();
    printf("rows/CTA %d: bulk smem->peer copies on the peer's mbarrier %.3f us per exchange (incl. a CTA barrier + a relaxed cluster barrier as WAR guard)\n", rows, out[0] / 1e3 / iters);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
