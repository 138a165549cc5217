// dev_utils.cuh -- sm_100a PTX wrappers: mbarrier, 1-D bulk (TMA engine) copies, ldmatrix / mma,
// cache-controlled loads, a watchdogged grid barrier, Philox4x32-10 and block reductions.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "layout.h"

namespace tts {

// Every device-side wait is bounded: a hung GPU box is a strike (no progress for ~2 s => give up,
// raise the watchdog flag, let every CTA fall through to the end of the kernel).
constexpr long long kWatchdogCycles = 4000000000LL;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// ---- mbarrier -----------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Returns false on watchdog expiry (or when *abort_flag is set by someone else).
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity, volatile int* abort_flag) {
  if (mbar_try_wait(bar, parity)) return true;
  long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > kWatchdogCycles || *abort_flag) {
      *abort_flag = 1;
      return false;
    }
  }
  return true;
}

// ---- bulk async copy global -> shared (UBLKCP), completion on an mbarrier ---------------------------------
__device__ __forceinline__ uint64_t make_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(dst_smem)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

// Bulk reduction shared -> global (UBLKRED): y[i] += buf[i] for `bytes` / 8 unsigned 64-bit words, each element an atomic add at L2,
// issued by one thread as one instruction; the group is committed at once (the source must stay untouched until the adds are done)
__device__ __forceinline__ void bulk_red_add_u64(void* gdst, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.u64 [%0], [%1], %2;\n\tcp.async.bulk.commit_group;" ::"l"(gdst),
               "r"(smem_u32(src_smem)), "r"(bytes)
               : "memory");
}

// 16-byte asynchronous copy global (L2, never L1) -> shared, and the wait for all of this thread's copies
__device__ __forceinline__ void cp_async_cg16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// ---- loads that must see other CTAs' writes (activations, KV written this launch): L2 only -----------------
__device__ __forceinline__ uint4 ld_cg_u4(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ float4 ld_cg_f4(const void* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float2 ld_cg_f2(const void* p) { return __ldcg(reinterpret_cast<const float2*>(p)); }
__device__ __forceinline__ float ld_cg_f(const float* p) { return __ldcg(p); }
__device__ __forceinline__ int ld_cg_i(const int* p) { return __ldcg(p); }
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
__device__ __forceinline__ void red_release_add_u32(uint32_t* p, uint32_t v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ---- thread-block cluster helpers (ar_decode.cu) ---------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_id_x() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}
// all (non-exited) threads of the cluster; used once at kernel start, before the producer warp goes its own way
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive (release, cluster scope) on the mbarrier at the same shared offset in CTA `rank` of this cluster
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* local_bar, uint32_t rank) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}" ::"r"(smem_u32(local_bar)),
      "r"(rank)
      : "memory");
}
// 8-byte store into the shared memory of CTA `rank` of this cluster, at the offset of `local_ptr`
__device__ __forceinline__ void st_cluster_u2(void* local_ptr, uint32_t rank, uint2 v) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "st.shared::cluster.v2.u32 [ra], {%2, %3};\n\t}" ::"r"(smem_u32(local_ptr)),
      "r"(rank), "r"(v.x), "r"(v.y)
      : "memory");
}
// 4-byte store into the shared memory of CTA `rank` that completes 4 bytes of the transaction count of the mbarrier at
// `local_bar`'s offset in the SAME CTA: the receiver waits on its own barrier, the sender needs no fence and no cluster barrier
__device__ __forceinline__ void st_async_u32(void* local_ptr, uint32_t rank, uint32_t v, uint64_t* local_bar) {
  asm volatile(
      "{\n\t.reg .b32 ra, rb;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mapa.shared::cluster.u32 rb, %3, %1;\n\t"
      "st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [ra], %2, [rb];\n\t}" ::"r"(smem_u32(local_ptr)),
      "r"(rank), "r"(v), "r"(smem_u32(local_bar))
      : "memory");
}
__device__ __forceinline__ void st_async_u2(void* local_ptr, uint32_t rank, uint2 v, uint64_t* local_bar) {
  asm volatile(
      "{\n\t.reg .b32 ra, rb;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mapa.shared::cluster.u32 rb, %4, %1;\n\t"
      "st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [ra], {%2, %3}, [rb];\n\t}" ::"r"(smem_u32(local_ptr)),
      "r"(rank), "r"(v.x), "r"(v.y), "r"(smem_u32(local_bar))
      : "memory");
}
__device__ __forceinline__ void st_cluster_u32(void* local_ptr, uint32_t rank, uint32_t v) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "st.shared::cluster.u32 [ra], %2;\n\t}" ::"r"(smem_u32(local_ptr)),
      "r"(rank), "r"(v)
      : "memory");
}
// fire-and-forget bulk prefetch of `bytes` (multiple of 16) of global memory into L2
__device__ __forceinline__ void bulk_prefetch_l2(const void* p, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// 2^-32 fixed point: integer adds commute, so sums accumulated with atomics do not depend on arrival order
constexpr float kFixScale = 4294967296.0f, kFixInv = 1.0f / 4294967296.0f;
__device__ __forceinline__ long long to_fixed(float v) { return __float2ll_rn(v * kFixScale); }
__device__ __forceinline__ float from_fixed(long long v) { return __ll2float_rn(v) * kFixInv; }
__device__ __forceinline__ void red_add_s64(long long* p, long long v) {
  asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ longlong2 ld_cg_ll2(const void* p) { return __ldcg(reinterpret_cast<const longlong2*>(p)); }

// ---- self-validating data (head-cluster decode kernels: two of the three grid barriers of a block are replaced by polling
// the data itself; ar_decode.cu) ------------------------------------------------------------------------------------------
// Counting accumulator: a 64-bit word = (2^-24 fixed-point value << 8) | count.  Every contribution is one red.add of
// (value << 8) + 1, so value and count change together, atomically, in L2: a reader that sees count == expected sees the
// complete sum -- no flag, no fence, no barrier.  Integer adds commute => the sum is still order-independent (deterministic).
constexpr float kCntScale = 16777216.0f, kCntInv = 1.0f / 16777216.0f;
__device__ __forceinline__ long long cnt_contrib(float v) { return (__float2ll_rn(v * kCntScale) << 8) + 1ll; }
__device__ __forceinline__ long long cnt_word(float v, int count) { return (__float2ll_rn(v * kCntScale) << 8) | (long long)count; }
__device__ __forceinline__ float cnt_value(long long w) { return __ll2float_rn(w >> 8) * kCntInv; }
__device__ __forceinline__ int cnt_count(long long w) { return (int)(w & 0xff); }
// strong (relaxed, gpu scope) 16-byte loads for polling: never cached in L1, single-copy atomic per 8-byte half
__device__ __forceinline__ longlong2 ld_relaxed_ll2(const void* p) {
  longlong2 v;
  asm volatile("ld.relaxed.gpu.global.v2.s64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_relaxed_u4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
// tagged fp32 pair (value, tag, value, tag): one 16-byte store, each (value, tag) half lands atomically
__device__ __forceinline__ void st_tagged2(void* p, float v0, float v1, uint32_t tag) {
  asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(__float_as_uint(v0)), "r"(tag), "r"(__float_as_uint(v1)), "r"(tag) : "memory");
}

// ---- tensor core, warp-level path of the HBM-bound decode GEMMs (the dense contractions use tcgen05: tc_gemm.cuh) ----
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
// four 8x8 b16 matrices, transposed on the way (B fragments of a row-major [k][n] tile: thread (g, t) gets rows 2t, 2t+1 of column g)
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t saddr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(saddr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint4& a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
// Two fp32 fused multiply-adds in one instruction (Blackwell FFMA2): (d0, d1) += (a0, a1) * (b0, b1), each IEEE fma.rn
__device__ __forceinline__ void ffma2(float& d0, float& d1, float a0, float a1, float b0, float b1) {
  asm("{\n\t.reg .b64 ra, rb, rc;\n\tmov.b64 ra, {%2,%3};\n\tmov.b64 rb, {%4,%5};\n\tmov.b64 rc, {%0,%1};\n\t"
      "fma.rn.f32x2 rc, ra, rb, rc;\n\tmov.b64 {%0,%1}, rc;\n\t}"
      : "+f"(d0), "+f"(d1)
      : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}
__device__ __forceinline__ float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

// ---- named barrier over the consumer warps only (the producer warp never joins) ------------------------------
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kConsumerThreads) : "memory"); }

// ---- warp / block reductions (fixed trees: deterministic) -------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// kConsumerThreads (256) consumer threads; `scratch` = 32 floats of shared memory.  All consumer threads must call.
__device__ __forceinline__ float block_sum(float v, float* scratch) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  v = warp_sum(v);
  consumer_sync();
  if (lane == 0) scratch[w] = v;
  consumer_sync();
  float r = (lane < kNW) ? scratch[lane] : 0.f;
  return warp_sum(r);
}
__device__ __forceinline__ float block_max(float v, float* scratch) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  v = warp_max(v);
  consumer_sync();
  if (lane == 0) scratch[w] = v;
  consumer_sync();
  float r = (lane < kNW) ? scratch[lane] : -INFINITY;
  return warp_max(r);
}
__device__ __forceinline__ uint32_t block_sum_u32(uint32_t v, float* scratch) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  v = __reduce_add_sync(0xffffffffu, v);
  consumer_sync();
  if (lane == 0) reinterpret_cast<uint32_t*>(scratch)[w] = v;
  consumer_sync();
  uint32_t r = (lane < kNW) ? reinterpret_cast<uint32_t*>(scratch)[lane] : 0u;
  return __reduce_add_sync(0xffffffffu, r);
}

// ---- Philox4x32-10 (oracle/philox.py is the checked definition) --------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}
__device__ __forceinline__ float philox_uniform(uint64_t seed, uint32_t seq_index, uint32_t step) {
  uint4 r = philox4x32_10(make_uint4(seq_index, step, 0u, 0u), make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
  return (float)(r.x >> 8) * (1.0f / 16777216.0f);
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ float gelu_new(float x) {
  // HF NewGELUActivation: 0.5 x (1 + tanh( sqrt(2/pi) (x + 0.044715 x^3) ))
  float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  return 0.5f * x * (1.0f + tanhf(u));
}

}  // namespace tts
