// full_attn.cuh -- full (non-causal) softmax attention over whole sequences, used by the CLVP encoders (clvp.cu) and the
// conditioning encoder (cond_enc.cu).
#pragma once
#include "dev_utils.cuh"

namespace tts {

constexpr int kFaKvStride = 72;           // padded K/V tile row in shared memory (bf16 elements): conflict-free fragment loads

// ---- full (non-causal) softmax attention of one (sequence, head, 64-query block) ------------------------------------------
// qkv: bf16 [rows][3 D] = q | k | v, heads of 64 dims (D = 768: CLVP, rotary already applied by the GEMM epilogue; D = 1024:
// conditioning encoder); ctx: bf16 [rows][D].
// 4 warps, 16 query rows each, flash-style over 64-position K/V tiles staged in shared memory (cp.async, double buffered);
// S = Q K^T and O += P V on mma.m16n8k16 (fragment scheme of ar_attn.cuh attn_tile16).  Scores are scaled by 64^-0.5 in
// fp32 after the product (xtransformers.py:672), positions >= len are masked, V rows >= len are zero.
template <int D>
__global__ void __launch_bounds__(128) full_attn_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ ctx, int len) {
  __shared__ __align__(16) __nv_bfloat16 ks[2][64 * kFaKvStride];
  __shared__ __align__(16) __nv_bfloat16 vs[2][64 * kFaKvStride];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int head = blockIdx.y, seq = blockIdx.z;
  const size_t row_base = (size_t)seq * len;
  const int q0 = blockIdx.x * 64 + warp * 16;
  const bool active = q0 < len;
  const __nv_bfloat16* kbase = qkv + row_base * (3 * D) + D + head * 64;
  const __nv_bfloat16* vbase = kbase + D;

  auto load_tile = [&](int tile, int buf) {
    const int p0 = tile * 64;
#pragma unroll
    for (int pass = 0; pass < 4; ++pass) {
      const int idx = pass * 128 + threadIdx.x;
      const int r = idx >> 3, seg = idx & 7;
      __nv_bfloat16* kd = &ks[buf][r * kFaKvStride + seg * 8];
      __nv_bfloat16* vd = &vs[buf][r * kFaKvStride + seg * 8];
      if (p0 + r < len) {
        cp_async_cg16(kd, kbase + (size_t)(p0 + r) * (3 * D) + seg * 8);
        cp_async_cg16(vd, vbase + (size_t)(p0 + r) * (3 * D) + seg * 8);
      } else {
        *reinterpret_cast<uint4*>(kd) = make_uint4(0u, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(vd) = make_uint4(0u, 0u, 0u, 0u);
      }
    }
  };

  uint4 qa[4];
  {
    const int ra = min(q0 + g, len - 1), rb = min(q0 + g + 8, len - 1);       // rows past the end: clamped, never stored
    const __nv_bfloat16* qpa = qkv + (row_base + ra) * (3 * D) + head * 64 + 2 * t;
    const __nv_bfloat16* qpb = qkv + (row_base + rb) * (3 * D) + head * 64 + 2 * t;
#pragma unroll
    for (int kb = 0; kb < 4; ++kb) {
      qa[kb].x = *reinterpret_cast<const uint32_t*>(qpa + kb * 16);
      qa[kb].y = *reinterpret_cast<const uint32_t*>(qpb + kb * 16);
      qa[kb].z = *reinterpret_cast<const uint32_t*>(qpa + kb * 16 + 8);
      qa[kb].w = *reinterpret_cast<const uint32_t*>(qpb + kb * 16 + 8);
    }
  }
  float o[8][4];
#pragma unroll
  for (int j = 0; j < 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;      // rows g and g + 8
  const float sc = 0.125f * 1.4426950408889634f;                  // 64^-0.5 * log2(e)
  const int n_tiles = (len + 63) >> 6;
  load_tile(0, 0);
  for (int tile = 0; tile < n_tiles; ++tile) {
    cp_async_wait_all();
    __syncthreads();                                              // tile landed; everyone is done with the other buffer
    if (tile + 1 < n_tiles) load_tile(tile + 1, (tile + 1) & 1);
    if (!active) continue;
    const __nv_bfloat16* kt = ks[tile & 1];
    const __nv_bfloat16* vt = vs[tile & 1];
    const int chunks = min(4, (len - tile * 64 + 15) >> 4);
    for (int c = 0; c < chunks; ++c) {
      const int c0 = tile * 64 + c * 16;
      // K: B fragments of S = Q K^T (k = dim, n = position): K is [position][dim] in shared memory, i.e. already "column-major"
      // for the mma -> plain ldmatrix; one x4 = (positions 0-7 | 8-15) x (dims kb*16 + 0-7 | 8-15)
      uint32_t kf[2][4][2];
      {
        const int mat = lane >> 3, r = lane & 7;
        const uint32_t kaddr = smem_u32(kt + (c * 16 + (mat >> 1) * 8 + r) * kFaKvStride + (mat & 1) * 8);
#pragma unroll
        for (int kb = 0; kb < 4; ++kb) ldmatrix_x4(kaddr + kb * 32, kf[0][kb][0], kf[0][kb][1], kf[1][kb][0], kf[1][kb][1]);
      }
      // V: B fragments of O += P V (k = position, n = dim): V is [position][dim] = row-major [k][n] -> ldmatrix.trans;
      // one x4 = (positions 0-7 | 8-15) x (dims jd*8 + 0-7 | + 8-15)
      uint32_t vf[8][2];
      {
        const int mat = lane >> 3, r = lane & 7;
        const uint32_t vaddr = smem_u32(vt + (c * 16 + (mat & 1) * 8 + r) * kFaKvStride + (mat >> 1) * 8);
#pragma unroll
        for (int jd = 0; jd < 8; jd += 2) ldmatrix_x4_trans(vaddr + jd * 16, vf[jd][0], vf[jd][1], vf[jd + 1][0], vf[jd + 1][1]);
      }
      float s[2][4];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
        for (int kb = 0; kb < 4; ++kb) mma_bf16_16816(s[j], qa[kb], kf[j][kb][0], kf[j][kb][1]);
      }
      float cm0 = -INFINITY, cm1 = -INFINITY;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int p = c0 + 8 * j + 2 * t;
        s[j][0] = (p < len) ? s[j][0] * sc : -INFINITY;
        s[j][1] = (p + 1 < len) ? s[j][1] * sc : -INFINITY;
        s[j][2] = (p < len) ? s[j][2] * sc : -INFINITY;
        s[j][3] = (p + 1 < len) ? s[j][3] * sc : -INFINITY;
        cm0 = fmaxf(cm0, fmaxf(s[j][0], s[j][1]));
        cm1 = fmaxf(cm1, fmaxf(s[j][2], s[j][3]));
      }
      cm0 = fmaxf(cm0, __shfl_xor_sync(0xffffffffu, cm0, 1));
      cm0 = fmaxf(cm0, __shfl_xor_sync(0xffffffffu, cm0, 2));
      cm1 = fmaxf(cm1, __shfl_xor_sync(0xffffffffu, cm1, 1));
      cm1 = fmaxf(cm1, __shfl_xor_sync(0xffffffffu, cm1, 2));
      const float mn0 = fmaxf(m0, cm0), mn1 = fmaxf(m1, cm1);     // finite: every chunk visited holds a position < len
      const float f0 = exp2f(m0 - mn0), f1 = exp2f(m1 - mn1);
      m0 = mn0; m1 = mn1;
      l0 *= f0; l1 *= f1;
#pragma unroll
      for (int jd = 0; jd < 8; ++jd) { o[jd][0] *= f0; o[jd][1] *= f0; o[jd][2] *= f1; o[jd][3] *= f1; }
      float pr[2][4];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        pr[j][0] = exp2f(s[j][0] - mn0); pr[j][1] = exp2f(s[j][1] - mn0);
        pr[j][2] = exp2f(s[j][2] - mn1); pr[j][3] = exp2f(s[j][3] - mn1);
        l0 += pr[j][0] + pr[j][1];
        l1 += pr[j][2] + pr[j][3];
      }
      const uint4 pa = make_uint4(pack_bf16x2(pr[0][0], pr[0][1]), pack_bf16x2(pr[0][2], pr[0][3]),
                                  pack_bf16x2(pr[1][0], pr[1][1]), pack_bf16x2(pr[1][2], pr[1][3]));
#pragma unroll
      for (int jd = 0; jd < 8; ++jd) mma_bf16_16816(o[jd], pa, vf[jd][0], vf[jd][1]);
    }
  }
  if (!active) return;
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float i0 = 1.0f / l0, i1 = 1.0f / l1;
#pragma unroll
  for (int jd = 0; jd < 8; ++jd) {
    const int col = head * 64 + 8 * jd + 2 * t;
    if (q0 + g < len) *reinterpret_cast<uint32_t*>(ctx + (row_base + q0 + g) * D + col) = pack_bf16x2(o[jd][0] * i0, o[jd][1] * i0);
    if (q0 + g + 8 < len) *reinterpret_cast<uint32_t*>(ctx + (row_base + q0 + g + 8) * D + col) = pack_bf16x2(o[jd][2] * i1, o[jd][3] * i1);
  }
}

}  // namespace tts
