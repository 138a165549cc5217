// tc_gemm.cuh -- the dense contraction of the prompt-shaped passes on the 5th-generation tensor cores.
//
//   D[M, N] (fp32, TMEM) = A[M, K] (bf16 row-major activations) . Wt[N, K]^T (bf16, K-major weight image)
//
// One persistent CTA per SM, 6 warps: warp 0 = TMA producer (cp.async.bulk.tensor.2d, 128-byte swizzle, completion on an
// mbarrier ring), warp 1 = MMA issuer (one elected thread issues tcgen05.mma.cta_group::1.kind::f16, 128 x BN x 16 per
// instruction, accumulators in TMEM, double buffered across output tiles), warps 2..5 = epilogue (tcgen05.ld of the
// accumulator lanes they own -> registers -> fused epilogue -> global).  The MMA of tile i+1 overlaps the epilogue of tile i.
// Used where the rows make the GEMMs real dense contractions: the prompt pass (reference GPT2InferenceModel.forward
// prefill branch, unified_voice.py:203-211) and the latent re-forward (UnifiedVoice.forward, unified_voice.py:544-600);
// the decode GEMMs (<= 64 rows) are HBM-bound and stay on the weight-streaming kernels.
#pragma once
#include <cuda.h>
#include "dev_utils.cuh"

namespace tts {

constexpr int kTcBM = 128;   // rows per output tile (UMMA M, cta_group::1)
constexpr int kTcBK = 64;    // k extent of one pipeline stage: 64 bf16 = one 128-byte swizzle row

enum TcEpi : int {
  TC_EPI_STORE = 0,   // out_f32[row][col] = acc + bias          (mel head logits; bring-up test)
  TC_EPI_RESID = 1,   // x[row][col] += acc + bias                (attn.c_proj, mlp.c_proj)
  TC_EPI_GELU = 2,    // out_bf16[row][col] = gelu_new(acc + bias) (c_fc)
  TC_EPI_QKV = 3,     // q (fp32, pre-scaled 1/8) and the K/V page slots of the row's position (c_attn)
  TC_EPI_ROT = 4,     // CLVP q|k|v: out_bf16[row][col] = acc, rotary embedding on the first 32 dims of every 64-dim head
  TC_EPI_GLU = 5      // CLVP feed-forward: out_bf16[row][u] = (acc_x + b) * gelu_erf(acc_gate + b); weight rows interleaved
                      // per 32 hidden units (x block, gate block) so both halves of a unit sit in one tile (clvp.cu)
};

struct TcGemmArgs {
  int M, N, K;                 // valid rows / columns; K multiple of 64
  const float* bias;           // [N] (zero-padded to the tile edge by the caller, or nullptr)
  float* out_f32;              // STORE / RESID
  int ldo;                     // row stride of out_f32 (floats)
  __nv_bfloat16* out_bf16;     // GELU: [M][N]
  // QKV
  float* q;                    // [M][1024]
  __nv_bfloat16* kv;           // page pool
  const int* row_utt;          // [M]
  const int* row_pos;          // [M]
  const int* prefix_pt;        // [utts][kMaxPagesPerSeg]
  int n_pages, layer;
  int* watchdog;               // device flag: a wait timed out
  // ROT
  const float* rot_tab;        // [seq_len][32]: cos(pos * inv_freq[i]) for i < 16, then sin
  int seq_len;                 // rows per sequence (position of a row = row % seq_len)
};

// ---- PTX wrappers ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem] (+)= A[smem desc] . B[smem desc]; issued by ONE thread on behalf of the CTA
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst_smem, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(smem_u32(dst_smem)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread i of the warp receives TMEM lane (lane base of the address) + i
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor of a K-major bf16 tile stored as [rows][64] with the 128-byte swizzle (what a
// cp.async.bulk.tensor box of {64, rows} with CU_TENSOR_MAP_SWIZZLE_128B writes): rows are 128 bytes, 8-row groups
// are 1024 bytes apart (stride byte offset), descriptor version 1 (sm_100), layout type 2 = SWIZZLE_128B.
// Advancing by one UMMA_K (16 bf16 = 32 bytes) inside the swizzle row adds 2 to the start-address field.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3fffu);        // start address, 16-byte units
  d |= (uint64_t)1u << 16;                            // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024u >> 4) << 32;                  // stride byte offset
  d |= (uint64_t)1u << 46;                            // descriptor version (Blackwell)
  d |= (uint64_t)2u << 61;                            // SWIZZLE_128B
  return d;
}
// Instruction descriptor, kind::f16: fp32 accumulate, bf16 x bf16, both operands K-major, M x N.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ bool tc_wait(uint64_t* bar, uint32_t parity, int* watchdog) {
  if (mbar_try_wait(bar, parity)) return true;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > kWatchdogCycles) {
      *reinterpret_cast<volatile int*>(watchdog) = 1;
      return false;
    }
  }
  return true;
}

template <int BN>
struct TcSmem {
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr uint32_t kABytes = kTcBM * kTcBK * 2;
  static constexpr uint32_t kBBytes = BN * kTcBK * 2;
  static constexpr uint32_t kStageBytes = kABytes + kBBytes;
  static constexpr uint32_t kBarOffset = kStages * kStageBytes;
  static constexpr uint32_t kBytes = kBarOffset + 256 + 1024;   // + barriers + alignment slack
};

// ---- epilogue of one thread: row `row`, 32 consecutive columns starting at `col` ------------------------------------
template <int EPI>
__device__ __forceinline__ void tc_epilogue(const TcGemmArgs& g, int row, int col, const uint32_t (&r)[32]) {
  if (row >= g.M || col >= g.N) return;
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
  if (g.bias) {
    if (col + 32 <= g.N) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const float4 b = *reinterpret_cast<const float4*>(g.bias + col + i);
        v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
      }
    } else {   // ragged last column chunk (mel head: N = 8194)
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (col + i < g.N) v[i] += g.bias[col + i];
    }
  }
  if (EPI == TC_EPI_STORE) {
    float* o = g.out_f32 + (size_t)row * g.ldo + col;
    if (col + 32 <= g.N && (g.ldo & 3) == 0) {   // 16-byte aligned rows (the product's logits rows are: v_stride % 8 == 0)
#pragma unroll
      for (int i = 0; i < 32; i += 4) *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
    } else {
      for (int i = 0; i < 32 && col + i < g.N; ++i) o[i] = v[i];
    }
  } else if (EPI == TC_EPI_RESID) {
    float* o = g.out_f32 + (size_t)row * g.ldo + col;
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      const float4 x = *reinterpret_cast<const float4*>(o + i);
      *reinterpret_cast<float4*>(o + i) = make_float4(x.x + v[i], x.y + v[i + 1], x.z + v[i + 2], x.w + v[i + 3]);
    }
  } else if (EPI == TC_EPI_GELU) {
    __nv_bfloat16* o = g.out_bf16 + (size_t)row * g.N + col;
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
      uint4 pk;
      pk.x = pack_bf16x2(gelu_new(v[i]), gelu_new(v[i + 1]));
      pk.y = pack_bf16x2(gelu_new(v[i + 2]), gelu_new(v[i + 3]));
      pk.z = pack_bf16x2(gelu_new(v[i + 4]), gelu_new(v[i + 5]));
      pk.w = pack_bf16x2(gelu_new(v[i + 6]), gelu_new(v[i + 7]));
      *reinterpret_cast<uint4*>(o + i) = pk;
    }
  } else if (EPI == TC_EPI_ROT) {
    // reference xtransformers.py:279-290, 660-664: rotary dims = the first 32 of each head, dim i pairs with dim i + 16,
    // applied to q, k AND v; a 32-column chunk starting at a multiple of 64 is exactly the rotary part of one head
    if (g.rot_tab && (col & 63) == 0) {   // rot_tab == nullptr: plain bf16 store (conditioning encoder q|k|v)
      const float* rt = g.rot_tab + (size_t)(row % g.seq_len) * 32;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float c = rt[i], sn = rt[16 + i];
        const float a = v[i], b = v[i + 16];
        v[i] = a * c - b * sn;
        v[i + 16] = b * c + a * sn;
      }
    }
    __nv_bfloat16* o = g.out_bf16 + (size_t)row * g.N + col;
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
      uint4 pk;
      pk.x = pack_bf16x2(v[i], v[i + 1]);
      pk.y = pack_bf16x2(v[i + 2], v[i + 3]);
      pk.z = pack_bf16x2(v[i + 4], v[i + 5]);
      pk.w = pack_bf16x2(v[i + 6], v[i + 7]);
      *reinterpret_cast<uint4*>(o + i) = pk;
    }
  } else if (EPI == TC_EPI_GLU) {
    // handled by tc_epilogue_glu (two chunks at a time)
  } else {   // TC_EPI_QKV: columns [0,1024) query, [1024,2048) key, [2048,3072) value; 32 columns never straddle a head
    const int part = col >> 10, cc = col & (kD - 1);
    if (part == 0) {
      float* o = g.q + (size_t)row * kD + cc;
#pragma unroll
      for (int i = 0; i < 32; i += 4)
        *reinterpret_cast<float4*>(o + i) = make_float4(v[i] * 0.125f, v[i + 1] * 0.125f, v[i + 2] * 0.125f, v[i + 3] * 0.125f);
    } else {
      const int pos = g.row_pos[row];
      if (pos < 0) return;
      const int page = g.prefix_pt[g.row_utt[row] * kMaxPagesPerSeg + (pos >> 5)];
      __nv_bfloat16* o = g.kv + kv_row_offset(g.layer, g.n_pages, page, part - 1, cc >> 6, pos & 31) + (cc & 63);
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        uint4 pk;
        pk.x = pack_bf16x2(v[i], v[i + 1]);
        pk.y = pack_bf16x2(v[i + 2], v[i + 3]);
        pk.z = pack_bf16x2(v[i + 4], v[i + 5]);
        pk.w = pack_bf16x2(v[i + 6], v[i + 7]);
        *reinterpret_cast<uint4*>(o + i) = pk;
      }
    }
  }
}

// GLU epilogue of one thread: x chunk r0 (columns col .. col + 31 of the interleaved image) and its gate chunk r1 (col + 32 ..):
// hidden units (col >> 1) .. + 31.  exact (erf) GELU as nn.GELU() (reference xtransformers.py:431-440, 455).
__device__ __forceinline__ void tc_epilogue_glu(const TcGemmArgs& g, int row, int col, const uint32_t (&r0)[32], const uint32_t (&r1)[32]) {
  if (row >= g.M || col >= g.N) return;
  __nv_bfloat16* o = g.out_bf16 + (size_t)row * (g.N >> 1) + (col >> 1);
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    float h[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float x = __uint_as_float(r0[i + k]) + g.bias[col + i + k];
      const float t = __uint_as_float(r1[i + k]) + g.bias[col + 32 + i + k];
      h[k] = x * (0.5f * t * (1.0f + erff(t * 0.70710678118654752f)));
    }
    uint4 pk;
    pk.x = pack_bf16x2(h[0], h[1]); pk.y = pack_bf16x2(h[2], h[3]); pk.z = pack_bf16x2(h[4], h[5]); pk.w = pack_bf16x2(h[6], h[7]);
    *reinterpret_cast<uint4*>(o + i) = pk;
  }
}

template <int BN, int EPI>
__global__ void __launch_bounds__(192, 1) tc_gemm_kernel(const __grid_constant__ CUtensorMap tma_a,
                                                         const __grid_constant__ CUtensorMap tma_b, const TcGemmArgs g) {
  using S = TcSmem<BN>;
  extern __shared__ uint8_t tc_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(tc_smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + S::kBarOffset);
  uint64_t* empty = full + S::kStages;
  uint64_t* tmem_full = empty + S::kStages;     // [2]
  uint64_t* tmem_empty = tmem_full + 2;         // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tiles = (g.M + kTcBM - 1) / kTcBM, n_tiles = (g.N + BN - 1) / BN;
  const int total = m_tiles * n_tiles, k_blocks = g.K / kTcBK;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    for (int i = 0; i < S::kStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 4); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 2 * BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {   // ===== TMA producer =====
      uint32_t stage = 0, phase = 0;
      bool ok = true;
      for (int tile = blockIdx.x; tile < total && ok; tile += gridDim.x) {
        const int m_idx = (tile / n_tiles) * kTcBM, n_idx = (tile % n_tiles) * BN;
        for (int kb = 0; kb < k_blocks; ++kb) {
          if (!(ok = tc_wait(&empty[stage], phase ^ 1u, g.watchdog))) break;
          uint8_t* sa = smem + stage * S::kStageBytes;
          mbar_arrive_expect_tx(&full[stage], S::kStageBytes);
          tma_load_2d(sa, &tma_a, &full[stage], kb * kTcBK, m_idx);
          tma_load_2d(sa + S::kABytes, &tma_b, &full[stage], kb * kTcBK, n_idx);
          if (++stage == S::kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {   // ===== MMA issuer =====
      constexpr uint32_t idesc = umma_idesc_bf16(kTcBM, BN);
      uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
      bool ok = true;
      for (int tile = blockIdx.x; tile < total && ok; tile += gridDim.x) {
        if (!(ok = tc_wait(&tmem_empty[acc], acc_phase ^ 1u, g.watchdog))) break;   // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < k_blocks; ++kb) {
          if (!(ok = tc_wait(&full[stage], phase, g.watchdog))) break;
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * S::kStageBytes);
          const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + S::kABytes);
#pragma unroll
          for (int k = 0; k < kTcBK / 16; ++k) umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (uint32_t)((kb | k) != 0));
          umma_commit(&empty[stage]);            // frees the smem stage once these MMAs have read it
          if (++stage == S::kStages) { stage = 0; phase ^= 1u; }
        }
        umma_commit(&tmem_full[acc]);            // accumulator complete -> epilogue
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else {             // ===== epilogue: warps 2..5, TMEM lanes 32 * (warp % 4) .. + 31 =====
    const int ew = warp & 3;
    uint32_t acc = 0, acc_phase = 0;
    bool ok = true;
    for (int tile = blockIdx.x; tile < total && ok; tile += gridDim.x) {
      const int m_idx = (tile / n_tiles) * kTcBM, n_idx = (tile % n_tiles) * BN;
      ok = tc_wait(&tmem_full[acc], acc_phase, g.watchdog);
      ok = __all_sync(0xffffffffu, ok);
      if (!ok) break;
      tc_fence_after();
      const int row = m_idx + ew * 32 + lane;
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(ew * 32) << 16);
      if (EPI == TC_EPI_GLU) {
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 64) {
          uint32_t r0[32], r1[32];
          tmem_ld_32x32(taddr + c0, r0);
          tmem_ld_32x32(taddr + c0 + 32, r1);
          tc_epilogue_glu(g, row, n_idx + c0, r0, r1);
        }
      } else {
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
          uint32_t r[32];
          tmem_ld_32x32(taddr + c0, r);
          tc_epilogue<EPI>(g, row, n_idx + c0, r);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

}  // namespace tts
