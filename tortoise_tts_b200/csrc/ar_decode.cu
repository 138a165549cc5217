// ar_decode.cu -- the head-cluster decode kernel: the sample loop for <= 16 live sequences.
//
// Same stages as the interpreter (ar_kernels.cu) but with 3 grid-wide barriers per GPT-2 block instead of 5:
// 32 thread-block clusters of 4 CTAs, cluster c owns attention head c >> 1 for 8 of the 16 rows, so
//     ln_1 -> c_attn (q,k,v columns of the head) -> [cluster barrier] -> attention of the head ->
//     [cluster barrier] -> split-K attn.c_proj (rows 64h..64h+63 of the weight) -> red.add into y
// runs without leaving the cluster (phase A); the 16 per-head partial products are summed in L2 by
// 64-bit integer atomics in 2^-32 fixed point (order-independent => deterministic) into the zeroed accumulator y (head 0 also adds the residual x and the bias).
// Phase B = ln_2 -> c_fc -> gelu_new (N split over the 128 CTAs), phase C = mlp.c_proj: x = y + . ; y = 0.
// Weights stream through the same bulk-copy ring as in the interpreter, ungated by any barrier.
// Reference path replaced: GPT2InferenceModel.forward cached branch (unified_voice.py:212-229) + HF GPT2Block.
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include "ar_gemm.cuh"
#include "ar_attn.cuh"
#include "ar_sample.cuh"
#include "ar_launch.h"

namespace tts {

constexpr int kU4PerChunk = kChunkBytes / 16;

// ---- barrier among the consumer warps of the 8 CTAs of a cluster (the producer warp never joins) ----------
__device__ __forceinline__ bool cluster_barrier(Ctrl* c, uint32_t& cphase) {
  consumer_sync();
  if (threadIdx.x < kClusterSize) mbar_arrive_remote(&c->cbar, threadIdx.x);
  bool ok = true;
  if (!mbar_try_wait_cluster(&c->cbar, cphase)) {
    const long long t0 = clock64();
    while (!mbar_try_wait_cluster(&c->cbar, cphase)) {
      if (c->abort_flag || clock64() - t0 > kWatchdogCycles) {
        c->abort_flag = 1;
        ok = false;
        break;
      }
    }
  }
  cphase ^= 1u;
  return ok;
}

// ---- producer: this CTA's weight chunks of the whole launch, in consumption order --------------------------
__device__ void dec_producer(const Params& P, Ctrl* c, uint8_t* ring, int head, int rank) {
  if ((threadIdx.x & 31) != 0) return;
  const uint64_t pol = make_evict_first_policy();
  const int cta = blockIdx.x;
  uint32_t pi = 0, slot = 0, round = 0;
  bool stopped = false;
  auto issue = [&](const uint4* src) {
    if (stopped) return;
    if (round > 0) {
      const uint32_t par = (round - 1) & 1;
      const long long t_start = clock64();
      while (!mbar_try_wait(&c->empty[slot], par)) {
        if (c->stop_flag || c->abort_flag) { stopped = true; return; }
        if (clock64() - t_start > kWatchdogCycles) { c->abort_flag = 1; stopped = true; return; }
      }
    }
    mbar_arrive_expect_tx(&c->full[slot], kChunkBytes);
    bulk_g2s(ring + (size_t)slot * kChunkBytes, src, kChunkBytes, &c->full[slot], pol);
    ++pi;
    if (++slot == (uint32_t)P.n_slots) { slot = 0; ++round; }
  };
  int ht0, hcnt;
  tile_range(P.v_tiles, P.G, 0, cta, &ht0, &hcnt);
  for (int step = 0; step < P.n_steps && !stopped; ++step) {
    if (P.skip_last_model && step == P.n_steps - 1) break;
    for (int layer = 0; layer < P.n_layers && !stopped; ++layer) {
      const LayerDev& L = P.layers[layer];
      for (int i = 0; i < 6; ++i) issue(L.w_qkv + (size_t)((i >> 1) * (kD / 8) + head * 8 + 2 * rank + (i & 1)) * kU4PerChunk);
      for (int i = 0; i < 2; ++i) issue(L.w_projk + (size_t)(head * 8 + 2 * rank + i) * kU4PerChunk);
      for (int i = 0; i < 4; ++i) issue(L.w_fc + (size_t)(cta * 4 + i) * kU4PerChunk);
      for (int i = 0; i < 4; ++i) issue(L.w_proj2 + (size_t)(cta * 4 + i) * kU4PerChunk);
    }
    for (int i = 0; i < hcnt; ++i) issue(P.w_head + (size_t)(ht0 + i) * kU4PerChunk);
  }
  // Drain: never leave the kernel with bulk copies still landing in this CTA's shared memory.
  const long long t_start = clock64();
  while (!c->stop_flag) {
    if (c->abort_flag || clock64() - t_start > kWatchdogCycles) return;
  }
  for (uint32_t k = c->stop_ci; k < pi; ++k) {
    if (!mbar_wait(&c->full[k % P.n_slots], (k / P.n_slots) & 1, &c->abort_flag)) return;
  }
}

// ---- attention of head `head` for rows rbase, rbase+1: 4 position ranges per row, merged in shared memory ----
__device__ void attn_local(const Params& P, int layer, int head, int rbase, const Smem& sm) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int rg = lane >> 3, dc = lane & 7;
  float* part = sm.red;                               // [8 warps][kPartStride]
  {
    const int row = rbase + (w >> 2), s = w & 3;
    const int len = (row < P.n_rows) ? ld_cg_i(P.row_pos + row) + 1 : 0;
    float m = -INFINITY, l = 0.f, acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    if (len > 0) {
      const int p0 = (s * len) >> 2, p1 = ((s + 1) * len) >> 2;
      if (p1 > p0) attn_unit(P, layer, row, head, p0, p1, m, l, acc);
    }
    float* pp = part + w * kPartStride;
    if (rg == 0) {
      *reinterpret_cast<float4*>(pp + 8 + dc * 8) = make_float4(acc[0], acc[1], acc[2], acc[3]);
      *reinterpret_cast<float4*>(pp + 8 + dc * 8 + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
    }
    if (lane == 0) { pp[0] = m; pp[1] = l; }
  }
  consumer_sync();
  if (w < 2) {
    const int row = rbase + w;
    if (row < kDecRows) {
      const int len = (row < P.n_rows) ? ld_cg_i(P.row_pos + row) + 1 : 0;
      float o0 = 0.f, o1 = 0.f;
      if (len > 0) {
        const float* pb = part + (4 * w) * kPartStride;
        float M = -INFINITY;
#pragma unroll
        for (int s = 0; s < 4; ++s) M = fmaxf(M, pb[s * kPartStride]);
        float L = 0.f;
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          const float wgt = exp2f(pb[s * kPartStride] - M);       // empty range: exp2(-inf) = 0
          L = fmaf(wgt, pb[s * kPartStride + 1], L);
          const float2 a = *reinterpret_cast<const float2*>(pb + s * kPartStride + 8 + 2 * lane);
          o0 = fmaf(wgt, a.x, o0);
          o1 = fmaf(wgt, a.y, o1);
        }
        const float inv = 1.0f / L;
        o0 *= inv;
        o1 *= inv;
      }
      write_ctx_pair(P, row, head * kHD + 2 * lane, o0, o1);
    }
  }
}

// ---- split-K attn.c_proj: y[rows, 256r .. 256r+256) += ctx[rows, 64h .. 64h+64) . W[64h .. 64h+64, 256r ..] ------
// rows = [row_lo, row_lo + 8): the half of the 16-row M tile this cluster owns (the other half of the A fragments
// is the sibling cluster's and is ignored).  Two 16 KiB chunks of 128 columns; warp w: n8-tiles 2w, 2w+1 of each.
__device__ void proj_splitk(const Params& P, int layer, int head, int rank, int row_lo, const Smem& sm, Pipe& pipe) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  Ctrl* c = sm.ctrl;
  const uint4* a_gl = reinterpret_cast<const uint4*>(P.ctx) + (size_t)(head * 4) * 32 + lane;   // m-tile 0, kb = 4h + j
  uint4 a[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) a[j] = ld_cg_u4(a_gl + j * 32);
  const int g = lane >> 2, t = lane & 3;
  const int hh = row_lo >> 3;                      // which half of the C fragment is ours: regs 2hh, 2hh+1
  const int row = row_lo + g;
  const bool row_ok = row < P.n_rows;
#pragma unroll
  for (int ch = 0; ch < 2; ++ch) {
    const int col0 = 256 * rank + 128 * ch + 16 * w + 2 * t;       // tile tt adds 8
    // residual + bias ride on head 0's partial product
    float2 add[2];
#pragma unroll
    for (int tt = 0; tt < 2; ++tt) {
      add[tt] = make_float2(0.f, 0.f);
      if (head == 0 && row_ok) {
        const float2 b = *reinterpret_cast<const float2*>(P.layers[layer].b_proj + col0 + 8 * tt);
        const float2 xo = ld_cg_f2(P.x + (size_t)row * kD + col0 + 8 * tt);
        add[tt] = make_float2(xo.x + b.x, xo.y + b.y);
      }
    }
    float acc[2][4];
#pragma unroll
    for (int tt = 0; tt < 2; ++tt)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[tt][j] = 0.f;
    if (mbar_wait(&c->full[pipe.slot], pipe.phase, &c->abort_flag)) {
      const uint4* bsm = reinterpret_cast<const uint4*>(sm.ring + (size_t)pipe.slot * kChunkBytes) + lane;
#pragma unroll
      for (int tt = 0; tt < 2; ++tt) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const uint4 b = bsm[(((2 * w + tt) * 2) + s) * 32];
          mma_bf16_16816(acc[tt], a[2 * s], b.x, b.y);
          mma_bf16_16816(acc[tt], a[2 * s + 1], b.z, b.w);
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&c->empty[pipe.slot]);
    pipe.advance();
    if (row_ok) {
#pragma unroll
      for (int tt = 0; tt < 2; ++tt)
      {
        long long* yp = P.y + (size_t)row * kD + col0 + 8 * tt;
        red_add_s64(yp, to_fixed((hh ? acc[tt][2] : acc[tt][0]) + add[tt].x));
        red_add_s64(yp + 1, to_fixed((hh ? acc[tt][3] : acc[tt][1]) + add[tt].y));
      }
    }
  }
}

// ---- the kernel ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 1) ar_decode_kernel(const __grid_constant__ Params P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const Smem sm = carve_smem(smem_raw, 1);
  Ctrl* c = sm.ctrl;
  const int rank = (int)cluster_ctarank();
  const int head = (int)cluster_id_x() >> 1;
  const int row_lo = ((int)cluster_id_x() & 1) * 8;
  if (threadIdx.x == 0) {
    for (int i = 0; i < P.n_slots; ++i) {
      mbar_init(&c->full[i], 1);
      mbar_init(&c->empty[i], kNW);
    }
    mbar_init(&c->cbar, kClusterSize);
    c->abort_flag = 0;
    c->ibuf[0] = 0;
    c->ibuf[1] = 0;
    c->stop_flag = 0;
    c->stop_ci = 0;
    fence_mbar_init();
  }
  __syncthreads();
  cluster_sync_all();          // every CTA's cbar is initialised before anyone arrives on it remotely

  if (threadIdx.x >= kConsumerThreads) {
    dec_producer(P, c, sm.ring, head, rank);
    return;
  }

  Pipe pipe;
  pipe.ci = 0; pipe.slot = 0; pipe.phase = 0; pipe.n_slots = P.n_slots;
  uint32_t epoch = 0, cphase = 0;
  int steps_done = 0;
  bool alive = true;
  const int n = P.n_rows;
  int ht0, hcnt;
  tile_range(P.v_tiles, P.G, 0, blockIdx.x, &ht0, &hcnt);
  Instr ins;
  ins.row0 = 0; ins.nrows = n; ins.flags = 0; ins.rot = 0; ins.pad0 = ins.pad1 = 0;
  for (int step = 0; step < P.n_steps && alive; ++step) {
    if (blockIdx.x == 0 && threadIdx.x == 0) c->ibuf[1] = (P.prof && step == P.prof_step) ? 1 : 0;
    prof_mark(P, c, 100 + OP_SAMPLE);
    alive = grid_barrier(P, c, epoch);
    if (!alive) break;
    prof_mark(P, c, 200 + OP_SAMPLE);
    ins.op = OP_SAMPLE; ins.layer = 0; ins.nrows = P.n_seqs;
    sample_stage(P, ins, sm, step);
    ins.nrows = n;
    steps_done = step + 1;
    if (P.skip_last_model && step == P.n_steps - 1) break;
    for (int layer = 0; layer < P.n_layers && alive; ++layer) {
      ins.layer = layer;
      // ---- phase A: ln_1 -> q,k,v of this head -> attention -> split-K c_proj -> y ----
      prof_mark(P, c, 100 + OP_LN_QKV);
      alive = grid_barrier(P, c, epoch);
      if (!alive) break;
      prof_mark(P, c, 200 + OP_LN_QKV);
      if (layer == 0 && ld_cg_i(P.n_active) == 0) { alive = false; break; }
      ins.op = OP_LN_QKV; ins.flags = F_ROW_RANGE; ins.pad0 = row_lo; ins.pad1 = row_lo + 8;
      gemm_tiles<1>(P, ins, sm, pipe, head * 8 + 2 * rank, 6, kD / 8, 2, P.x);
      ins.flags = 0;
      if (!(alive = cluster_barrier(c, cphase))) break;
      prof_mark(P, c, 200 + OP_ATTN);
      attn_local(P, layer, head, row_lo + 2 * rank, sm);
      if (!(alive = cluster_barrier(c, cphase))) break;
      prof_mark(P, c, 200 + OP_PROJ);
      proj_splitk(P, layer, head, rank, row_lo, sm, pipe);
      // ---- phase B: ln_2 -> c_fc -> gelu_new ----
      prof_mark(P, c, 100 + OP_LN_FC);
      alive = grid_barrier(P, c, epoch);
      if (!alive) break;
      prof_mark(P, c, 200 + OP_LN_FC);
      ins.op = OP_LN_FC; ins.flags = F_FIXED_SRC;
      gemm_tiles<1>(P, ins, sm, pipe, blockIdx.x * 4, 4, 0, 1 << 30, P.x);
      // ---- phase C: mlp.c_proj, x = y + . ----
      prof_mark(P, c, 100 + OP_PROJ2);
      alive = grid_barrier(P, c, epoch);
      if (!alive) break;
      prof_mark(P, c, 200 + OP_PROJ2);
      ins.op = OP_PROJ2; ins.flags = 0;
      gemm_tiles<1>(P, ins, sm, pipe, blockIdx.x, 1, 0, 1 << 30, nullptr);
      if (c->abort_flag) alive = false;
    }
    if (!alive) break;
    prof_mark(P, c, 100 + OP_HEAD);
    alive = grid_barrier(P, c, epoch);
    if (!alive) break;
    prof_mark(P, c, 200 + OP_HEAD);
    ins.op = OP_HEAD; ins.layer = 0;
    gemm_tiles<1>(P, ins, sm, pipe, ht0, hcnt, 0, 1 << 30, P.x);
    if (c->abort_flag) alive = false;
  }
  consumer_sync();
  if (threadIdx.x == 0) {
    if (blockIdx.x == 0) *P.steps_done = steps_done;
    c->stop_ci = pipe.ci;
    __threadfence_block();
    c->stop_flag = 1;
  }
}

// ---- attn.c_proj -> split-K chunks (layout.h projk_elem_offset); src = HF Conv1D weight [1024 in, 1024 out] ----
template <typename T>
__global__ void pack_projk_kernel(const T* __restrict__ src, uint4* __restrict__ dst) {
  uint32_t* out = reinterpret_cast<uint32_t*>(dst);
  const size_t total_words = (size_t)kProjkChunks * (kChunkBytes / 4);
  for (size_t wi = (size_t)blockIdx.x * blockDim.x + threadIdx.x; wi < total_words; wi += (size_t)gridDim.x * blockDim.x) {
    const int chunk = (int)(wi / (kChunkBytes / 4));
    const int in_chunk = (int)(wi % (kChunkBytes / 4));
    const int h = chunk >> 3, r = chunk & 7;
    const int word = in_chunk & 3, lane = (in_chunk >> 2) & 31, s = (in_chunk >> 7) & 1, tile = in_chunk >> 8;
    const int odd = word >> 1, hi = word & 1;
    const int t = lane & 3, g = lane >> 2;
    const int k = h * 64 + (2 * s + odd) * 16 + hi * 8 + 2 * t;
    const int col = 128 * r + 8 * tile + g;
    out[wi] = pack_bf16x2((float)src[(size_t)k * kD + col], (float)src[(size_t)(k + 1) * kD + col]);
  }
}

void launch_pack_projk(const void* src, int dtype, uint4* dst, cudaStream_t st) {
  if (dtype == 0)
    pack_projk_kernel<float><<<512, 256, 0, st>>>((const float*)src, dst);
  else
    pack_projk_kernel<__nv_bfloat16><<<512, 256, 0, st>>>((const __nv_bfloat16*)src, dst);
}

static cudaError_t decode_config(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attrs, int smem_bytes, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(ar_decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(kDecCtas);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = (size_t)smem_bytes;
  cfg.stream = st;
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = kClusterSize;
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributeCooperative;
  attrs[1].val.cooperative = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = 2;
  return cudaSuccess;
}

cudaError_t launch_decode(const Params& P, int smem_bytes, cudaStream_t st) {
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attrs[2];
  cudaError_t e = decode_config(cfg, attrs, smem_bytes, st);
  if (e != cudaSuccess) return e;
  return cudaLaunchKernelEx(&cfg, ar_decode_kernel, P);
}

// Clusters of kClusterSize CTAs (1 CTA / SM at this shared-memory size) that can be co-resident; the kernel needs kDecClusters.
int decode_max_clusters(int smem_bytes) {
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attrs[2];
  if (decode_config(cfg, attrs, smem_bytes, 0) != cudaSuccess) { cudaGetLastError(); return 0; }
  cfg.numAttrs = 1;   // the occupancy query takes the cluster dimension only
  int n = 0;
  cudaError_t e = cudaOccupancyMaxActiveClusters(&n, ar_decode_kernel, &cfg);
  if (e != cudaSuccess) {
    if (getenv("TTS_AR_VERBOSE")) fprintf(stderr, "[tts_ar] cudaOccupancyMaxActiveClusters: %s\n", cudaGetErrorString(e));
    cudaGetLastError();
    return 0;
  }
  return n;
}

}  // namespace tts
