// ar_core.cuh -- control block, weight-stream pipeline, grid barrier and LayerNorm staging of the
// persistent interpreter kernel.
#pragma once
#include "ar_params.h"
#include "dev_utils.cuh"

namespace tts {

constexpr int kMaxSlots = 12;
constexpr int kLnSplitMinRows = 16;   // interpreter: LayerNorm as its own grid-wide phase above this many rows

struct Ctrl {
  uint64_t full[kMaxSlots];    // producer -> consumers: chunk landed (tx-count barrier)
  uint64_t empty[kMaxSlots];   // consumers -> producer: slot free again (16 warp arrivals)
  uint64_t cbar;               // ar_decode.cu: the head's context rows of all 4 CTAs have landed (st.async transaction count)
  uint64_t qbar;               // ar_decode.cu: q and the new K/V rows of this CTA's attention rows have landed
  uint64_t xbar[2];            // ar_decode.cu: the 4 K-partials of this CTA's mlp.c_proj tile have landed (row groups alternate)
  float scratch[32];
  volatile int abort_flag;     // a wait timed out somewhere: fall through everything
  volatile int stop_flag;      // consumers are done (normal end or early exit)
  volatile uint32_t stop_ci;   // chunks the consumers actually consumed
  int ibuf[8];
};

struct Smem {
  Ctrl* ctrl;
  uint8_t* a_rm;     // LN'd A operand, bf16 row-major, rows padded to kARowBytes (ldmatrix source)
  float* red;        // cross-warp reduction buffer / sampler scratch / attention plan copy
  uint8_t* ring;     // n_slots * kChunkBytes of streamed weights
};

__device__ __forceinline__ Smem carve_smem(uint8_t* base, int mt_max) {
  Smem s;
  s.ctrl = reinterpret_cast<Ctrl*>(base);
  s.a_rm = base + kSmemCtrlBytes;
  s.red = reinterpret_cast<float*>(s.a_rm + smem_a_bytes(mt_max));
  s.ring = reinterpret_cast<uint8_t*>(s.red) + kSmemRedBytes;
  return s;
}

// Bring-up profiling: CTA 0 / thread 0 appends (id, ns) marks while ctrl->ibuf[1] is set (see ar_kernels.cu).
__device__ __forceinline__ void prof_mark(const Params& P, Ctrl* c, int id) {
  if (P.prof && blockIdx.x == 0 && threadIdx.x == 0 && c->ibuf[1]) {
    const int k = c->ibuf[0]++;
    if (k < 2048) {
      P.prof[1024 + 2 * k] = (unsigned long long)id;
      P.prof[1024 + 2 * k + 1] = globaltimer_ns();
    }
  }
}

// Consumer-side view of the weight stream: every consumer warp consumes every chunk, in order.
struct Pipe {
  uint32_t ci;      // chunks consumed so far
  uint32_t slot;    // ci % n_slots
  uint32_t phase;   // (ci / n_slots) & 1
  int n_slots;
  __device__ __forceinline__ void advance() {
    ++ci;
    if (++slot == (uint32_t)n_slots) { slot = 0; phase ^= 1; }
  }
};

// What a GEMM-type instruction streams.
struct GemmDesc {
  const uint4* w;     // packed chunks
  int K;              // 1024 or 4096
  int n_tiles;        // N / 8 (head: ceil(V/8))
};

__device__ __forceinline__ bool gemm_desc(const Params& P, const Instr& ins, GemmDesc& d) {
  switch (ins.op) {
    case OP_LN_QKV: d.w = P.layers[ins.layer].w_qkv; d.K = kD; d.n_tiles = kD3 / 8; return true;
    case OP_PROJ:   d.w = P.layers[ins.layer].w_proj; d.K = kD; d.n_tiles = kD / 8; return true;
    case OP_LN_FC:  d.w = P.layers[ins.layer].w_fc; d.K = kD; d.n_tiles = kD4 / 8; return true;
    case OP_PROJ2:  d.w = P.layers[ins.layer].w_proj2; d.K = kD4; d.n_tiles = kD / 8; return true;
    case OP_HEAD:   d.w = P.w_head; d.K = kD; d.n_tiles = P.v_tiles; return true;
    default: return false;
  }
}

// The per-step instruction sequence, generated on the fly (no table in global memory on the critical path):
//   [SAMPLE]  { LN_QKV g.. | ATTN | PROJ g.. | LN_FC g.. | PROJ2 g.. } x layers  [HEAD g.. | LATENT]
// Row groups of kRowGroup rows; the first instruction of every op carries the grid barrier.
struct ProgIter {
  int layer, op, group, rot, phase;   // phase 0: sample, 1: layers, 2: tail, 3: done
  int n_groups, n_head_groups;
  int row_group;   // rows per GEMM instruction: 128 (8 m-tiles, one per warp) when the A operand comes from L2 (LayerNorm
                   // phase active), else kRowGroup = 64 (the LayerNorm'd rows must fit the shared-memory A buffer)
  __device__ __forceinline__ void reset(const Params& P) {
    phase = P.has_sample ? 0 : 1;
    layer = 0; op = OP_LN_QKV; group = 0; rot = 0;
    row_group = (P.xn != nullptr && P.n_rows > kLnSplitMinRows) ? 2 * kRowGroup : kRowGroup;
    n_groups = (P.n_rows + row_group - 1) / row_group;
    n_head_groups = (P.n_head_rows + kRowGroup - 1) / kRowGroup;
  }
  __device__ __forceinline__ bool next(const Params& P, Instr& in) {
    in.pad0 = in.pad1 = 0;
    if (phase == 0) {
      in.op = OP_SAMPLE; in.layer = 0; in.row0 = 0; in.nrows = P.n_seqs; in.flags = F_BARRIER_BEFORE; in.rot = 0;
      phase = 1;
      return true;
    }
    if (phase == 1) {
      in.op = op; in.layer = layer; in.rot = rot;
      if (op == OP_ATTN) {
        in.row0 = 0; in.nrows = P.n_rows; in.flags = F_BARRIER_BEFORE;
        op = OP_PROJ;
        return true;
      }
      in.row0 = group * row_group;
      in.nrows = min(row_group, P.n_rows - in.row0);
      in.flags = (group == 0) ? F_BARRIER_BEFORE : 0;
      if (group == 0 && op == OP_LN_QKV && layer == 0 && P.has_sample) in.flags |= F_CHECK_EXIT | F_PLAN;
      if (++group == n_groups) {
        group = 0;
        const int nt = (op == OP_LN_QKV) ? kD3 / 8 : (op == OP_LN_FC ? kD4 / 8 : kD / 8);
        rot = next_rot(nt, P.G, rot);
        if (op == OP_PROJ2) {
          op = OP_LN_QKV;
          if (++layer == P.n_layers) phase = 2;
        } else {
          op = op + 1;
        }
      }
      return true;
    }
    if (phase == 2) {
      if (P.tail == 1) {
        in.op = OP_HEAD; in.layer = 0; in.rot = rot;
        in.row0 = group * kRowGroup;
        in.nrows = min(kRowGroup, P.n_head_rows - in.row0);
        in.flags = (group == 0 ? F_BARRIER_BEFORE : 0) | (P.decode_mode ? 0 : F_ROW_MAP);
        if (++group == n_head_groups) phase = 3;
        return true;
      }
      if (P.tail == 2) {
        in.op = OP_LATENT; in.layer = 0; in.rot = rot; in.row0 = 0; in.nrows = P.n_rows; in.flags = F_BARRIER_BEFORE;
        phase = 3;
        return true;
      }
      phase = 3;
    }
    return false;
  }
};

// Is instruction `pc` of step `step` executed?  (generate() does not run the model after the last token.)
__device__ __forceinline__ bool instr_live(const Params& P, int step, const Instr& ins) {
  return !(P.skip_last_model && step == P.n_steps - 1 && ins.op != OP_SAMPLE);
}

// ---- producer warp: streams this CTA's weight chunks for the whole launch, ahead of the barriers --------
static __device__ void producer_loop(const Params& P, Ctrl* c, uint8_t* ring) {
  if ((threadIdx.x & 31) != 0) return;
  const uint64_t pol = make_evict_first_policy();
  const int cta = blockIdx.x;
  uint32_t pi = 0, slot = 0, round = 0;
  bool stopped = false;
  for (int step = 0; step < P.n_steps && !stopped; ++step) {
    ProgIter it;
    it.reset(P);
    Instr ins;
    while (!stopped && it.next(P, ins)) {
      GemmDesc gd;
      if (!instr_live(P, step, ins) || !gemm_desc(P, ins, gd)) continue;
      int t0, cnt;
      tile_range(gd.n_tiles, P.G, ins.rot, cta, &t0, &cnt);
      const int kq = gd.K / kChunkK;
      const uint4* src = gd.w + (size_t)t0 * kq * (kChunkBytes / 16);
      for (int j = 0; j < cnt * kq; ++j) {
        if (round > 0) {
          const uint32_t par = (round - 1) & 1;
          long long t_start = clock64();
          while (!mbar_try_wait(&c->empty[slot], par)) {
            if (c->stop_flag || c->abort_flag) { stopped = true; break; }
            if (clock64() - t_start > kWatchdogCycles) { c->abort_flag = 1; stopped = true; break; }
          }
          if (stopped) break;
        }
        mbar_arrive_expect_tx(&c->full[slot], kChunkBytes);
        bulk_g2s(ring + (size_t)slot * kChunkBytes, src + (size_t)j * (kChunkBytes / 16), kChunkBytes, &c->full[slot], pol);
        ++pi;
        if (++slot == (uint32_t)P.n_slots) { slot = 0; ++round; }
      }
    }
  }
  // Drain: never leave the kernel with bulk copies still landing in this CTA's shared memory.
  long long t_start = clock64();
  while (!c->stop_flag) {
    if (c->abort_flag || clock64() - t_start > kWatchdogCycles) return;
  }
  for (uint32_t k = c->stop_ci; k < pi; ++k) {
    if (!mbar_wait(&c->full[k % P.n_slots], (k / P.n_slots) & 1, &c->abort_flag)) return;
  }
}

// ---- grid-wide barrier over the consumer threads of all CTAs --------------------------------------------
__device__ __forceinline__ bool grid_barrier(const Params& P, Ctrl* c, uint32_t& epoch) {
  consumer_sync();
  ++epoch;
  if (threadIdx.x == 0) {
    // bar.sync above orders the CTA's writes before this release (cumulativity); the poll is relaxed and one
    // acquire fence follows it, so the L1 invalidate happens once per barrier instead of once per poll.
    red_release_add_u32(P.bar_counter, 1u);
    const uint32_t target = epoch * (uint32_t)P.G;
    long long t0 = clock64();
    uint32_t spins = 0;
    while (ld_relaxed_u32(P.bar_counter) < target) {
      if ((++spins & 255u) == 0) {
        if (c->abort_flag || __ldcg(P.watchdog) || clock64() - t0 > kWatchdogCycles) {
          c->abort_flag = 1;
          *reinterpret_cast<volatile int*>(P.watchdog) = 1;
          break;
        }
      }
    }
    fence_acq_rel_gpu();
  }
  consumer_sync();
  return c->abort_flag == 0;
}

// ---- LayerNorm of up to 64 rows into the shared bf16 A operand ------------------------------------------------
// Warp w normalises rows w, w+8, ... two at a time (all loads of both rows are issued before the first
// reduction; the gain/bias lines are prefetched into L1 up front, so a 16-row stage costs one L2 round trip).
// fp32 statistics, two-pass, eps 1e-5 = nn.LayerNorm default used by HF GPT2Block and by final_norm.
// kDouble: ln_f then final_norm (reference quirk Q1, unified_voice.py:106,239); CTA 0 also publishes the fp32
// result as the step latent (stream_generator.py:1172).  Rows >= nrows of the last M-tile are zero-filled.
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

__device__ __forceinline__ void ln_load_row(const Params& P, const Instr& ins, const float* src, int xrow, int lane, float4 (&v)[8]) {
  if (ins.flags & F_FIXED_SRC) {
    const long long* yp = P.y + (size_t)xrow * kD;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const longlong2 a = ld_cg_ll2(yp + 128 * j + 4 * lane), b = ld_cg_ll2(yp + 128 * j + 4 * lane + 2);
      v[j] = make_float4(from_fixed(a.x), from_fixed(a.y), from_fixed(b.x), from_fixed(b.y));
    }
  } else {
    const float* xp = src + (size_t)xrow * kD;
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = ld_cg_f4(xp + 128 * j + 4 * lane);
  }
}

__device__ __forceinline__ void ln_normalise(float4 (&v)[8], const float* g, const float* b, int lane) {
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
  const float mean = warp_sum(s) * (1.0f / kD);
  float ss = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float a = v[j].x - mean, bq = v[j].y - mean, c2 = v[j].z - mean, d = v[j].w - mean;
    ss += (a * a + bq * bq) + (c2 * c2 + d * d);
  }
  const float rstd = rsqrtf(warp_sum(ss) * (1.0f / kD) + 1e-5f);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float4 gg = *reinterpret_cast<const float4*>(g + 128 * j + 4 * lane);
    const float4 bb = *reinterpret_cast<const float4*>(b + 128 * j + 4 * lane);
    v[j].x = (v[j].x - mean) * rstd * gg.x + bb.x;
    v[j].y = (v[j].y - mean) * rstd * gg.y + bb.y;
    v[j].z = (v[j].z - mean) * rstd * gg.z + bb.z;
    v[j].w = (v[j].w - mean) * rstd * gg.w + bb.w;
  }
}

template <bool kDouble>
__device__ __forceinline__ void ln_finish_row(const Params& P, const Instr& ins, uint8_t* dst, int lrow, int lane, float4 (&v)[8],
                                              const float* g1, const float* b1, const float* g2, const float* b2) {
  ln_normalise(v, g1, b1, lane);
  if (kDouble) {
    ln_normalise(v, g2, b2, lane);
    if (ins.op == OP_HEAD && blockIdx.x == 0) {
      float* lp = P.cur_latent + (size_t)lrow * kD;
#pragma unroll
      for (int j = 0; j < 8; ++j) *reinterpret_cast<float4*>(lp + 128 * j + 4 * lane) = v[j];
    }
    if (ins.op == OP_LATENT) {
      const int drow = P.latent_dst[lrow];
      if (drow >= 0) {
        float* lp = P.latent_out + (size_t)drow * kD;
#pragma unroll
        for (int j = 0; j < 8; ++j) *reinterpret_cast<float4*>(lp + 128 * j + 4 * lane) = v[j];
      }
    }
  }
  if (ins.op != OP_LATENT) {
#pragma unroll
    for (int j = 0; j < 8; ++j)
      *reinterpret_cast<uint2*>(dst + (128 * j + 4 * lane) * 2) =
          make_uint2(pack_bf16x2(v[j].x, v[j].y), pack_bf16x2(v[j].z, v[j].w));
  }
}

template <bool kDouble>
__device__ void ln_rows_to_smem(const Params& P, const Instr& ins, uint8_t* a_rm, const float* src, const float* g1, const float* b1,
                                const float* g2, const float* b2, int mt_count) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#ifdef TTS_LN_PREFETCH
  prefetch_l1(g1 + 32 * lane);
  prefetch_l1(b1 + 32 * lane);
  if (kDouble) {
    prefetch_l1(g2 + 32 * lane);
    prefetch_l1(b2 + 32 * lane);
  }
#endif
  for (int r0 = w; r0 < mt_count * 16; r0 += 2 * kNW) {
    const int r1 = r0 + kNW;
    const bool ok0 = r0 < ins.nrows, ok1 = r1 < ins.nrows;
    // rows past the end are computed on a clamped (valid) row and stored as zeros: no divergent load paths
    const int l0 = ins.row0 + (ok0 ? r0 : 0), l1 = ins.row0 + (ok1 ? r1 : 0);   // logical rows (logits / latent rows)
    float4 v0[8], v1[8];
    ln_load_row(P, ins, src, (ins.flags & F_ROW_MAP) ? P.head_rows[l0] : l0, lane, v0);
    ln_load_row(P, ins, src, (ins.flags & F_ROW_MAP) ? P.head_rows[l1] : l1, lane, v1);
    uint8_t* d0 = a_rm + (size_t)r0 * kARowBytes;
    uint8_t* d1 = a_rm + (size_t)r1 * kARowBytes;
    if (ok0) {
      ln_finish_row<kDouble>(P, ins, d0, l0, lane, v0, g1, b1, g2, b2);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) *reinterpret_cast<uint2*>(d0 + (128 * j + 4 * lane) * 2) = make_uint2(0u, 0u);
    }
    if (ok1) {
      ln_finish_row<kDouble>(P, ins, d1, l1, lane, v1, g1, b1, g2, b2);
    } else if (r1 < mt_count * 16) {
#pragma unroll
      for (int j = 0; j < 8; ++j) *reinterpret_cast<uint2*>(d1 + (128 * j + 4 * lane) * 2) = make_uint2(0u, 0u);
    }
  }
}

// ---- LayerNorm of ALL rows of an ln_1 / ln_2 stage, once, grid-wide (interpreter, more than 16 rows) ---------------
// Warp (cta, w) normalises rows gw, gw + 8 G, ... and stores them as bf16 in mma A-fragment order (P.xn); after the grid
// barrier that follows, the GEMM reads its A fragments from L2 exactly like the attn.c_proj stage reads the context.
// Without this every CTA re-normalises all rows for its own column slice: 148 x the loads and the arithmetic.
__device__ __forceinline__ bool ln_split_active(const Params& P) { return P.xn != nullptr && P.n_rows > kLnSplitMinRows; }

static __device__ void ln_rows_to_global(const Params& P, const Instr& ins, const float* g, const float* b) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int row = w * P.G + blockIdx.x; row < P.n_rows; row += kNW * P.G) {     // CTA-major: one row per CTA first
    float4 v[8];
    ln_load_row(P, ins, P.x, row, lane, v);
    ln_normalise(v, g, b, lane);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int col = 128 * j + 4 * lane;
      __nv_bfloat16* d = P.xn + afrag_elem_offset(row, col, kD);
      *reinterpret_cast<uint32_t*>(d) = pack_bf16x2(v[j].x, v[j].y);
      *reinterpret_cast<uint32_t*>(P.xn + afrag_elem_offset(row, col + 2, kD)) = pack_bf16x2(v[j].z, v[j].w);
    }
  }
}

}  // namespace tts
