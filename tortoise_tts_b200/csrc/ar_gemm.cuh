// ar_gemm.cuh -- the weight-streaming skinny GEMM stage:  out[rows<=64, n-range of this CTA] = A[rows, K] . W[K, n-range]
//
// HBM-bound by construction (every weight byte is used by <= 64 rows): weights arrive as 16 KiB
// fragment-ordered chunks through the bulk-copy ring (ar_core.cuh producer_loop), each consumer warp owns
// a k-slice of every chunk (split-K inside the CTA), multiplies with mma.m16n8k16 (bf16 x bf16 -> fp32)
// and the 16 partial tiles are summed in a fixed order through shared memory (deterministic).
// Dense contractions (prompt pass at large row counts) are a different regime; see DESIGN.md.
#pragma once
#include "ar_core.cuh"

namespace tts {

// Epilogue of one output pair (row, col..col+1) with the fp32 sums v0, v1.
__device__ __forceinline__ void gemm_epilogue(const Params& P, const Instr& ins, int row_in_group, int col, float v0,
                                              float v1) {
  if (row_in_group >= ins.nrows) return;
  const int row = ins.row0 + row_in_group;
  if ((ins.flags & F_ROW_RANGE) && (row < ins.pad0 || row >= ins.pad1)) return;
  switch (ins.op) {
    case OP_LN_QKV: {
      const float* bias = P.layers[ins.layer].b_qkv;
      v0 += bias[col];
      v1 += bias[col + 1];
      const int part = col >> 10, c = col & (kD - 1);
      if (part == 0) {   // query, pre-scaled by 1/sqrt(64) (exact: power of two)
        *reinterpret_cast<float2*>(P.q + (size_t)row * kD + c) = make_float2(v0 * 0.125f, v1 * 0.125f);
      } else {           // key / value -> page slot of this row's position
        const int pos = ld_cg_i(P.row_pos + row);   // written by OP_SAMPLE on another CTA
        if (pos < 0) return;
        const int u = P.row_utt[row];
        const int plen = P.utt_plen[u];
        int page, slot;
        if (pos < plen) {
          page = P.prefix_pt[u * kMaxPagesPerSeg + (pos >> 5)];
          slot = pos & 31;
        } else {
          const int pp = pos - plen;
          page = P.priv_pt[row * kMaxPagesPerSeg + (pp >> 5)];
          slot = pp & 31;
        }
        __nv_bfloat16* dst = P.kv + kv_row_offset(ins.layer, P.n_pages, page, part - 1, c >> 6, slot) + (c & 63);
        *reinterpret_cast<uint32_t*>(dst) = pack_bf16x2(v0, v1);
      }
      break;
    }
    case OP_PROJ:
    case OP_PROJ2: {
      const float* bias = (ins.op == OP_PROJ) ? P.layers[ins.layer].b_proj : P.layers[ins.layer].b_proj2;
      float* xp = P.x + (size_t)row * kD + col;
      if (P.y && ins.op == OP_PROJ2) {   // cluster decode kernel: x = y + mlp ; y is handed back zeroed
        long long* yp = P.y + (size_t)row * kD + col;
        const longlong2 old = ld_cg_ll2(yp);
        *reinterpret_cast<float2*>(xp) = make_float2(from_fixed(old.x) + v0 + bias[col], from_fixed(old.y) + v1 + bias[col + 1]);
        *reinterpret_cast<longlong2*>(yp) = make_longlong2(0ll, 0ll);
        break;
      }
      const float2 old = ld_cg_f2(xp);
      *reinterpret_cast<float2*>(xp) = make_float2(old.x + v0 + bias[col], old.y + v1 + bias[col + 1]);
      break;
    }
    case OP_LN_FC: {
      const float* bias = P.layers[ins.layer].b_fc;
      const float h0 = gelu_new(v0 + bias[col]), h1 = gelu_new(v1 + bias[col + 1]);
      *reinterpret_cast<uint32_t*>(P.hbuf + afrag_elem_offset(row, col, kD4)) = pack_bf16x2(h0, h1);
      break;
    }
    case OP_HEAD: {
      float* lp = P.logits + (size_t)row * P.v_stride + col;
      if (col < P.V) lp[0] = v0 + P.b_head[col];
      if (col + 1 < P.V) lp[1] = v1 + P.b_head[col + 1];
      break;
    }
    default: break;
  }
}

// MT = M-tiles of 16 rows in this instruction (1, 2, 4 or 8; 8 only with the A operand in L2).  Warp w: m-tile w % MT, k-slice w / MT.
// Tiles of this CTA: t0 + (i / tinner) * tstride + i % tinner, i < cnt (tinner >= cnt = the contiguous schedule
// of tile_range).
template <int MT>
__device__ void gemm_tiles(const Params& P, const Instr& ins, const Smem& sm, Pipe& pipe, int t0, int cnt, int tstride,
                           int tinner, const float* ln_src, bool a_staged = false) {
  constexpr int KS = kNW / MT;             // k-slices
  constexpr int KPW = kKbPerChunk / KS;    // k16-blocks per warp per chunk (4 * MT)
  constexpr int KB = KPW > 16 ? 16 : KPW;  // A-fragment batch
  constexpr int NT = 4;                    // n8-tiles accumulated per pass
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int mt = w % MT, ks = w / MT;
  Ctrl* c = sm.ctrl;

  GemmDesc gd;
  gemm_desc(P, ins, gd);
  const int KQ = gd.K / kChunkK;
  const bool ln_split = ln_split_active(P) && (ins.op == OP_LN_QKV || ins.op == OP_LN_FC);
  const bool a_from_smem = !ln_split && (ins.op == OP_LN_QKV || ins.op == OP_LN_FC || ins.op == OP_HEAD);

  // ---- A operand ----
  if (a_from_smem && !a_staged) {
    if (ins.op == OP_LN_QKV)
      ln_rows_to_smem<false>(P, ins, sm.a_rm, ln_src, P.layers[ins.layer].ln1_g, P.layers[ins.layer].ln1_b, nullptr, nullptr, MT);
    else if (ins.op == OP_LN_FC)
      ln_rows_to_smem<false>(P, ins, sm.a_rm, ln_src, P.layers[ins.layer].ln2_g, P.layers[ins.layer].ln2_b, nullptr, nullptr, MT);
    else
      ln_rows_to_smem<true>(P, ins, sm.a_rm, ln_src, P.lnf_g, P.lnf_b, P.fn_g, P.fn_b, MT);
    consumer_sync();
  }
  prof_mark(P, c, 1);
  // ldmatrix row address of this lane inside m-tile `mt` (see layout notes in DESIGN.md)
  const uint32_t a_sm_base = smem_u32(sm.a_rm) + (uint32_t)((mt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * kARowBytes) +
                             (uint32_t)((lane >> 4) * 16);
  const uint4* a_gl = nullptr;
  if (!a_from_smem) {
    const __nv_bfloat16* base = ln_split ? P.xn : (ins.op == OP_PROJ) ? P.ctx : P.hbuf;
    a_gl = reinterpret_cast<const uint4*>(base) + ((size_t)((ins.row0 >> 4) + mt) * (gd.K >> 4)) * 32 + lane;
  }

  for (int tp = 0; tp < cnt; tp += NT) {
    const int ntc = (cnt - tp < NT) ? (cnt - tp) : NT;
    float acc2[NT][2][4];                    // two independent accumulator chains per tile (mma latency)
#pragma unroll
    for (int i = 0; i < NT; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc2[i][0][j] = acc2[i][1][j] = 0.f;

#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      if (nt < ntc) {
        for (int kq = 0; kq < KQ; ++kq) {
          // A fragments of this warp's k-slice, in batches of KB k16-blocks (first batch issued before
          // waiting for the weights)
          uint4 a[KB];
          const int kb0 = kq * kKbPerChunk + ks * KPW;
          bool ok = false;
#pragma unroll
          for (int hb = 0; hb < KPW; hb += KB) {
            if (a_from_smem) {
#pragma unroll
              for (int i = 0; i < KB; ++i)
                ldmatrix_x4(a_sm_base + (uint32_t)((kb0 + hb + i) * 32), a[i].x, a[i].y, a[i].z, a[i].w);
            } else {
#pragma unroll
              for (int i = 0; i < KB; ++i) a[i] = ld_cg_u4(a_gl + (size_t)(kb0 + hb + i) * 32);
            }
            if (hb == 0) ok = mbar_wait(&c->full[pipe.slot], pipe.phase, &c->abort_flag);
            if (ok) {
              const uint4* bsm = reinterpret_cast<const uint4*>(sm.ring + (size_t)pipe.slot * kChunkBytes) +
                                 (ks * (KPW / 2) + hb / 2) * 32 + lane;
#pragma unroll
              for (int s = 0; s < KB / 2; ++s) {
                const uint4 b = bsm[s * 32];
                mma_bf16_16816(acc2[nt][s & 1], a[2 * s], b.x, b.y);
                mma_bf16_16816(acc2[nt][s & 1], a[2 * s + 1], b.z, b.w);
              }
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&c->empty[pipe.slot]);
          pipe.advance();
        }
      }
    }
    prof_mark(P, c, 2);
    // ---- cross-warp (split-K) reduction, fixed order ----
    float4* red4 = reinterpret_cast<float4*>(sm.red);
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
      if (nt < ntc)
        red4[(w * NT + nt) * 32 + lane] = make_float4(acc2[nt][0][0] + acc2[nt][1][0], acc2[nt][0][1] + acc2[nt][1][1],
                                                      acc2[nt][0][2] + acc2[nt][1][2], acc2[nt][0][3] + acc2[nt][1][3]);
    consumer_sync();
    const float2* red2 = reinterpret_cast<const float2*>(sm.red);
    for (int e = threadIdx.x; e < MT * ntc * 64; e += kConsumerThreads) {
      const int emt = e / (ntc * 64);
      const int rem = e - emt * (ntc * 64);
      const int nt = rem >> 6;
      const int ln = (rem & 63) >> 1, pr = rem & 1;
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int k = 0; k < KS; ++k) {
        const float2 v = red2[(((k * MT + emt) * NT + nt) * 32 + ln) * 2 + pr];
        s0 += v.x;
        s1 += v.y;
      }
      gemm_epilogue(P, ins, emt * 16 + (ln >> 2) + 8 * pr, (t0 + ((tp + nt) / tinner) * tstride + (tp + nt) % tinner) * 8 + 2 * (ln & 3), s0, s1);
    }
    consumer_sync();
    prof_mark(P, c, 3);
  }
}

template <int MT>
__device__ void gemm_stage(const Params& P, const Instr& ins, const Smem& sm, Pipe& pipe) {
  GemmDesc gd;
  gemm_desc(P, ins, gd);
  int t0, cnt;
  tile_range(gd.n_tiles, P.G, ins.rot, blockIdx.x, &t0, &cnt);
  gemm_tiles<MT>(P, ins, sm, pipe, t0, cnt, 0, 1 << 30, P.x);
}

}  // namespace tts
