// ar_attn.cuh -- split-KV attention of one query token per row over the paged bf16 KV cache.
//
// HBM/L2-bound streaming kernel: one warp per work unit (row, head, position range); K and V rows are
// read with coalesced 128-bit loads (8 lanes per 128-byte row, 4 rows per warp instruction), the dot
// product is finished with 3 shuffles, the softmax is the online (running max / running sum) form in
// fp32, and when a (row, head) is split over several warps the last warp to arrive merges the partials in
// a fixed order (deterministic).  Replaces HF GPT2Attention + DynamicCache.update (torch.cat per step) that
// the reference calls through GPT2InferenceModel.forward (unified_voice.py:216-229).
#pragma once
#include "ar_core.cuh"

namespace tts {

constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ void write_ctx_pair(const Params& P, int row, int col, float v0, float v1) {
  const int64_t off = P.ctx_row_major ? (int64_t)row * kD + col : afrag_elem_offset(row, col, kD);
  *reinterpret_cast<uint32_t*>(P.ctx + off) = pack_bf16x2(v0, v1);
}

// Builds the decode split plan for rows [0, nrows): executed by ONE warp.
//   plan[0] = 1 (split mode)  plan[1] = n_units  plan[2] = tp  plan[4 + r] = first unit of row r (nrows + 1 entries)
// Row r gets ns_r = min(32, ceil(kvlen_r / tp)) splits; its units are [split][head].
static __device__ void build_attn_plan(const Params& P, int nrows) {
  const int lane = threadIdx.x & 31;
  int total = 0, maxlen = 0;
  for (int r = lane; r < nrows; r += 32) {
    const int len = ld_cg_i(P.row_pos + r) + 1;
    total += len > 0 ? len : 0;
    maxlen = max(maxlen, len);
  }
  total = __reduce_add_sync(0xffffffffu, total);
  maxlen = __reduce_max_sync(0xffffffffu, maxlen);
  const int cap = min(P.G * kNW, P.plan_cap_units);
  int tp = max(8, (int)(((long long)total * kH + cap - 1) / cap));
  int units = 0;
  for (int it = 0; it < 64; ++it) {
    units = 0;
    for (int r = lane; r < nrows; r += 32) {
      const int len = ld_cg_i(P.row_pos + r) + 1;
      if (len > 0) units += kH * min(32, (len + tp - 1) / tp);
    }
    units = __reduce_add_sync(0xffffffffu, units);
    if (units <= cap || tp >= maxlen) break;
    tp += max(1, tp >> 3);
  }
  int running = 0;
  for (int r0 = 0; r0 < nrows; r0 += 32) {
    const int r = r0 + lane;
    int n = 0;
    if (r < nrows) {
      const int len = ld_cg_i(P.row_pos + r) + 1;
      if (len > 0) n = kH * min(32, (len + tp - 1) / tp);
    }
    int incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (r < nrows) P.plan[4 + r] = running + incl - n;
    running += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) {
    P.plan[4 + nrows] = running;
    P.plan[0] = 1;
    P.plan[1] = running;
    P.plan[2] = tp;
  }
}

// One work unit: positions [p0, p1) of `row`'s sequence for `head`.  Result: running max m (log2 domain),
// sum l and the un-normalised output (dims 8*dc .. 8*dc+7 in lanes with rg == 0 after the final reduce).
__device__ __forceinline__ void attn_unit(const Params& P, int layer, int row, int head, int p0, int p1, float& m_out,
                                          float& l_out, float (&acc)[8]) {
  const int lane = threadIdx.x & 31;
  const int rg = lane >> 3, dc = lane & 7;
  float qv[8];
  {
    const float* qp = P.q + (size_t)row * kD + head * kHD + dc * 8;
    const float4 a = ld_cg_f4(qp), b = ld_cg_f4(qp + 4);
    qv[0] = a.x * kLog2e; qv[1] = a.y * kLog2e; qv[2] = a.z * kLog2e; qv[3] = a.w * kLog2e;
    qv[4] = b.x * kLog2e; qv[5] = b.y * kLog2e; qv[6] = b.z * kLog2e; qv[7] = b.w * kLog2e;
  }
  const int u = P.row_utt[row];
  const int plen = P.utt_plen[u];
  float m = -INFINITY, l = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  int p = p0;
  while (p < p1) {
    int page, s0, seg_end;
    if (p < plen) {
      page = P.prefix_pt[u * kMaxPagesPerSeg + (p >> 5)];
      s0 = p & 31;
      seg_end = min(p1, plen);
    } else {
      const int pp = p - plen;
      page = P.priv_pt[row * kMaxPagesPerSeg + (pp >> 5)];
      s0 = pp & 31;
      seg_end = p1;
    }
    const int run = min(seg_end - p, kPage - s0);
    const __nv_bfloat16* kb = P.kv + kv_row_offset(layer, P.n_pages, page, 0, head, s0) + dc * 8;
    const __nv_bfloat16* vb = kb + kKvVOffset;
    uint4 kf[8], vf[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int r = 4 * j + rg;
      kf[j] = (r < run) ? ld_cg_u4(kb + r * kHD) : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int r = 4 * j + rg;
      vf[j] = (r < run) ? ld_cg_u4(vb + r * kHD) : make_uint4(0u, 0u, 0u, 0u);
    }
    float sc[8];
    float cm = -INFINITY;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float s = qv[0] * bf16_lo(kf[j].x);
      s = fmaf(qv[1], bf16_hi(kf[j].x), s);
      s = fmaf(qv[2], bf16_lo(kf[j].y), s);
      s = fmaf(qv[3], bf16_hi(kf[j].y), s);
      s = fmaf(qv[4], bf16_lo(kf[j].z), s);
      s = fmaf(qv[5], bf16_hi(kf[j].z), s);
      s = fmaf(qv[6], bf16_lo(kf[j].w), s);
      s = fmaf(qv[7], bf16_hi(kf[j].w), s);
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      s += __shfl_xor_sync(0xffffffffu, s, 4);
      sc[j] = (4 * j + rg < run) ? s : -INFINITY;
      cm = fmaxf(cm, sc[j]);
    }
    cm = fmaxf(cm, __shfl_xor_sync(0xffffffffu, cm, 8));
    cm = fmaxf(cm, __shfl_xor_sync(0xffffffffu, cm, 16));
    const float m_new = fmaxf(m, cm);                 // finite: run >= 1
    const float scale = exp2f(m - m_new);             // m == -inf -> 0
    l *= scale;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] *= scale;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float pj = exp2f(sc[j] - m_new);          // masked rows: exp2(-inf) = 0
      l += pj;
      acc[0] = fmaf(pj, bf16_lo(vf[j].x), acc[0]);
      acc[1] = fmaf(pj, bf16_hi(vf[j].x), acc[1]);
      acc[2] = fmaf(pj, bf16_lo(vf[j].y), acc[2]);
      acc[3] = fmaf(pj, bf16_hi(vf[j].y), acc[3]);
      acc[4] = fmaf(pj, bf16_lo(vf[j].z), acc[4]);
      acc[5] = fmaf(pj, bf16_hi(vf[j].z), acc[5]);
      acc[6] = fmaf(pj, bf16_lo(vf[j].w), acc[6]);
      acc[7] = fmaf(pj, bf16_hi(vf[j].w), acc[7]);
    }
    m = m_new;
    p += run;
  }
  l += __shfl_xor_sync(0xffffffffu, l, 8);
  l += __shfl_xor_sync(0xffffffffu, l, 16);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 8);
    acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 16);
  }
  m_out = m;
  l_out = l;
}

// Merge the `ns` partials of (row, head); first partial index u0, stride kH units between splits.
__device__ __forceinline__ void attn_combine(const Params& P, int row, int head, int u0, int ns) {
  const int lane = threadIdx.x & 31;
  float ms = -INFINITY, ls = 0.f;
  if (lane < ns) {
    const float2 ml = ld_cg_f2(P.part + (size_t)(u0 + lane * kH) * kPartStride);
    ms = ml.x;
    ls = ml.y;
  }
  const float M = warp_max(ms);
  const float wgt = (lane < ns) ? exp2f(ms - M) : 0.f;
  const float L = warp_sum(wgt * ls);
  float o0 = 0.f, o1 = 0.f;
  for (int s = 0; s < ns; ++s) {
    const float ws = __shfl_sync(0xffffffffu, wgt, s);
    const float2 a = ld_cg_f2(P.part + (size_t)(u0 + s * kH) * kPartStride + 8 + 2 * lane);
    o0 = fmaf(ws, a.x, o0);
    o1 = fmaf(ws, a.y, o1);
  }
  const float inv = 1.0f / L;
  write_ctx_pair(P, row, head * kHD + 2 * lane, o0 * inv, o1 * inv);
}

// ---- prompt pass / latent re-forward: one warp = 16 consecutive positions of ONE sequence x one head ---------------
// Causal attention of a 16-row query tile over positions [0, pos0 + nq) with mma.m16n8k16 (QK^T and PV), online
// softmax in fp32 (log2 domain).  The 16 queries share every K/V chunk they read: 16 x fewer L2 bytes than one
// decode-shaped unit per (row, head).  All K/V of a prompt-shaped sequence lives in its prefix pages.
//   S (16 x 16 chunk) = Q (A fragments: bf16(q * log2 e), q is already scaled by 1/8) . K^T (B = K[pos][d], 32-bit loads)
//   O (16 x 64)      += P (A fragments straight from the S accumulators) . V (B = V[pos][d] transposed: 16-bit loads)
__device__ __forceinline__ uint32_t ld_cg_u32(const void* p) { return __ldcg(reinterpret_cast<const unsigned int*>(p)); }
__device__ __forceinline__ uint32_t ld_cg_u16(const void* p) { return (uint32_t)__ldcg(reinterpret_cast<const unsigned short*>(p)); }

static __device__ void attn_tile16(const Params& P, int layer, int r0, int nq, int head) {
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int u = P.row_utt[r0];
  const int pos0 = ld_cg_i(P.row_pos + r0);
  const int kv_len = pos0 + nq;                      // positions 0 .. pos0 + nq - 1
  const int* pt = P.prefix_pt + u * kMaxPagesPerSeg;
  // Q fragments: 4 k16-blocks of the head's 64 dims
  uint4 qa[4];
  {
    const int ra = r0 + min(g, nq - 1), rb = r0 + min(g + 8, nq - 1);       // rows past the tile end: clamped, never stored
    const float* qpa = P.q + (size_t)ra * kD + head * kHD + 2 * t;
    const float* qpb = P.q + (size_t)rb * kD + head * kHD + 2 * t;
#pragma unroll
    for (int kb = 0; kb < 4; ++kb) {
      const float2 a0 = ld_cg_f2(qpa + kb * 16), a1 = ld_cg_f2(qpb + kb * 16);
      const float2 a2 = ld_cg_f2(qpa + kb * 16 + 8), a3 = ld_cg_f2(qpb + kb * 16 + 8);
      qa[kb] = make_uint4(pack_bf16x2(a0.x * kLog2e, a0.y * kLog2e), pack_bf16x2(a1.x * kLog2e, a1.y * kLog2e),
                          pack_bf16x2(a2.x * kLog2e, a2.y * kLog2e), pack_bf16x2(a3.x * kLog2e, a3.y * kLog2e));
    }
  }
  float o[8][4];
#pragma unroll
  for (int j = 0; j < 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;      // rows g and g + 8
  for (int c0 = 0; c0 < kv_len; c0 += 16) {
    const int page = pt[c0 >> 5];
    const __nv_bfloat16* kb_ = P.kv + kv_row_offset(layer, P.n_pages, page, 0, head, c0 & 31);
    const __nv_bfloat16* vb_ = kb_ + kKvVOffset;
    // K: B fragments of the two position n-tiles; V: B fragments of the eight dim n-tiles (issued together)
    uint32_t kf[2][4][2];
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int kb = 0; kb < 4; ++kb) {
        const __nv_bfloat16* kp = kb_ + (8 * j + g) * kHD + kb * 16 + 2 * t;
        kf[j][kb][0] = ld_cg_u32(kp);
        kf[j][kb][1] = ld_cg_u32(kp + 8);
      }
    uint32_t vf[8][2];
#pragma unroll
    for (int jd = 0; jd < 8; ++jd) {
      const __nv_bfloat16* vp = vb_ + (2 * t) * kHD + 8 * jd + g;
      vf[jd][0] = ld_cg_u16(vp) | (ld_cg_u16(vp + kHD) << 16);
      vf[jd][1] = ld_cg_u16(vp + 8 * kHD) | (ld_cg_u16(vp + 9 * kHD) << 16);
    }
    {   // slots past the end of the sequence hold stale bits (0 x NaN = NaN in the PV product): zero them
      const int pv = c0 + 2 * t;
      const uint32_t k0 = (pv < kv_len ? 0x0000ffffu : 0u) | (pv + 1 < kv_len ? 0xffff0000u : 0u);
      const uint32_t k1 = (pv + 8 < kv_len ? 0x0000ffffu : 0u) | (pv + 9 < kv_len ? 0xffff0000u : 0u);
#pragma unroll
      for (int jd = 0; jd < 8; ++jd) { vf[jd][0] &= k0; vf[jd][1] &= k1; }
    }
    float s[2][4];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
      for (int kb = 0; kb < 4; ++kb) mma_bf16_16816(s[j], qa[kb], kf[j][kb][0], kf[j][kb][1]);
    }
    // causal mask: position <= pos0 + row index in the tile  (slots past the sequence end hold stale data: masked too)
    float cm0 = -INFINITY, cm1 = -INFINITY;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int p = c0 + 8 * j + 2 * t;
      if (p > pos0 + g) s[j][0] = -INFINITY;
      if (p + 1 > pos0 + g) s[j][1] = -INFINITY;
      if (p > pos0 + g + 8) s[j][2] = -INFINITY;
      if (p + 1 > pos0 + g + 8) s[j][3] = -INFINITY;
      cm0 = fmaxf(cm0, fmaxf(s[j][0], s[j][1]));
      cm1 = fmaxf(cm1, fmaxf(s[j][2], s[j][3]));
    }
    cm0 = fmaxf(cm0, __shfl_xor_sync(0xffffffffu, cm0, 1));
    cm0 = fmaxf(cm0, __shfl_xor_sync(0xffffffffu, cm0, 2));
    cm1 = fmaxf(cm1, __shfl_xor_sync(0xffffffffu, cm1, 1));
    cm1 = fmaxf(cm1, __shfl_xor_sync(0xffffffffu, cm1, 2));
    const float mn0 = fmaxf(m0, cm0), mn1 = fmaxf(m1, cm1);      // finite from the first chunk on (position 0 is allowed)
    const float sc0 = exp2f(m0 - mn0), sc1 = exp2f(m1 - mn1);
    m0 = mn0; m1 = mn1;
    l0 *= sc0; l1 *= sc1;
#pragma unroll
    for (int jd = 0; jd < 8; ++jd) { o[jd][0] *= sc0; o[jd][1] *= sc0; o[jd][2] *= sc1; o[jd][3] *= sc1; }
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
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float i0 = 1.0f / l0, i1 = 1.0f / l1;
#pragma unroll
  for (int jd = 0; jd < 8; ++jd) {
    if (g < nq) write_ctx_pair(P, r0 + g, head * kHD + 8 * jd + 2 * t, o[jd][0] * i0, o[jd][1] * i0);
    if (g + 8 < nq) write_ctx_pair(P, r0 + g + 8, head * kHD + 8 * jd + 2 * t, o[jd][2] * i1, o[jd][3] * i1);
  }
}

static __device__ void attn_stage(const Params& P, const Instr& ins, const Smem& sm) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int rg = lane >> 3, dc = lane & 7;
  const int mode = ld_cg_i(P.plan + 0);
  const int nrows = ins.nrows;
  int n_units;
  int* unit_off = reinterpret_cast<int*>(sm.red);
  if (mode == 1) {
    n_units = ld_cg_i(P.plan + 1);
    for (int i = threadIdx.x; i <= nrows; i += kConsumerThreads) unit_off[i] = ld_cg_i(P.plan + 4 + i);
    consumer_sync();
  } else {
    n_units = nrows * kH;
  }
  prof_mark(P, sm.ctrl, 11);
  // prompt-shaped rows: 16-row query tiles once there are enough of them to occupy every warp of the grid (a short prompt
  // is latency-bound: one decode-shaped unit per (row, head) spreads it over 16 x more warps and is faster, 8.8 vs 10.0 ms
  // at 540 rows); tiles that straddle two sequences fall back to one unit per row
  const int n_tiles = (nrows + 15) >> 4;
  if (mode == 0 && P.decode_mode == 0 && P.attn_tiles_mode != 1 && (n_tiles * kH >= P.G * kNW || P.attn_tiles_mode == 2)) {
    for (int ut = blockIdx.x * kNW + w; ut < n_tiles * kH; ut += P.G * kNW) {
      const int tile = ut / kH, head = ut - tile * kH;
      const int r0 = ins.row0 + 16 * tile, nq = min(16, nrows - 16 * tile);
      const int pa = ld_cg_i(P.row_pos + r0), pb = ld_cg_i(P.row_pos + r0 + nq - 1);
      if (pa >= 0 && P.row_utt[r0] == P.row_utt[r0 + nq - 1] && pb == pa + nq - 1) {
        attn_tile16(P, ins.layer, r0, nq, head);
      } else {
        for (int row = r0; row < r0 + nq; ++row) {
          const int p1 = ld_cg_i(P.row_pos + row) + 1;
          if (p1 <= 0) continue;
          float m, l, acc[8];
          attn_unit(P, ins.layer, row, head, 0, p1, m, l, acc);
          if (rg == 0) {
            const float inv = 1.0f / l;
#pragma unroll
            for (int i = 0; i < 8; i += 2) write_ctx_pair(P, row, head * kHD + dc * 8 + i, acc[i] * inv, acc[i + 1] * inv);
          }
        }
      }
    }
    return;
  }
  for (int u = blockIdx.x * kNW + w; u < n_units; u += P.G * kNW) {
    int row, head, p0, p1, ns, split;
    if (mode == 1) {
      int lo = 0, hi = nrows;                        // unit_off[lo] <= u < unit_off[hi]
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (unit_off[mid] <= u) lo = mid; else hi = mid;
      }
      row = lo;
      const int local = u - unit_off[row];
      ns = (unit_off[row + 1] - unit_off[row]) / kH;
      split = local / kH;
      head = local - split * kH;
      const int len = ld_cg_i(P.row_pos + ins.row0 + row) + 1;
      p0 = (int)(((long long)split * len) / ns);
      p1 = (int)(((long long)(split + 1) * len) / ns);
      row += ins.row0;
    } else {
      row = ins.row0 + u / kH;
      head = u % kH;
      ns = 1;
      split = 0;
      p0 = 0;
      p1 = ld_cg_i(P.row_pos + row) + 1;
      if (p1 <= 0) continue;
    }
    float m, l, acc[8];
    attn_unit(P, ins.layer, row, head, p0, p1, m, l, acc);
    if (ns == 1) {
      if (rg == 0) {
        const float inv = 1.0f / l;
#pragma unroll
        for (int i = 0; i < 8; i += 2) write_ctx_pair(P, row, head * kHD + dc * 8 + i, acc[i] * inv, acc[i + 1] * inv);
      }
    } else {
      float* pp = P.part + (size_t)u * kPartStride;
      if (rg == 0) {
        *reinterpret_cast<float4*>(pp + 8 + dc * 8) = make_float4(acc[0], acc[1], acc[2], acc[3]);
        *reinterpret_cast<float4*>(pp + 8 + dc * 8 + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
      }
      if (lane == 0) *reinterpret_cast<float2*>(pp) = make_float2(m, l);
      __threadfence();
      __syncwarp();
      int old = 0;
      if (lane == 0) old = atomicAdd(P.attn_cnt + row * kH + head, 1);
      old = __shfl_sync(0xffffffffu, old, 0);
      if (old == ns - 1) {                          // last split of this (row, head): merge, fixed order
        __threadfence();
        attn_combine(P, row, head, u - split * kH, ns);
        if (lane == 0) P.attn_cnt[row * kH + head] = 0;
      }
    }
  }
}

}  // namespace tts
