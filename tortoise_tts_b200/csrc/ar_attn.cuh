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
  *reinterpret_cast<uint32_t*>(P.ctx + afrag_elem_offset(row, col, kD)) = pack_bf16x2(v0, v1);
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
    const __nv_bfloat16* vb = kb + (size_t)kH * kPage * kHD;
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
