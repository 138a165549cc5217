// ar_decode.cu -- the head-cluster decode kernels: the sample loop for <= 16, <= 32 or <= 64 live sequences per launch.
//
// Same stages as the interpreter (ar_kernels.cu) but with 3 grid-wide barriers per GPT-2 block instead of 5:
// 32 thread-block clusters of 4 CTAs, cluster c owns attention head c >> 1 for one half of the launch's rows, so
//     ln_1 -> c_attn (q,k,v columns of the head) -> [cluster barrier] -> attention of the head ->
//     [cluster barrier] -> split-K attn.c_proj (rows 64h..64h+63 of the weight) -> red.add into y
// runs without leaving the cluster (phase A); the 16 per-head partial products are summed in L2 by 64-bit integer
// atomics in 2^-32 fixed point (order-independent => deterministic) into the accumulator y, which the previous stage
// left initialised to x + bias (the residual), so all 16 heads do the same work.
// Phase B = ln_2 -> c_fc -> gelu_new (N split over the 128 CTAs), phase C = mlp.c_proj: x = y + . ; y = x + next bias.
// Weights stream through the same bulk-copy ring as in the interpreter, ungated by any barrier, and every streamed
// chunk is multiplied with ALL rows of the launch: the template parameter MODE = 16-row m-tiles per launch
//     MODE 1: <= 16 rows; a cluster owns 8 rows of its head (half an m-tile), a CTA attends 2 rows
//     MODE 2: <= 32 rows; a cluster owns one m-tile (16 rows), a CTA attends 4 rows
//     MODE 4: <= 64 rows; a cluster owns two m-tiles (32 rows), a CTA attends 8 rows; phases B / C / head run the 64
//             rows as two 32-row groups over the same resident weight chunks (the LayerNorm'd A operand of 32 rows
//             is what fits next to the ring)
// so that 32 (64) sequences cost one pass over the 773 MB of weights per step, not two (four).
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

constexpr int kDecThreads = kConsumerThreads;     // 8 warps; there is no dedicated producer warp in this kernel

template <int MODE>
struct DecCfg {
  static_assert(MODE == 1 || MODE == 2 || MODE == 4, "16, 32 or 64 rows");
  static constexpr bool kHalf = MODE == 1;
  static constexpr int kMtA = MODE == 4 ? 2 : 1;          // m-tiles a cluster multiplies in phase A
  static constexpr int kRowsA = kHalf ? 8 : 16 * kMtA;    // rows of its head a cluster owns in phase A
#ifdef TTS_LN1_FULL_TILE   // compile-time A/B only: MODE 1 normalises the whole m-tile as before
  static constexpr int kLnRowsA = 16 * kMtA;
#else
  // rows it normalises for that: its own rows.  MODE 1: HALF an m-tile -- the c_attn epilogue keeps only the cluster's 8 rows
  // (the sibling cluster produces the other 8), and an mma row depends on the same row of A only, so the other 8 rows of the
  // A-operand buffer may hold anything: half the row loads and half the distributed-shared bytes of ln_1
  // (same-box A/B 624.0 -> 614.6 us/step, profiles/r02_ab_ln1_half_tile.log)
  static constexpr int kLnRowsA = kRowsA;
#endif
  static constexpr int kAttnRows = kRowsA / kClusterSize; // attention rows per CTA: 2, 4, 8
  static constexpr int kWarpsPerRow = kNW / kAttnRows;    // 4, 2, 1
  static constexpr int kMtB = MODE == 1 ? 1 : 2;          // m-tiles per row group in phases B, C and the head
  static constexpr int kGroups = MODE == 4 ? 2 : 1;       // row groups of 16 * kMtB rows
  static constexpr int kMtSmem = kMtB;                    // m-tiles of the shared-memory A operand
};

// Shared-memory reduction buffer (Smem::red, 33 KiB) of these kernels, in floats:
//   [0, 6144)      cross-warp split-K reduction of one m-tile (<= 6 tiles x 8 warps x 512 B); LayerNorm gains staged at
//                  0 (ln_1, head) or 4096 (ln_2: above the 16 KiB its 4-tile reduction uses, so two row groups share them)
//   [4096, 6144)   mlp.c_proj: K-partials received from the cluster's CTAs, [group][m-tile][source rank][64 pairs] float2
//   [6144, 6784)   per-step row facts of this CTA (rinfo)
//   [7168, 8192)   attention context of the cluster's head, <= 2 m-tiles x 64 dims bf16 in mma A-fragment order
constexpr int kStageOffLn2 = 4096;
constexpr int kXredOffsetFloats = 4096;
constexpr int kRinfoOffsetFloats = 6144;
constexpr int kCtxOffsetFloats = 7168;
// MODE 1 / 2 (<= 4 attention rows per CTA): q and the new position's K / V rows of a CTA's attention rows arrive in its shared
// memory (st.async from the 4 c_attn epilogues of the cluster) instead of through global memory behind a release fence
constexpr int kKvNewOffsetFloats = 7680;   // bf16 [K | V][attention row][64]  (after the 2 KiB context tile of MODE 1 / 2)
constexpr int kQsOffsetFloats = 8192;      // fp32 [attention row][64]
template <int MODE>
struct QkvAsync {
#ifdef TTS_QKV_SYNC
  static constexpr bool on = false;
#else
  static constexpr bool on = MODE <= 2;    // MODE 4 has no shared memory left for the 4 KiB it would need
#endif
};

// ---- hardware cluster barrier (all 256 threads of the 4 CTAs; ~0.2 us).  release/acquire at cluster scope: global
// and distributed-shared writes made before it are visible to the cluster's threads after it. -----------------------
__device__ __forceinline__ void cluster_barrier() {
#ifdef TTS_CLUSTER_RELAXED_DIAG   // timing diagnostic only (results are NOT valid): what the release fence of the arrive costs
  asm volatile("barrier.cluster.arrive.relaxed.aligned;\n\tbarrier.cluster.wait.aligned;" ::: "memory");
#else
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
#endif
}

// Timing diagnostic (TTS_CLUSTER_RELAXED_SITES = bit mask of call sites; results are NOT valid): which release fences cost what.
template <int SITE>
__device__ __forceinline__ void cluster_barrier_at() {
#ifdef TTS_CLUSTER_RELAXED_SITES
  if (TTS_CLUSTER_RELAXED_SITES & SITE) {
    asm volatile("barrier.cluster.arrive.relaxed.aligned;\n\tbarrier.cluster.wait.aligned;" ::: "memory");
    return;
  }
#endif
  cluster_barrier();
}

// ---- weight feed: warp 0 of the CTA keeps the bulk-copy ring full --------------------------------------------------
// The chunk sequence of the whole launch is fixed (per layer: 6 c_attn, 2 split-K c_proj, 4 c_fc, 4 mlp.c_proj chunks;
// then this CTA's mel-head chunks, once per row group), so instead of a producer warp waiting on "slot empty" barriers,
// every stage that has finished reading k ring slots (after a CTA barrier) lets warp 0 issue the next k chunks of the
// sequence into exactly those slots.  The ring is therefore always n_slots chunks (half a layer or more) ahead of the
// consumers, across grid barriers, steps and the sampler.
struct Feed {
  uint32_t issued;           // chunks issued so far (uniform across the feed warp)
  uint32_t per_step, total;  // chunks per model pass / in the whole launch
  uint64_t pol;
};

__device__ __forceinline__ void feed_init(const Params& P, Feed& f, int hcnt, int groups) {
  f.issued = 0;
  f.per_step = 16u * (uint32_t)P.n_layers + (uint32_t)(groups * hcnt);
  f.total = f.per_step * (uint32_t)(P.n_steps - (P.skip_last_model ? 1 : 0));
  f.pol = make_evict_first_policy();
}

// Issue the next n (<= 32) chunks: called by the whole feed warp (warp 0), lane k issues chunk k.  The n slots must be
// free.  (Issuing a 16 KiB bulk copy costs ~0.25-0.3 us whichever thread or warp does it: the SM ingests ~55 GB/s.)
__device__ __forceinline__ void feed(const Params& P, Ctrl* c, uint8_t* ring, Feed& f, int n, int head, int rank, int ht0,
                                     int hcnt) {
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t left = f.total - f.issued;
  const uint32_t m = min((uint32_t)n, left);
  if (lane < m) {
    const uint32_t p = f.issued + lane;
    const uint32_t r = p % f.per_step;
    const uint32_t slot = p % (uint32_t)P.n_slots;
    const uint4* src;
    if (r < 16u * (uint32_t)P.n_layers) {
      const uint4* L = layer_w(P, (int)(r >> 4));
      const int i = (int)(r & 15u);
      if (i < 6) src = L + (size_t)((i >> 1) * (kD / 8) + head * 8 + 2 * rank + (i & 1)) * kU4PerChunk;
      else if (i < 8) src = L + kOffProjk + (size_t)(head * 8 + 2 * rank + (i - 6)) * kU4PerChunk;
      else if (i < 12) src = L + kOffFc + (size_t)(blockIdx.x * 4 + (i - 8)) * kU4PerChunk;
      else src = L + kOffProj2 + (size_t)(((blockIdx.x & ~3) + (i - 12)) * 4 + rank) * kU4PerChunk;   // tile 4c+i, K quarter = rank
    } else {
      src = P.w_head + (size_t)(ht0 + (int)((r - 16u * (uint32_t)P.n_layers) % (uint32_t)hcnt)) * kU4PerChunk;
    }
    mbar_arrive_expect_tx(&c->full[slot], kChunkBytes);
    bulk_g2s(ring + (size_t)slot * kChunkBytes, src, kChunkBytes, &c->full[slot], f.pol);
  }
  f.issued += m;
}

// Refills are issued by warp 0 at the grid barriers (dec_grid_barrier), right after the CTA's arrival and at most 6
// chunks at a time: the transfer then mostly overlaps the wait for the other CTAs instead of the latency-critical loads
// of the stage bodies (LayerNorm rows, K/V, A fragments) on the SM's L2 read port.  (Measured alternatives: a producer
// warp refilling as slots free up 744 us/step; refills in the LayerNorm -> GEMM window 747 us; this 684 us.)
struct Feeder {
  Feed f;                    // meaningful in warp 0 only
  int head, rank, ht0, hcnt;
  int pending;               // ring slots read to the end since the last refill (uniform across the CTA's threads)
};

// ---- per-step row facts of this CTA, kept in shared memory ---------------------------------------------------------
// Positions and page tables do not change inside a step, so the dependent global loads
// (row_pos -> utterance -> prompt length -> page table -> page) are done once per step, not once per layer.
//   [i]      page of the new position of row tile_row0 + i (i < 32; -1: inactive row)   [32 + i] its slot in the page
//   [64 + j] kv length of attention row j (j < 8, 0: inactive)      [72 + j] prompt length of its utterance
//   [80 + 64 j + i] prefix page table (i < 32) / private page table (i >= 32) of attention row j
__device__ __forceinline__ int* rinfo_ptr(const Smem& sm) { return reinterpret_cast<int*>(sm.red + kRinfoOffsetFloats); }
// attention context of this cluster's head in mma A-fragment order (2 KiB per m-tile): every CTA of the cluster holds a
// full copy, written through distributed shared memory by the CTA that merged the row
__device__ __forceinline__ uint4* ctx_smem(const Smem& sm) { return reinterpret_cast<uint4*>(sm.red + kCtxOffsetFloats); }

// Called by `nthr` (>= 72) threads (index t); the caller synchronises them afterwards.
template <int MODE>
__device__ __forceinline__ void build_rinfo(const Params& P, int tile_row0, int attn_row0, const Smem& sm, int t, int nthr) {
  using C = DecCfg<MODE>;
  int* ri = rinfo_ptr(sm);
  for (int idx = t; idx < C::kAttnRows * 64; idx += nthr) {
    const int j = idx >> 6, i = idx & 63, row = attn_row0 + j;
    int v = 0;
    if (row < P.n_rows) v = (i < 32) ? P.prefix_pt[P.row_utt[row] * kMaxPagesPerSeg + i] : P.priv_pt[row * kMaxPagesPerSeg + (i - 32)];
    ri[80 + idx] = v;
  }
  if (t < 16 * C::kMtA) {
    const int row = tile_row0 + t;
    const int pos = (row < P.n_rows) ? ld_cg_i(P.row_pos + row) : -1;
    int page = -1, slot = 0;
    if (pos >= 0) {
      const int pp = pos - P.utt_plen[P.row_utt[row]];          // decode rows are past their prompt
      page = P.priv_pt[row * kMaxPagesPerSeg + (pp >> 5)];
      slot = pp & 31;
    }
    ri[t] = page;
    ri[32 + t] = slot;
  } else if (t >= 64 && t < 64 + C::kAttnRows) {
    const int j = t - 64, row = attn_row0 + j;
    const int pos = (row < P.n_rows) ? ld_cg_i(P.row_pos + row) : -1;
    ri[64 + j] = pos + 1;
    ri[72 + j] = (row < P.n_rows) ? P.utt_plen[P.row_utt[row]] : 0;
  }
}

// Pull the NEXT layer's K/V pages of this CTA's attention rows and its small fp32 parameters into L2 (bulk prefetch), so
// the attention phase and the LayerNorm / bias loads are not HBM stalls.  Called by `nthr` threads (index t).
__device__ __forceinline__ void prefetch_next(const Params& P, const Smem& sm, int head, int pf_layer, int attn_rows, int t, int nthr) {
  if (pf_layer < 0 || pf_layer > P.n_layers) return;
  // the small fp32 parameters (LN gains/biases, GEMM biases: one contiguous block per layer, tts_ar.cu) are
  // read once per step and do not survive 1.3 GB of streaming in L2: pull the next block in, 1/128 per CTA
  if (t == 0) {
    if (pf_layer < P.n_layers) {
      constexpr int kPer = ((4 * kD + kD3 + kD + kD4 + kD) * 4 / kDecCtas + 15) / 16 * 16;   // 416 B
      bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(layer_f(P, pf_layer)) + (size_t)blockIdx.x * kPer, kPer);
    } else {
      const int per = ((4 * kD + P.v_stride) * 4 / kDecCtas + 15) / 16 * 16;
      bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(P.lnf_g) + (size_t)blockIdx.x * per, per);
    }
  }
  if (pf_layer >= P.n_layers) return;
  // One thread per (attention row, K|V, page).  Same-box A/B at 16 rows (profiles/r02_ab_kv_prefetch.log): per-page
  // prefetches 654.9 us/step, one merged prefetch per segment (the pool is head-major: a segment is one contiguous range)
  // 659.7, none 662.3; at 64 rows the prefetch LOSES (1526 vs 1461 us/step without: a layer's K/V is 70+ MB there, two
  // layers do not fit L2 next to the weight stream), so MODE 4 skips it (the caller).
  const int* ri = rinfo_ptr(sm);
  for (int idx = t; idx < attn_rows * 64; idx += nthr) {
    const int j = idx >> 6, i = idx & 63, pg = i >> 1, kv = i & 1;
    const int len = ri[64 + j], plen = ri[72 + j];
    if (len <= 0) continue;
    const int* pt = ri + 80 + 64 * j;
    const __nv_bfloat16* base = P.kv + kv_row_offset(pf_layer, P.n_pages, 0, kv, head, 0);
    if (pg * kPage < plen) bulk_prefetch_l2(base + (size_t)pt[pg] * kKvPageStride, kPage * kHD * 2);
    if (pg * kPage < len - plen) bulk_prefetch_l2(base + (size_t)pt[32 + pg] * kKvPageStride, kPage * kHD * 2);
  }
}

// ---- cluster-cooperative LayerNorm: CTA rank r normalises rows rbase + npc r .. + npc - 1 (one warp per row) and
// writes the bf16 rows into rows npc r .. of the A-operand buffer of all 4 CTAs of the cluster through distributed shared
// memory, so the rows are read from L2 once per cluster instead of once per CTA.  The caller follows with cluster_barrier().
// `gb` = the staged (gain, bias) vectors: one pair, or two for the head (ln_f then final_norm, `dbl`); the head is the
// same call site as ln_1 on purpose -- code that runs once per step is fetched cold, code shared with the layers is not.
// Source of the rows (`src`):
//   LN_SRC_X    plain fp32 x, complete behind a grid barrier (layer 0: written by the sampler)
//   LN_SRC_XT   the tagged residual stream xt (layers >= 1 and the head): written by the mlp.c_proj owners of the previous
//               block as (value, tag) pairs; the warp POLLS its row until every element carries `expect` (the block's tag) --
//               this replaces the grid barrier between mlp.c_proj and ln_1
//   LN_SRC_Y    the counting accumulator y (ln_2): polled until every word has received `expect` contributions (residual
//               + 16 heads) -- this replaces the grid barrier between the attention block and ln_2
// Polling reads one 16-byte piece per lane until it is complete, then the whole row once, and re-reads the row only if a
// piece is still incomplete (contributions land in any order).  Every wait is watchdogged.
enum LnSrc : int { LN_SRC_X = 0, LN_SRC_XT = 1, LN_SRC_Y = 2 };

// `dst0`: row of the A-operand buffer that receives row `rbase` (MODE 1 ln_1 normalises only the half m-tile its cluster owns).
__device__ __forceinline__ void ln_cluster(const Params& P, const Smem& sm, int src, uint32_t expect, const float* gb, bool dbl, int rank,
                                           int rbase, int npc, int dst0 = 0) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (w >= npc) return;
  const int r = npc * rank + w;
  const int grow = rbase + r;
  const bool ok = grow < P.n_rows;
  const int lrow = ok ? grow : 0;
  float4 v[8];
  if (src == LN_SRC_Y) {
    const long long* yp = P.y + (size_t)lrow * kD;
    const long long t0 = clock64();
    for (;;) {
      const longlong2 p0 = ld_relaxed_ll2(yp + 4 * lane);
      if (__all_sync(0xffffffffu, cnt_count(p0.x) == (int)expect && cnt_count(p0.y) == (int)expect)) break;
      if (clock64() - t0 > kWatchdogCycles || sm.ctrl->abort_flag) { sm.ctrl->abort_flag = 1; *reinterpret_cast<volatile int*>(P.watchdog) = 1; break; }
    }
    for (;;) {
      bool good = true;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const longlong2 a = ld_relaxed_ll2(yp + 128 * j + 4 * lane), b = ld_relaxed_ll2(yp + 128 * j + 4 * lane + 2);
        good = good && cnt_count(a.x) == (int)expect && cnt_count(a.y) == (int)expect && cnt_count(b.x) == (int)expect &&
               cnt_count(b.y) == (int)expect;
        v[j] = make_float4(cnt_value(a.x), cnt_value(a.y), cnt_value(b.x), cnt_value(b.y));
      }
      if (__all_sync(0xffffffffu, good)) break;
      if (clock64() - t0 > kWatchdogCycles || sm.ctrl->abort_flag) { sm.ctrl->abort_flag = 1; *reinterpret_cast<volatile int*>(P.watchdog) = 1; break; }
    }
  } else if (src == LN_SRC_XT) {
    const uint2* xp = P.xt + (size_t)lrow * kD;
    const long long t0 = clock64();
    for (;;) {
      const uint4 p0 = ld_relaxed_u4(xp + 2 * lane);
      if (__all_sync(0xffffffffu, p0.y == expect && p0.w == expect)) break;
      if (clock64() - t0 > kWatchdogCycles || sm.ctrl->abort_flag) { sm.ctrl->abort_flag = 1; *reinterpret_cast<volatile int*>(P.watchdog) = 1; break; }
    }
    for (;;) {
      bool good = true;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint4 a = ld_relaxed_u4(xp + 128 * j + 4 * lane), b = ld_relaxed_u4(xp + 128 * j + 4 * lane + 2);
        good = good && a.y == expect && a.w == expect && b.y == expect && b.w == expect;
        v[j] = make_float4(__uint_as_float(a.x), __uint_as_float(a.z), __uint_as_float(b.x), __uint_as_float(b.z));
      }
      if (__all_sync(0xffffffffu, good)) break;
      if (clock64() - t0 > kWatchdogCycles || sm.ctrl->abort_flag) { sm.ctrl->abort_flag = 1; *reinterpret_cast<volatile int*>(P.watchdog) = 1; break; }
    }
  } else {
    const float* xp = P.x + (size_t)lrow * kD;
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = ld_cg_f4(xp + 128 * j + 4 * lane);
  }
  prof_mark(P, sm.ctrl, 21);
  const int npass = dbl ? 2 : 1;
#pragma unroll 1
  for (int p = 0; p < npass; ++p) ln_normalise(v, gb + 2 * p * kD, gb + (2 * p + 1) * kD, lane);
  prof_mark(P, sm.ctrl, 22);
  if (dbl && blockIdx.x < kClusterSize && ok) {      // final_norm(ln_f(x)): the latent of this step
    float* lp = P.cur_latent + (size_t)lrow * kD;
#pragma unroll
    for (int j = 0; j < 8; ++j) *reinterpret_cast<float4*>(lp + 128 * j + 4 * lane) = v[j];
  }
  uint8_t* dst = sm.a_rm + (size_t)(dst0 + r) * kARowBytes;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint2 pk = ok ? make_uint2(pack_bf16x2(v[j].x, v[j].y), pack_bf16x2(v[j].z, v[j].w)) : make_uint2(0u, 0u);
#pragma unroll
    for (int d = 0; d < kClusterSize; ++d) st_cluster_u2(dst + (128 * j + 4 * lane) * 2, d, pk);
  }
  prof_mark(P, sm.ctrl, 23);
}

// ---- attention of head `head` for this CTA's rows --------------------------------------------------------------------
// Warp w: attention row w / kWarpsPerRow, share (w % kWarpsPerRow) of the row's 16-position granules (a granule never
// straddles a page).  K/V granules are triple-buffered in registers so the next granules' 128-bit loads are in flight
// while earlier ones are reduced; the partial (max, sum, acc) states of a row are merged in shared memory.
struct Gran {
  uint4 k[4], v[4];
  int cnt;
};

// pnew >= 0 (last private granule only, MODE 1 / 2): index in the granule of the position written by THIS layer's c_attn;
// its K / V (8 dims of this lane: knew / vnew) are taken from shared memory, the pool copy is for the steps to come
__device__ __forceinline__ void gran_load(Gran& G, const __nv_bfloat16* kvb, const int* pt, int g, int Gp, int plen, int len,
                                          int rg, int pnew = -1, const uint4* knew = nullptr, const uint4* vnew = nullptr) {
  int pos0, cnt, page;
#if defined(TTS_ATTN_DIAG) && TTS_ATTN_DIAG == 3   // timing diagnostic only (results are NOT valid): all K/V reads hit the shared prompt pages
  if (g >= Gp) g = (g - Gp) % (Gp > 1 ? Gp - 1 : 1);
#endif
  if (g < Gp) {
    pos0 = 16 * g;
    cnt = min(16, plen - pos0);
    page = pt[pos0 >> 5];
  } else {
    pos0 = 16 * (g - Gp);
    cnt = min(16, len - plen - pos0);
    page = pt[32 + (pos0 >> 5)];
  }
  const __nv_bfloat16* kb = kvb + (size_t)page * kKvPageStride + (pos0 & 31) * kHD;
  const __nv_bfloat16* vb = kb + kKvVOffset;
  G.cnt = cnt;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int r = 4 * j + rg;
#if defined(TTS_ATTN_DIAG) && TTS_ATTN_DIAG == 1   // timing diagnostic only (results are NOT valid): no K/V traffic
    G.k[j] = make_uint4(0x3c003c00u + r, 0x3c003c00u, 0x3c003c00u + g, 0x3c003c00u);
#else
    G.k[j] = (r < cnt) ? (r == pnew ? *knew : ld_cg_u4(kb + r * kHD)) : make_uint4(0u, 0u, 0u, 0u);
#endif
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int r = 4 * j + rg;
#if defined(TTS_ATTN_DIAG) && TTS_ATTN_DIAG == 1
    G.v[j] = make_uint4(0x3c003c00u + r, 0x3c003c00u, 0x3c003c00u + g, 0x3c003c00u);
#else
    G.v[j] = (r < cnt) ? (r == pnew ? *vnew : ld_cg_u4(vb + r * kHD)) : make_uint4(0u, 0u, 0u, 0u);
#endif
  }
}

__device__ __forceinline__ void gran_reduce(const Gran& G, const float (&qv)[8], int rg, float& m, float& l, float (&acc)[8]) {
#if defined(TTS_ATTN_DIAG) && TTS_ATTN_DIAG == 2   // timing diagnostic only (results are NOT valid): the K/V traffic without the arithmetic
  uint32_t x = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) x ^= G.k[j].x ^ G.k[j].y ^ G.k[j].z ^ G.k[j].w ^ G.v[j].x ^ G.v[j].y ^ G.v[j].z ^ G.v[j].w;
  acc[0] += __uint_as_float(x & 0x3fffffffu) * qv[0];
  m = 0.f; l = 1.f;
  return;
#endif
  float sc[4];
  float cm = -INFINITY;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
#ifndef TTS_ATTN_QK_FFMA2
    float s = qv[0] * bf16_lo(G.k[j].x);
    s = fmaf(qv[1], bf16_hi(G.k[j].x), s);
    s = fmaf(qv[2], bf16_lo(G.k[j].y), s);
    s = fmaf(qv[3], bf16_hi(G.k[j].y), s);
    s = fmaf(qv[4], bf16_lo(G.k[j].z), s);
    s = fmaf(qv[5], bf16_hi(G.k[j].z), s);
    s = fmaf(qv[6], bf16_lo(G.k[j].w), s);
    s = fmaf(qv[7], bf16_hi(G.k[j].w), s);
#else   // A/B: even / odd dims in two chains (different summation order: NOT bit-identical to the scalar chain)
    float s = qv[0] * bf16_lo(G.k[j].x), s1 = qv[1] * bf16_hi(G.k[j].x);
    ffma2(s, s1, qv[2], qv[3], bf16_lo(G.k[j].y), bf16_hi(G.k[j].y));
    ffma2(s, s1, qv[4], qv[5], bf16_lo(G.k[j].z), bf16_hi(G.k[j].z));
    ffma2(s, s1, qv[6], qv[7], bf16_lo(G.k[j].w), bf16_hi(G.k[j].w));
    s += s1;
#endif
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    sc[j] = (4 * j + rg < G.cnt) ? s : -INFINITY;
    cm = fmaxf(cm, sc[j]);
  }
  cm = fmaxf(cm, __shfl_xor_sync(0xffffffffu, cm, 8));
  cm = fmaxf(cm, __shfl_xor_sync(0xffffffffu, cm, 16));
  const float m_new = fmaxf(m, cm);                 // finite: cnt >= 1
  const float scale = exp2f(m - m_new);             // m == -inf -> 0
  l *= scale;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] *= scale;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float pj = exp2f(sc[j] - m_new);          // masked rows: exp2(-inf) = 0
    l += pj;
#ifdef TTS_ATTN_NO_FFMA2
    acc[0] = fmaf(pj, bf16_lo(G.v[j].x), acc[0]);
    acc[1] = fmaf(pj, bf16_hi(G.v[j].x), acc[1]);
    acc[2] = fmaf(pj, bf16_lo(G.v[j].y), acc[2]);
    acc[3] = fmaf(pj, bf16_hi(G.v[j].y), acc[3]);
    acc[4] = fmaf(pj, bf16_lo(G.v[j].z), acc[4]);
    acc[5] = fmaf(pj, bf16_hi(G.v[j].z), acc[5]);
    acc[6] = fmaf(pj, bf16_lo(G.v[j].w), acc[6]);
    acc[7] = fmaf(pj, bf16_hi(G.v[j].w), acc[7]);
#else   // the same eight fma.rn as four paired instructions (bit-identical)
    ffma2(acc[0], acc[1], pj, pj, bf16_lo(G.v[j].x), bf16_hi(G.v[j].x));
    ffma2(acc[2], acc[3], pj, pj, bf16_lo(G.v[j].y), bf16_hi(G.v[j].y));
    ffma2(acc[4], acc[5], pj, pj, bf16_lo(G.v[j].z), bf16_hi(G.v[j].z));
    ffma2(acc[6], acc[7], pj, pj, bf16_lo(G.v[j].w), bf16_hi(G.v[j].w));
#endif
  }
  m = m_new;
}

// A row's positions are ALWAYS cut into 4 quarter ranges of 16-position granules, each reduced to its own partial
// (max, sum, acc) and merged in quarter order, whatever the number of warps per row: warp w streams the quarters
// (w % WPR) * 4 / WPR .. of row w / WPR back to back (the loads run on across quarter boundaries), so the result of a
// row is bit-identical in every MODE.
// ctx_row0: row index (inside the cluster's m-tiles) of attention row 0 of this CTA
template <int MODE>
__device__ void attn_local(const Params& P, int layer, int head, int attn_row0, int ctx_row0, const Smem& sm, uint32_t q_phase) {
  using C = DecCfg<MODE>;
  constexpr int WPR = C::kWarpsPerRow;
  constexpr int QPW = 4 / WPR;                        // quarters per warp: 1, 2, 4
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int rg = lane >> 3, dc = lane & 7;
  const int* ri = rinfo_ptr(sm);
  float* part = sm.red;                               // [attention row][quarter][kPartStride]
  {
    const int j = w / WPR, s = w % WPR, row = attn_row0 + j;
    const int len = ri[64 + j], plen = ri[72 + j];
    float m = -INFINITY, l = 0.f, acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    int q = s * QPW;                                  // quarter being accumulated
    auto flush = [&]() {                              // finish quarter q: reduce over the 4 row groups, park the partial, reset
      float ll = l;
      ll += __shfl_xor_sync(0xffffffffu, ll, 8);
      ll += __shfl_xor_sync(0xffffffffu, ll, 16);
      float* pp = part + (j * 4 + q) * kPartStride;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 8);
        acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 16);
      }
      if (rg == 0) {
        *reinterpret_cast<float4*>(pp + 8 + dc * 8) = make_float4(acc[0], acc[1], acc[2], acc[3]);
        *reinterpret_cast<float4*>(pp + 8 + dc * 8 + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
      }
      if (lane == 0) { pp[0] = m; pp[1] = ll; }
      m = -INFINITY; l = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = 0.f;
      ++q;
    };
    if (len > 0) {
      const int Gp = (plen + 15) >> 4, Gt = Gp + ((len - plen + 15) >> 4);
      const int g0 = (s * QPW * Gt) >> 2, g1 = ((s + 1) * QPW * Gt) >> 2;
      if (g1 > g0) {
        const int* pt = ri + 80 + 64 * j;
#if defined(TTS_ATTN_DIAG) && TTS_ATTN_DIAG == 4   // timing diagnostic only (results are NOT valid): every layer reads layer 0's K/V (L2-resident)
        const __nv_bfloat16* kvb = P.kv + kv_row_offset(0, P.n_pages, 0, 0, head, 0) + dc * 8;
#else
        const __nv_bfloat16* kvb = P.kv + kv_row_offset(layer, P.n_pages, 0, 0, head, 0) + dc * 8;
#endif
#ifndef TTS_ATTN_DEPTH
#define TTS_ATTN_DEPTH 2
#endif
        // granules per warp in the register ring (4 KiB each).  2 = the next granule's loads are in flight while one is reduced.
        // Same-box A/B on the final build of round 2 (profiles/r02_ab_attention_depth_final.log): depth 2 / 3 / 4 = 597.4 / 600.4 /
        // 606.2 us/step at 16 rows, 835.7 / 844.4 at 32, 1379 / 1418 at 64, depth 1 670.5 vs 658.6 (520 steps): the phase is not
        // short of bytes in flight, and 32 registers fewer per ring slot help everything around it.
        constexpr int NG = TTS_ATTN_DEPTH;
        Gran G[NG];
        // the granule that holds the position this layer's c_attn has just produced (always the last private one)
        const int g_new = QkvAsync<MODE>::on ? Gt - 1 : -2, p_new = (len - 1 - plen) & 15;
        const uint4* knew = reinterpret_cast<const uint4*>(sm.red + kKvNewOffsetFloats) + j * 8 + dc;
        const uint4* vnew = knew + C::kAttnRows * 8;
        auto load = [&](Gran& Gx, int g) { gran_load(Gx, kvb, pt, g, Gp, plen, len, rg, g == g_new ? p_new : -1, knew, vnew); };
        // (requesting the first granules of old positions before this wait was measured neutral: 622.6 vs 622.0 us/step)
        if (QkvAsync<MODE>::on && !mbar_wait(&sm.ctrl->qbar, q_phase, &sm.ctrl->abort_flag)) *reinterpret_cast<volatile int*>(P.watchdog) = 1;
        if (P.prof && layer == 3 && sm.ctrl->ibuf[2] && lane == 0) P.prof[10240 + blockIdx.x * 16 + 8 + w] = globaltimer_ns();
#pragma unroll
        for (int i = 0; i < NG - 1; ++i)
          if (g0 + i < g1) load(G[i], g0 + i);
        float qv[8];
        {
          float4 a, b;
          if (QkvAsync<MODE>::on) {
            const float4* qs = reinterpret_cast<const float4*>(sm.red + kQsOffsetFloats + j * kHD + dc * 8);
            a = qs[0]; b = qs[1];
          } else {
            const float* qp = P.q + (size_t)row * kD + head * kHD + dc * 8;
            a = ld_cg_f4(qp); b = ld_cg_f4(qp + 4);
          }
          qv[0] = a.x * kLog2e; qv[1] = a.y * kLog2e; qv[2] = a.z * kLog2e; qv[3] = a.w * kLog2e;
          qv[4] = b.x * kLog2e; qv[5] = b.y * kLog2e; qv[6] = b.z * kLog2e; qv[7] = b.w * kLog2e;
        }
        int qend = ((q + 1) * Gt) >> 2;               // first granule of the next quarter
        for (int g = g0; g < g1; g += NG) {
#pragma unroll
          for (int k = 0; k < NG; ++k) {
            if (g + k < g1) {
              if (g + k + NG - 1 < g1) load(G[(k + NG - 1) % NG], g + k + NG - 1);
              if (QPW > 1) while (g + k >= qend) { flush(); qend = ((q + 1) * Gt) >> 2; }
              gran_reduce(G[k], qv, rg, m, l, acc);
            }
          }
        }
      }
    }
    prof_mark(P, sm.ctrl, 12);
    if (P.prof && layer == 3 && sm.ctrl->ibuf[2] && lane == 0) P.prof[10240 + blockIdx.x * 16 + w] = globaltimer_ns();
    while (q < (s + 1) * QPW) flush();                 // the last quarter, and quarters without a granule (empty partials)
  }
  consumer_sync();
  if (w < C::kAttnRows) {
    const int len = ri[64 + w];
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
    // the split-K c_proj of all 4 CTAs reads this row: push the pair into everyone's copy
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(ctx_smem(sm)) + afrag_elem_offset(ctx_row0 + w, 2 * lane, kHD);
    const uint32_t pk = pack_bf16x2(o0, o1);
#ifndef TTS_CTX_SYNC
    // every store completes 4 bytes on the receiver's context barrier (armed by its thread 0 below): the split-K c_proj waits
    // for that barrier instead of a cluster barrier behind a release fence
#pragma unroll
    for (int d = 0; d < kClusterSize; ++d) st_async_u32(dst, d, pk, &sm.ctrl->cbar);
#else
#pragma unroll
    for (int d = 0; d < kClusterSize; ++d) st_cluster_u32(dst, d, pk);
#endif
  }
#ifndef TTS_CTX_SYNC
  if (threadIdx.x == 0) mbar_arrive_expect_tx(&sm.ctrl->cbar, kClusterSize * C::kAttnRows * kHD * 2);
#endif
  prof_mark(P, sm.ctrl, 13);
}

// Have the next cnt (<= 32) chunks of the ring landed?  Lane k probes chunk k's barrier (one try_wait for the warp
// instead of cnt serial ones, ~0.05 us each); if one has not, every lane blocks on each in turn (watchdogged).
// A landed chunk's bytes are in shared memory before its barrier phase flips, so any lane may read any chunk after.
__device__ __forceinline__ void wait_chunks(Ctrl* c, const Pipe& pipe, int cnt) {
  const int lane = threadIdx.x & 31;
  bool mine = true;
  if (lane < cnt) {
    uint32_t sl = pipe.slot + (uint32_t)lane, ph = pipe.phase;
    if (sl >= (uint32_t)pipe.n_slots) { sl -= (uint32_t)pipe.n_slots; ph ^= 1u; }
    mine = mbar_try_wait(&c->full[sl], ph);
  }
  if (!__all_sync(0xffffffffu, mine)) {
    Pipe p2 = pipe;
    for (int t = 0; t < cnt; ++t) {
      mbar_wait(&c->full[p2.slot], p2.phase, &c->abort_flag);
      p2.advance();
    }
  }
}

// ---- GEMM of MT 16-row m-tiles over this CTA's next `cnt` (<= NT) streamed K=1024 chunks ------------------------------
// A = the LayerNorm'd rows staged in sm.a_rm; warp w owns k16-blocks 8w..8w+7 of every chunk (its A fragments
// are loaded once), the 8 partial tiles are summed through shared memory in a fixed order (deterministic), one m-tile per
// round.  HALF (MT == 1 only): only rows [row_lo, row_lo + 8) are produced (MODE 1 c_attn: the other half belongs to the
// sibling cluster).  epi(row, col, v0, v1) receives the fp32 sums of columns col, col+1 (col = 8 * tile_of(i) + even
// offset) of row `row` of the A operand (0 .. 16 MT - 1).  `bias` (indexed by output column) is fetched before the mma
// loop for the (at most two) pairs a thread finishes.  `commit`: the chunks are done with (false: another row group of
// the same stage re-reads them).
template <int NT, int MT, bool HALF, bool TAILSYNC, typename TileFn, typename Epi>
__device__ __forceinline__ void gemm16(const Params& P, const Smem& sm, Pipe& pipe, Feeder& fd, int cnt, int row_lo, const float* bias,
                                       bool commit, TileFn tile_of, Epi epi) {
  static_assert(!HALF || MT == 1, "half tiles are a MODE 1 notion");
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  Ctrl* c = sm.ctrl;
  constexpr int kPerTile = HALF ? 32 : 64;
  constexpr int kMaxE = (NT * kPerTile + kConsumerThreads - 1) / kConsumerThreads;
  constexpr int CH = 2;                             // two accumulator chains per (tile, m-tile) in EVERY mode: a row's fp32 summation
                                                    // order does not depend on how many rows share the launch (bit-identical results)
  float2 bpre[kMaxE];
#pragma unroll
  for (int k = 0; k < kMaxE; ++k) {
    const int e = threadIdx.x + k * kConsumerThreads;
    bpre[k] = make_float2(0.f, 0.f);
    if (e < cnt * kPerTile) {
      const int t = e / kPerTile, rem = e - t * kPerTile;
      const int ln = HALF ? rem : (rem >> 1);
      bpre[k] = *reinterpret_cast<const float2*>(bias + tile_of(t) * 8 + 2 * (ln & 3));
    }
  }
  uint4 a[MT][8];
  {
    const uint32_t base = smem_u32(sm.a_rm) + (uint32_t)(((lane & 7) + ((lane >> 3) & 1) * 8) * kARowBytes) +
                          (uint32_t)((lane >> 4) * 16) + (uint32_t)(w * 8 * 32);
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int i = 0; i < 8; ++i)
        ldmatrix_x4(base + (uint32_t)(mt * 16 * kARowBytes) + (uint32_t)(i * 32), a[mt][i].x, a[mt][i].y, a[mt][i].z, a[mt][i].w);
  }
  float acc[NT][MT][CH][4];
#pragma unroll
  for (int t = 0; t < NT; ++t)
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int ch = 0; ch < CH; ++ch)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[t][mt][ch][j] = 0.f;
  prof_mark(P, c, 5);
  wait_chunks(c, pipe, cnt);     // all chunks of the pass have normally landed long ago
  Pipe p2 = pipe;
  prof_mark(P, c, 6);
#pragma unroll
  for (int t = 0; t < NT; ++t)
    if (t < cnt) {
      const uint4* bsm = reinterpret_cast<const uint4*>(sm.ring + (size_t)p2.slot * kChunkBytes) + (w * 4) * 32 + lane;
#pragma unroll
      for (int s2 = 0; s2 < 4; ++s2) {
        const uint4 b = bsm[s2 * 32];
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          mma_bf16_16816(acc[t][mt][s2 & (CH - 1)], a[mt][2 * s2], b.x, b.y);
          mma_bf16_16816(acc[t][mt][s2 & (CH - 1)], a[mt][2 * s2 + 1], b.z, b.w);
        }
      }
      p2.advance();
    }
  if (commit) pipe = p2;
  prof_mark(P, c, 2);
  float4* red4 = reinterpret_cast<float4*>(sm.red);
  const float2* red2 = reinterpret_cast<const float2*>(sm.red);
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
    for (int t = 0; t < NT; ++t)
      if (t < cnt)
        red4[(w * NT + t) * 32 + lane] = make_float4(acc[t][mt][0][0] + acc[t][mt][1][0], acc[t][mt][0][1] + acc[t][mt][1][1],
                                                     acc[t][mt][0][2] + acc[t][mt][1][2], acc[t][mt][0][3] + acc[t][mt][1][3]);
    consumer_sync();
    prof_mark(P, c, 4);
#pragma unroll
    for (int q = 0; q < kMaxE; ++q) {
      const int e = threadIdx.x + q * kConsumerThreads;
      if (e >= cnt * kPerTile) break;
      const int t = e / kPerTile, rem = e - t * kPerTile;
      const int ln = HALF ? rem : (rem >> 1);
      const int hh = HALF ? (row_lo >> 3) : (rem & 1);
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int k = 0; k < kNW; ++k) {
        const float2 v = red2[((k * NT + t) * 32 + ln) * 2 + hh];
        s0 += v.x;
        s1 += v.y;
      }
      epi(16 * mt + (ln >> 2) + 8 * hh, tile_of(t) * 8 + 2 * (ln & 3), s0 + bpre[q].x, s1 + bpre[q].y);
    }
    if (TAILSYNC || mt + 1 < MT) consumer_sync();   // else: the caller's next cluster / grid barrier protects the reduction buffer
  }
  if (commit) fd.pending += cnt;         // every warp is done with the cnt ring slots; refilled at the next grid barrier
  prof_mark(P, c, 3);
}

// ---- split-K attn.c_proj: y[rows, 256r .. 256r+256) += ctx[rows, 64h .. 64h+64) . W[64h .. 64h+64, 256r ..] ------
// rows = the rows of its head this cluster owns (MODE 1: half of the single m-tile, the other half of the A fragments is
// the sibling cluster's and is ignored).  Two 16 KiB chunks of 128 columns; warp w: n8-tiles 2w, 2w+1 of each.
// The residual and the bias are already in y (left there by the previous stage), so all 16 heads do the same work.
template <int MODE>
__device__ void proj_splitk(const Params& P, int rank, int row_lo, int tile_row0, const Smem& sm, Pipe& pipe) {
  using C = DecCfg<MODE>;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  Ctrl* c = sm.ctrl;
  uint4 a[C::kMtA][4];                             // A fragments of the head's context: kb = 0..3 of the local copy
#pragma unroll
  for (int mt = 0; mt < C::kMtA; ++mt)
#pragma unroll
    for (int j = 0; j < 4; ++j) a[mt][j] = ctx_smem(sm)[(mt * 4 + j) * 32 + lane];
  const int g = lane >> 2, t = lane & 3;
  // MODE 1 / 2: the contributions of the cluster's rows go to y as ONE bulk reduction per row (cp.reduce.async.bulk .add.u64:
  // 2 KiB = this CTA's 256 columns of a row) from a staging tile in the A-operand buffer, instead of 2048 (4096) 64-bit red.global
  // per CTA -- an SM issues those at ~1.3 clocks per lane, which was ~1 us of every block's attention -> ln_2 hand-off.  Same
  // integer adds at L2, same words (value + count): bit-identical.  The staging row of global row R is row R of the A-operand
  // buffer: ln_1's rows went into registers at c_attn, and ln_2 (this CTA's or a peer's distributed-shared store) writes row R
  // only after y's row R is complete, i.e. after this reduction has read its source.
#ifdef TTS_PROJ_RED
  constexpr bool kBulk = false;
#else
  constexpr bool kBulk = C::kMtA == 1;
#endif
  constexpr int kStageRow = kARowBytes;
  wait_chunks(c, pipe, 2);
#pragma unroll
  for (int ch = 0; ch < 2; ++ch) {
    const int col0 = 256 * rank + 128 * ch + 16 * w + 2 * t;       // tile tt adds 8
    float acc[C::kMtA][2][4];
#pragma unroll
    for (int mt = 0; mt < C::kMtA; ++mt)
#pragma unroll
      for (int tt = 0; tt < 2; ++tt)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[mt][tt][j] = 0.f;
    {
      const uint4* bsm = reinterpret_cast<const uint4*>(sm.ring + (size_t)pipe.slot * kChunkBytes) + lane;
#pragma unroll
      for (int tt = 0; tt < 2; ++tt) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const uint4 b = bsm[(((2 * w + tt) * 2) + s) * 32];
#pragma unroll
          for (int mt = 0; mt < C::kMtA; ++mt) {
            mma_bf16_16816(acc[mt][tt], a[mt][2 * s], b.x, b.y);
            mma_bf16_16816(acc[mt][tt], a[mt][2 * s + 1], b.z, b.w);
          }
        }
      }
    }
    pipe.advance();
#pragma unroll
    for (int mt = 0; mt < C::kMtA; ++mt)
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        if (C::kHalf && hh != (row_lo >> 3)) continue;             // the other half of the tile is the sibling cluster's
        const int row = tile_row0 + 16 * mt + 8 * hh + g;
        if (kBulk) {
          const int lr = row_lo + (C::kHalf ? 0 : 8 * hh) + g;         // == the global row (phase-B row groups start at 0 in these modes)
#pragma unroll
          for (int tt = 0; tt < 2; ++tt)
            *reinterpret_cast<longlong2*>(sm.a_rm + (size_t)lr * kStageRow + (size_t)(128 * ch + 16 * w + 2 * t + 8 * tt) * 8) =
                make_longlong2(cnt_contrib(acc[mt][tt][2 * hh]), cnt_contrib(acc[mt][tt][2 * hh + 1]));
          continue;
        }
        if (row >= P.n_rows) continue;
#pragma unroll
        for (int tt = 0; tt < 2; ++tt) {
          long long* yp = P.y + (size_t)row * kD + col0 + 8 * tt;
          red_add_s64(yp, cnt_contrib(acc[mt][tt][2 * hh]));          // value and contribution count in one atomic
          red_add_s64(yp + 1, cnt_contrib(acc[mt][tt][2 * hh + 1]));
        }
      }
  }
  if (kBulk) {
    fence_proxy_async();                 // the staged words (generic-proxy stores) are read by the bulk engine
    consumer_sync();
    if (w == 0 && lane < C::kRowsA) {
      const int row = row_lo + lane;
      if (row < P.n_rows) bulk_red_add_u64(P.y + (size_t)row * kD + 256 * rank, sm.a_rm + (size_t)row * kStageRow, 2048u);
    }
  }
}

// ---- mlp.c_proj of row group `grp`: xt = (y + h . W + bias, tag) ; y = x + next layer's attn.c_proj bias, count 1 --------
// A cluster owns 4 n8-tiles (32 output columns); CTA rank r multiplies the K range [1024r, 1024r+1024) of h with the 4
// tiles (4 streamed chunks), so it reads only its quarter of h (32 KiB of bf16 A fragments per m-tile from L2 instead of
// 128 KiB: the SM's ~55 GB/s L2 read port is what this stage waits for).  The four K-partials of a tile are sent to the
// tile's owner (rank t owns tile t) through distributed shared memory and summed there in rank order (deterministic).
template <int MODE>
__device__ void proj2_stage(const Params& P, int layer, const Smem& sm, Pipe& pipe, Feeder& fd, int rank, int grp, bool commit, uint32_t xtag,
                            uint32_t& x_phase) {
  using C = DecCfg<MODE>;
  constexpr int MT = C::kMtB;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  Ctrl* c = sm.ctrl;
  uint4 a[MT][8];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) {
    const uint4* a_gl = reinterpret_cast<const uint4*>(P.hbuf) + ((size_t)(grp * MT + mt) * (kD4 >> 4) + rank * kKbPerChunk + w * 8) * 32 + lane;
#pragma unroll
    for (int i = 0; i < 8; ++i) a[mt][i] = ld_cg_u4(a_gl + (size_t)i * 32);
  }
  // epilogue operands of the threads that own an output pair of tile `rank` (64 pairs per m-tile: 16 rows x 4 column pairs)
  const int emt = threadIdx.x >> 6, erem = threadIdx.x & 63;
  const int col = ((blockIdx.x & ~3) + rank) * 8 + 2 * ((erem >> 1) & 3);
  const int erow = 16 * (grp * MT + emt) + (erem >> 3) + 8 * (erem & 1);
  const bool owner = emt < MT && erow < P.n_rows;
  longlong2 old = make_longlong2(0ll, 0ll);
  float2 bias = make_float2(0.f, 0.f), nbias = make_float2(0.f, 0.f);
  if (owner) {
    old = ld_cg_ll2(P.y + (size_t)erow * kD + col);
    bias = *reinterpret_cast<const float2*>(layer_f(P, layer) + kFOffBproj2 + col);
    if (layer + 1 < P.n_layers) nbias = *reinterpret_cast<const float2*>(layer_f(P, layer + 1) + kFOffBproj + col);
  }
  float acc[4][MT][2][4];
#pragma unroll
  for (int t = 0; t < 4; ++t)
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[t][mt][0][j] = acc[t][mt][1][j] = 0.f;
  wait_chunks(c, pipe, 4);
  Pipe p2 = pipe;
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const uint4* bsm = reinterpret_cast<const uint4*>(sm.ring + (size_t)p2.slot * kChunkBytes) + (w * 4) * 32 + lane;
#pragma unroll
    for (int s2 = 0; s2 < 4; ++s2) {
      const uint4 b = bsm[s2 * 32];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        mma_bf16_16816(acc[t][mt][s2 & 1], a[mt][2 * s2], b.x, b.y);
        mma_bf16_16816(acc[t][mt][s2 & 1], a[mt][2 * s2 + 1], b.z, b.w);
      }
    }
    p2.advance();
  }
  if (commit) pipe = p2;
  float4* red4 = reinterpret_cast<float4*>(sm.red);
  float2* xred = reinterpret_cast<float2*>(sm.red + kXredOffsetFloats) + (size_t)grp * MT * kClusterSize * 64;
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) {
    if (mt) consumer_sync();             // the previous round's sums have been read
#pragma unroll
    for (int t = 0; t < 4; ++t)
      red4[(w * 4 + t) * 32 + lane] = make_float4(acc[t][mt][0][0] + acc[t][mt][1][0], acc[t][mt][0][1] + acc[t][mt][1][1],
                                                  acc[t][mt][0][2] + acc[t][mt][1][2], acc[t][mt][0][3] + acc[t][mt][1][3]);
    consumer_sync();
    // this CTA's K-partial of pair (tile t, ln, hh) -> owner CTA t
    const float2* red2 = reinterpret_cast<const float2*>(sm.red);
    const int t = threadIdx.x >> 6, rem = threadIdx.x & 63, ln = rem >> 1, hh = rem & 1;
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int k = 0; k < kNW; ++k) {
      const float2 v = red2[((k * 4 + t) * 32 + ln) * 2 + hh];
      s0 += v.x;
      s1 += v.y;
    }
#ifndef TTS_X_SYNC
    // counted on the owner's barrier of this row group (st.async): no release fence, no cluster barrier.  A peer is at most one
    // row group ahead (its next wait needs this CTA's next partials), so two alternating barriers keep the phases apart.
    st_async_u2(xred + (mt * kClusterSize + rank) * 64 + rem, (uint32_t)t, make_uint2(__float_as_uint(s0), __float_as_uint(s1)), &c->xbar[grp & 1]);
#else
    st_cluster_u2(xred + (mt * kClusterSize + rank) * 64 + rem, (uint32_t)t, make_uint2(__float_as_uint(s0), __float_as_uint(s1)));
#endif
  }
  if (commit) fd.pending += 4;
#ifndef TTS_X_SYNC
  if (threadIdx.x == 0) mbar_arrive_expect_tx(&c->xbar[grp & 1], kClusterSize * MT * 64 * 8);
  if (!mbar_wait(&c->xbar[grp & 1], (x_phase >> (grp & 1)) & 1u, &c->abort_flag)) {
    *reinterpret_cast<volatile int*>(P.watchdog) = 1;
    return;
  }
  x_phase ^= 1u << (grp & 1);
#else
  cluster_barrier();
#endif
  if (owner) {
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int r = 0; r < kClusterSize; ++r) {
      const float2 v = xred[(emt * kClusterSize + r) * 64 + erem];
      s0 += v.x;
      s1 += v.y;
    }
    const float x0 = cnt_value(old.x) + s0 + bias.x, x1 = cnt_value(old.y) + s1 + bias.y;
    // the residual stream for the next block, self-validating: (value, tag) pairs polled by its ln_1 (no grid barrier)
    st_tagged2(P.xt + (size_t)erow * kD + col, x0, x1, xtag);
    // the next layer's attention block accumulates onto residual + attn.c_proj bias, contribution count 1.  Written as an
    // atomic ADD of (target - what was read): it commutes with the next block's head contributions, so no ordering between
    // this CTA and the clusters that are already ahead is needed.
    long long* yp = P.y + (size_t)erow * kD + col;
    red_add_s64(yp, cnt_word(x0 + nbias.x, 1) - old.x);
    red_add_s64(yp + 1, cnt_word(x1 + nbias.y, 1) - old.y);
  }
}

// ---- grid-wide barrier of this kernel ---------------------------------------------------------------------------------
// Same counter protocol as ar_core.cuh grid_barrier.  Warp 0 also refills ring slots right after the arrival.
// Acquire side: thread 0 polls the counter with ld.acquire.gpu; the poll that returns the full count synchronizes-with
// the red.release of every producer CTA, and the bar.sync that follows extends that to all threads of this CTA
// (PTX memory model: causality order through the CTA barrier).  Same-box A/B on C2 (profiles/r02_ab_grid_barrier_acquire.log):
// relaxed poll without any acquire 629.9 us/step (round 1's variant, formally racy), relaxed poll + fence.acq_rel.gpu
// 678.7, acquire poll 635.6.
// `stage` (optional): `nstage` contiguous fp32 vectors of kD floats (LayerNorm gains / biases of the next phase) that
// warps 4..7 copy into the reduction buffer (at float offset stage_off) while thread 0 polls, so the LayerNorm after
// the barrier reads them from shared memory instead of paying an L2 round trip between its statistics and its output.
__device__ __forceinline__ bool dec_grid_barrier(const Params& P, const Smem& sm, uint32_t& epoch, Feeder& fd, int extra,
                                                 int prof_slot = -1, const float* stage = nullptr, int nstage = 0, int stage_off = 0) {
  Ctrl* c = sm.ctrl;
  // bring-up: arrival / release time of every CTA at the three barriers of one layer (scripts/stage_prof.py)
  const bool rec = P.prof && prof_slot >= 0 && threadIdx.x == 0;
  if (rec) P.prof[8192 + blockIdx.x * 8 + prof_slot] = globaltimer_ns();
  consumer_sync();
  ++epoch;
  // at most 6 chunks (96 KiB) per barrier: a larger burst is still arriving when the LayerNorm loads that follow
  // the barrier are issued, and those then queue behind it
  fd.pending += extra;
  const int refill_n = min(fd.pending, 6);
  fd.pending -= refill_n;
  if (threadIdx.x < 32) {
    if (threadIdx.x == 0) red_release_add_u32(P.bar_counter, 1u);
    __syncwarp();
    if (refill_n) feed(P, c, sm.ring, fd.f, refill_n, fd.head, fd.rank, fd.ht0, fd.hcnt);
  }
  if (threadIdx.x >= 128 && nstage) {
    const int t = threadIdx.x - 128;                  // 0..127: 16-byte pieces t, t + 128, ...
#pragma unroll 1
    for (int k = t; k < nstage * (kD / 4); k += 128) cp_async_cg16(sm.red + stage_off + 4 * k, stage + 4 * k);
    cp_async_wait_all();
  }
  if (threadIdx.x == 0) {
    const uint32_t target = epoch * (uint32_t)P.G;
    const long long t0 = clock64();
    uint32_t spins = 0;
#ifdef TTS_DEC_RELAXED_POLL   // compile-time A/B only (scripts/ab_build.sh): no acquire at all, relies on every cross-CTA read being ld.cg
    while (ld_relaxed_u32(P.bar_counter) < target) {
#else
    while (ld_acquire_u32(P.bar_counter) < target) {
#endif
      if ((++spins & 255u) == 0) {
        if (c->abort_flag || __ldcg(P.watchdog) || clock64() - t0 > kWatchdogCycles) {
          c->abort_flag = 1;
          *reinterpret_cast<volatile int*>(P.watchdog) = 1;
          break;
        }
      }
    }
  }
  consumer_sync();
  if (rec) P.prof[8192 + blockIdx.x * 8 + prof_slot + 3] = globaltimer_ns();
  return c->abort_flag == 0;
}

// ---- phase boundary WITHOUT a grid barrier: the next stage polls its (self-validating) input itself ------------------
// What is left of the barrier is CTA-local: every warp is done with the previous stage's ring slots and reduction buffer,
// warp 0 refills the ring, warps 4..7 stage the next LayerNorm's gains / biases.
__device__ __forceinline__ bool dec_soft_boundary(const Params& P, const Smem& sm, Feeder& fd, int extra, const float* stage, int nstage,
                                                  int stage_off) {
  Ctrl* c = sm.ctrl;
  consumer_sync();
  fd.pending += extra;
  const int refill_n = min(fd.pending, 6);
  fd.pending -= refill_n;
  if (threadIdx.x < 32 && refill_n) feed(P, c, sm.ring, fd.f, refill_n, fd.head, fd.rank, fd.ht0, fd.hcnt);
  if (threadIdx.x >= 128 && nstage) {
    const int t = threadIdx.x - 128;
#pragma unroll 1
    for (int k = t; k < nstage * (kD / 4); k += 128) cp_async_cg16(sm.red + stage_off + 4 * k, stage + 4 * k);
    cp_async_wait_all();
  }
  consumer_sync();
  return c->abort_flag == 0;
}

// ---- the kernel ------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(kDecThreads, 1) ar_decode_kernel(const __grid_constant__ Params P) {
  using C = DecCfg<MODE>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const Smem sm = carve_smem(smem_raw, C::kMtSmem);
  Ctrl* c = sm.ctrl;
  const int rank = (int)cluster_ctarank();
  const int head = (int)cluster_id_x() >> 1;
  const int row_lo = ((int)cluster_id_x() & 1) * C::kRowsA;       // first row of its head this cluster owns
  const int tile_row0 = C::kHalf ? 0 : row_lo;                    // first row of the m-tile(s) it multiplies
  const int attn_row0 = row_lo + C::kAttnRows * rank;             // first attention row of this CTA
  const int ctx_row0 = attn_row0 - tile_row0;
  Feeder fd;
  fd.head = head; fd.rank = rank; fd.pending = 0;
  tile_range(P.v_tiles, P.G, 0, blockIdx.x, &fd.ht0, &fd.hcnt);
  if (threadIdx.x == 0) {
    for (int i = 0; i < P.n_slots; ++i) mbar_init(&c->full[i], 1);
    mbar_init(&c->cbar, 1);
    mbar_init(&c->qbar, 1);
    mbar_init(&c->xbar[0], 1);
    mbar_init(&c->xbar[1], 1);
    c->abort_flag = 0;
    c->ibuf[0] = 0;
    c->ibuf[1] = 0;
    fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    feed_init(P, fd.f, fd.hcnt, C::kGroups);
    feed(P, c, sm.ring, fd.f, P.n_slots, head, rank, fd.ht0, fd.hcnt);
  }
  cluster_sync_all();          // all CTAs of the cluster are running before anyone writes into a peer's shared memory

  Pipe pipe;
  pipe.ci = 0; pipe.slot = 0; pipe.phase = 0; pipe.n_slots = P.n_slots;
  uint32_t epoch = 0;
  uint32_t ctx_phase = 0, q_phase = 0, x_phase = 0;
  int steps_done = 0;
  bool alive = true;
  const int n = P.n_rows;
  const int ht0 = fd.ht0, hcnt = fd.hcnt;
  constexpr int kNpcB = 4 * C::kMtB;            // rows per CTA in the LayerNorm of a phase-B / head row group
  constexpr int kNpcA = C::kLnRowsA / kClusterSize;
  Instr ins;
  ins.row0 = 0; ins.nrows = n; ins.flags = 0; ins.rot = 0; ins.pad0 = ins.pad1 = 0;
  for (int step = 0; step < P.n_steps && alive; ++step) {
    if (blockIdx.x == 0 && threadIdx.x == 0) c->ibuf[1] = (P.prof && step == P.prof_step) ? 1 : 0;
    if (threadIdx.x == 0) c->ibuf[2] = (P.prof && step == P.prof_step) ? 1 : 0;   // every CTA: per-warp attention times of layer 3
    prof_mark(P, c, 100 + OP_SAMPLE);
    alive = dec_grid_barrier(P, sm, epoch, fd, 0);
    if (!alive) break;
    prof_mark(P, c, 200 + OP_SAMPLE);
    ins.op = OP_SAMPLE; ins.layer = 0; ins.nrows = P.n_seqs;
    sample_stage(P, ins, sm, step);
    ins.nrows = n;
    steps_done = step + 1;
    if (P.skip_last_model && step == P.n_steps - 1) break;
    // layer == n_layers is the head: ln_f + final_norm through the same LayerNorm code as ln_1, then the vocabulary GEMM
    for (int layer = 0; layer <= P.n_layers && alive; ++layer) {
      const bool is_head = layer == P.n_layers;
      // ---- phase A: ln_1 -> q,k,v of this head -> attention -> split-K c_proj -> y ----
      prof_mark(P, c, 100 + (is_head ? OP_HEAD : OP_LN_QKV));
      const int pslot = (step == P.prof_step && layer == 3) ? 0 : -4;
      // gains / biases to stage: ln_1 (g, b), or lnf_g, lnf_b, fn_g, fn_b (contiguous, tts_ar.cu) for the head.
      // Layer 0 follows the sampler (plain x, step bookkeeping): a real grid barrier.  Every later block, and the head, POLLS
      // the tagged residual stream the previous block's mlp.c_proj owners write (ln_cluster LN_SRC_XT): no grid barrier.
      const uint32_t xtag = (uint32_t)(step * (P.n_layers + 1) + layer);      // >= 1 for layer >= 1; xt is zeroed per launch
      if (layer == 0)
        alive = dec_grid_barrier(P, sm, epoch, fd, 0, pslot, layer_f(P, layer) + kFOffLn1g, 2);
      else
        alive = dec_soft_boundary(P, sm, fd, 0, is_head ? P.lnf_g : layer_f(P, layer) + kFOffLn1g, is_head ? 4 : 2, 0);
      if (!alive) break;
      prof_mark(P, c, 200 + (is_head ? OP_HEAD : OP_LN_QKV));
      if (layer == 0 && ld_cg_i(P.n_active) == 0) { alive = false; break; }
      {
        // row facts of the step (layer 0) and the L2 prefetch of the next layer's K/V segments and small parameters (a handful
        // of bulk-prefetch instructions): by the warps the LayerNorm leaves idle (4 rows per CTA: warps 4..7), else by
        // everyone (row facts before the LayerNorm, prefetch after its stores)
        const int pf_layer = is_head ? 0 : layer + 1;
        const int ln_warps = is_head ? kNpcB : kNpcA;
        const bool side = ln_warps <= 4;
        if (side) {
          if (threadIdx.x >= 128) {
            if (layer == 0) {
              build_rinfo<MODE>(P, tile_row0, attn_row0, sm, threadIdx.x - 128, 128);
              asm volatile("bar.sync 2, 128;" ::: "memory");
            }
            if (MODE != 4) prefetch_next(P, sm, head, pf_layer, C::kAttnRows, threadIdx.x - 128, 128);
          }
        } else if (layer == 0) {
          build_rinfo<MODE>(P, tile_row0, attn_row0, sm, threadIdx.x, kDecThreads);
          consumer_sync();
        }
        // ln_1: the rows this cluster owns (MODE 1: rows row_lo.. of the single m-tile, placed at the same rows of the A operand)
        const int ln_row0 = (is_head || C::kLnRowsA == 16 * C::kMtA) ? (is_head ? 0 : tile_row0) : row_lo;
        ln_cluster(P, sm, layer == 0 ? LN_SRC_X : LN_SRC_XT, xtag, sm.red, is_head, rank, ln_row0, ln_warps, ln_row0 - (is_head ? 0 : tile_row0));
        if (!side && MODE != 4) prefetch_next(P, sm, head, pf_layer, C::kAttnRows, threadIdx.x, kDecThreads);
      }
      cluster_barrier_at<1>();
      prof_mark(P, c, 1);
      if (is_head) break;
      if (QkvAsync<MODE>::on && threadIdx.x == 0)      // q (256 B) + new K + new V (128 B each) per attention row
        mbar_arrive_expect_tx(&c->qbar, C::kAttnRows * (kHD * 4 + 2 * kHD * 2));
      {
        const float* bias = layer_f(P, layer) + kFOffBqkv;
        const int* ri = rinfo_ptr(sm);
        const int t0 = head * 8 + 2 * rank;
        gemm16<6, C::kMtA, C::kHalf, false>(P, sm, pipe, fd, 6, row_lo, bias, true, [&](int i) { return t0 + (i >> 1) * (kD / 8) + (i & 1); },
                        [&](int lrow, int col, float v0, float v1) {
                          const int row = tile_row0 + lrow;
                          const int part = col >> 10, cc = col & (kD - 1);
                          if (QkvAsync<MODE>::on) {
                            // to the CTA that attends this row, ALWAYS (its barrier counts bytes, also for rows past the end)
                            const int rel = row - row_lo, dst = rel / C::kAttnRows, j = rel - dst * C::kAttnRows, d = cc & (kHD - 1);
                            if (part == 0) {
                              st_async_u2(sm.red + kQsOffsetFloats + j * kHD + d, (uint32_t)dst,
                                          make_uint2(__float_as_uint(v0 * 0.125f), __float_as_uint(v1 * 0.125f)), &c->qbar);
                            } else {
                              __nv_bfloat16* kvs = reinterpret_cast<__nv_bfloat16*>(sm.red + kKvNewOffsetFloats);
                              st_async_u32(kvs + ((part - 1) * C::kAttnRows + j) * kHD + d, (uint32_t)dst, pack_bf16x2(v0, v1), &c->qbar);
                            }
                          }
                          if (row >= n) return;
                          if (part == 0) {   // query, pre-scaled by 1/sqrt(64) (exact: power of two)
                            *reinterpret_cast<float2*>(P.q + (size_t)row * kD + cc) = make_float2(v0 * 0.125f, v1 * 0.125f);
                          } else {           // key / value -> page slot of this row's new position (row facts of the step)
                            const int page = ri[lrow];
                            if (page < 0) return;
                            __nv_bfloat16* dst = P.kv + kv_row_offset(layer, P.n_pages, page, part - 1, cc >> 6, ri[32 + lrow]) + (cc & 63);
                            *reinterpret_cast<uint32_t*>(dst) = pack_bf16x2(v0, v1);
                          }
                        });
      }
      if (!QkvAsync<MODE>::on) cluster_barrier_at<2>();      // MODE 1 / 2: attn_local waits for q / K / V on the CTA's own barrier
      prof_mark(P, c, 200 + OP_ATTN);
      attn_local<MODE>(P, layer, head, attn_row0, ctx_row0, sm, q_phase);
      q_phase ^= 1u;
#ifndef TTS_CTX_SYNC
      if (!mbar_wait(&c->cbar, ctx_phase, &c->abort_flag)) {
        *reinterpret_cast<volatile int*>(P.watchdog) = 1;
        alive = false;
        break;
      }
      ctx_phase ^= 1u;
#else
      cluster_barrier_at<4>();
#endif
      prof_mark(P, c, 200 + OP_PROJ);
      proj_splitk<MODE>(P, rank, row_lo, tile_row0, sm, pipe);
      // ---- phase B: ln_2 -> c_fc -> gelu_new ----
      prof_mark(P, c, 100 + OP_LN_FC);
      // no grid barrier: ln_2 polls the counting accumulator y until residual + 16 heads have landed (ln_cluster LN_SRC_Y)
      alive = dec_soft_boundary(P, sm, fd, 2, layer_f(P, layer) + kFOffLn2g, 2, kStageOffLn2);   // + the two split-K c_proj slots
      if (!alive) break;
      prof_mark(P, c, 200 + OP_LN_FC);
#pragma unroll 1
      for (int grp = 0; grp < C::kGroups; ++grp) {
        if (grp) cluster_barrier();       // every CTA of the cluster has multiplied the previous group's rows
        const int rbase = 16 * C::kMtB * grp;
        ln_cluster(P, sm, LN_SRC_Y, 1 + kH, sm.red + kStageOffLn2, false, rank, rbase, kNpcB);
        cluster_barrier_at<8>();
        prof_mark(P, c, 1);
        const float* bias = layer_f(P, layer) + kFOffBfc;
        const int t0 = blockIdx.x * 4;
        gemm16<4, C::kMtB, false, false>(P, sm, pipe, fd, 4, 0, bias, grp == C::kGroups - 1, [&](int i) { return t0 + i; },
                                         [&](int lrow, int col, float v0, float v1) {
                                           const int row = rbase + lrow;
                                           if (row >= n) return;
                                           const float h0 = gelu_new(v0), h1 = gelu_new(v1);
                                           *reinterpret_cast<uint32_t*>(P.hbuf + afrag_elem_offset(row, col, kD4)) = pack_bf16x2(h0, h1);
                                         });
      }
      // ---- phase C: mlp.c_proj, x = y + . ----
      prof_mark(P, c, 100 + OP_PROJ2);
      alive = dec_grid_barrier(P, sm, epoch, fd, 0, pslot + 2);
      if (!alive) break;
      prof_mark(P, c, 200 + OP_PROJ2);
#pragma unroll 1
      for (int grp = 0; grp < C::kGroups; ++grp) proj2_stage<MODE>(P, layer, sm, pipe, fd, rank, grp, grp == C::kGroups - 1, xtag + 1, x_phase);
      if (c->abort_flag) alive = false;
    }
    if (!alive) break;
    // ---- mel head: logits of every row group (group 0 was normalised by the layer loop's last iteration) ----
#pragma unroll 1
    for (int grp = 0; grp < C::kGroups; ++grp) {
      const int rbase = 16 * C::kMtB * grp;
      if (grp) {
        cluster_barrier();                // the previous group's rows have been multiplied by every CTA of the cluster
        // the staged ln_f / final_norm gains were overwritten by the reduction: stage them again
        for (int k = threadIdx.x; k < 4 * (kD / 4); k += kDecThreads) cp_async_cg16(sm.red + 4 * k, P.lnf_g + 4 * k);
        cp_async_wait_all();
        consumer_sync();
        ln_cluster(P, sm, LN_SRC_XT, (uint32_t)(step * (P.n_layers + 1) + P.n_layers), sm.red, true, rank, rbase, kNpcB);
        cluster_barrier();
      }
      for (int tb = 0; tb < hcnt; tb += 6) {
        const int t0 = ht0 + tb;
        gemm16<6, C::kMtB, false, true>(P, sm, pipe, fd, min(6, hcnt - tb), 0, P.b_head, true, [&](int i) { return t0 + i; },
                         [&](int lrow, int col, float v0, float v1) {
                           const int row = rbase + lrow;
                           if (row >= n) return;
                           float* lp = P.logits + (size_t)row * P.v_stride + col;     // b_head is zero-padded to v_stride
                           if (col < P.V) lp[0] = v0;
                           if (col + 1 < P.V) lp[1] = v1;
                         });
        // a CTA's head tiles (8 or 9 chunks, per row group) are more than the ring holds ahead at this point when it has 8
        // slots: refill here, not only at the next grid barrier (the reduction's trailing CTA barrier has passed: the
        // slots are free)
        const int refill_n = min(fd.pending, 6);
        fd.pending -= refill_n;
        if (threadIdx.x < 32 && refill_n) feed(P, c, sm.ring, fd.f, refill_n, fd.head, fd.rank, fd.ht0, fd.hcnt);
      }
    }
    if (c->abort_flag) alive = false;
  }
  consumer_sync();
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    *P.steps_done = steps_done;
    if (P.stream_done) {         // streaming: the launch has ended after steps_done steps (every flag of those steps was published before)
      __threadfence_system();
      *reinterpret_cast<volatile int*>(P.stream_done) = steps_done + 1;
    }
  }
  if (threadIdx.x == 0) {
    // never leave the kernel with bulk copies still landing in this CTA's shared memory
    for (uint32_t k = pipe.ci; k < fd.f.issued; ++k)
      if (!mbar_wait(&c->full[k % P.n_slots], (k / P.n_slots) & 1, &c->abort_flag)) break;
  }
  __syncthreads();
  cluster_sync_all();          // nobody exits while a peer could still address its shared memory
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

// ---- launch ----------------------------------------------------------------------------------------------------------
typedef void (*DecKernel)(const Params);
static DecKernel decode_kernel_of(int mode) {
  return mode == 1 ? (DecKernel)ar_decode_kernel<1> : mode == 2 ? (DecKernel)ar_decode_kernel<2> : (DecKernel)ar_decode_kernel<4>;
}

int decode_mode_for(int n_rows) { return n_rows <= 16 ? 1 : n_rows <= 32 ? 2 : 4; }
int decode_mt_smem(int mode) { return mode == 1 ? 1 : 2; }

static cudaError_t decode_config(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attrs, int mode, int smem_bytes, int device, bool cooperative,
                                 cudaStream_t st) {
  static unsigned long long attr_set[3] = {0, 0, 0};   // one bit per device ordinal: the opt-in is per device
  const int mi = mode == 1 ? 0 : mode == 2 ? 1 : 2;
  if (!(attr_set[mi] >> (device & 63) & 1ull)) {
    cudaError_t e = cudaFuncSetAttribute((const void*)decode_kernel_of(mode), cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
    if (e != cudaSuccess) return e;
    attr_set[mi] |= 1ull << (device & 63);
  }
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(kDecCtas);
  cfg.blockDim = dim3(kDecThreads);
  cfg.dynamicSmemBytes = (size_t)smem_bytes;
  cfg.stream = st;
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = kClusterSize;
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributeCooperative;
  attrs[1].val.cooperative = 1;
  cfg.attrs = attrs;
  // !cooperative: plain cluster launch (tts_ar_run.cu decode_cooperative: profilers cannot replay cooperative launches)
  cfg.numAttrs = cooperative ? 2 : 1;
  return cudaSuccess;
}

cudaError_t launch_decode(const Params& P, int mode, int smem_bytes, int device, bool cooperative, cudaStream_t st) {
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attrs[2];
  cudaError_t e = decode_config(cfg, attrs, mode, smem_bytes, device, cooperative, st);
  if (e != cudaSuccess) return e;
  void* args[] = {const_cast<Params*>(&P)};
  return cudaLaunchKernelExC(&cfg, (const void*)decode_kernel_of(mode), args);
}

// Clusters of kClusterSize CTAs (1 CTA / SM at this shared-memory size) that can be co-resident; the kernels need kDecClusters.
int decode_max_clusters(int mode, int smem_bytes, int device) {
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attrs[2];
  if (decode_config(cfg, attrs, mode, smem_bytes, device, true, 0) != cudaSuccess) { cudaGetLastError(); return 0; }
  cfg.numAttrs = 1;   // the occupancy query takes the cluster dimension only
  int n = 0;
  cudaError_t e = cudaOccupancyMaxActiveClusters(&n, (const void*)decode_kernel_of(mode), &cfg);
  if (e != cudaSuccess) {
    if (getenv("TTS_AR_VERBOSE")) fprintf(stderr, "[tts_ar] cudaOccupancyMaxActiveClusters: %s\n", cudaGetErrorString(e));
    cudaGetLastError();
    return 0;
  }
  return n;
}

}  // namespace tts
