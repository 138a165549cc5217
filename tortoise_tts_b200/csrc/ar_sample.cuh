// ar_sample.cuh -- logits processors, warpers, multinomial and the sample-loop bookkeeping, one CTA per
// sequence.  Restates, without a sort:
//   HF RepetitionPenaltyLogitsProcessor (over ALL input ids incl. the dummy prompt ids 1 and 8192 -- quirk Q4,
//   reference unified_voice.py:647-649), TemperatureLogitsWarper, TopKLogitsWarper, TopPLogitsWarper in the
//   order of the reference's _get_logits_warper (stream_generator.py:56-101), softmax + multinomial and the
//   pad-after-stop / unfinished bookkeeping of sample_stream (stream_generator.py:1160-1190).
// Top-k / top-p thresholds come from ONE search routine over the monotone uint image ("key") of the fp32 scores:
//   keep key >= T,  T = smallest key with  weight{key > T} < limit
//   top-k: weight = 1, limit = k  (T = k-th largest score, ties kept);
//   top-p: weight = exp(score - max), limit = p * Z  (token i is kept iff mass{score > score_i} < p, which is HF's
//   "ascending cumsum <= 1-p -> remove" rule; tests/test_oracle_golden.py proves the identity).
// Each round cuts the score interval in four with three probes and also counts the tokens left inside; once there
// are <= 32 the tokens themselves are gathered and a warp-wide sort + prefix sum ends the search.
// The multinomial is an inverse CDF in vocabulary order driven by Philox4x32-10 (oracle/philox.py).
#pragma once
#include "ar_core.cuh"

namespace tts {

constexpr int kVPerThread = (8448 + kConsumerThreads - 1) / kConsumerThreads;   // covers n_mel_codes <= 8448

__device__ __forceinline__ uint32_t key_of(float v) { return float_key(__float_as_uint(v)); }

// Sorts one 64-bit item per lane in descending order across the warp (bitonic, shuffles only).
__device__ __forceinline__ unsigned long long warp_sort_desc(unsigned long long v) {
  const int lane = threadIdx.x & 31;
#pragma unroll 1
  for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll 1
    for (int j = k >> 1; j > 0; j >>= 1) {
      const unsigned long long o = __shfl_xor_sync(0xffffffffu, v, j);
      const bool keep_max = ((lane & j) == 0) == ((lane & k) == 0);
      v = keep_max ? (v > o ? v : o) : (v < o ? v : o);
    }
  }
  return v;
}

__device__ __forceinline__ float key_to_float(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// Block maximum and minimum with one pair of barriers.
__device__ __forceinline__ void block_maxmin(float& mx, float& mn, float* scratch) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  mx = warp_max(mx);
  mn = -warp_max(-mn);
  consumer_sync();
  if (lane == 0) { scratch[w] = mx; scratch[kNW + w] = mn; }
  consumer_sync();
  mx = warp_max((lane < kNW) ? scratch[lane] : -INFINITY);
  mn = -warp_max((lane < kNW) ? -scratch[kNW + lane] : -INFINITY);
}

// Block reduction of K floats per thread with ONE bar.sync: per-warp shuffle trees, lane 0 parks the K warp sums in
// `buf` ([kNW][K] floats; the caller alternates two buffers between consecutive calls so no second barrier is needed
// against the next call's writes), then every thread adds the 8 warp sums in the fixed tree
// ((w0+w4)+(w2+w6)) + ((w1+w5)+(w3+w7)).  Deterministic, and identical on every thread.
template <int K>
__device__ __forceinline__ void block_sum_k(float (&v)[K], float* buf) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < K; ++k) v[k] = warp_sum(v[k]);
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < K; ++k) buf[w * K + k] = v[k];
  }
  consumer_sync();
#pragma unroll
  for (int k = 0; k < K; ++k)
    v[k] = ((buf[0 * K + k] + buf[4 * K + k]) + (buf[2 * K + k] + buf[6 * K + k])) +
           ((buf[1 * K + k] + buf[5 * K + k]) + (buf[3 * K + k] + buf[7 * K + k]));
}
static_assert(kNW == 8, "block_sum_k hard-codes the 8-warp tree");

// One sequence.  All consumer threads call.  `sc` / `ev` (shared, V floats each) hold the processed scores and their
// exponentials; thread t owns tokens t, t+256, ... in the selection passes (coalesced, conflict-free) and tokens
// 33t .. 33t+32 in the vocabulary-ordered multinomial.  The passes are deliberately compact loops (unrolled x11, not
// x33): this code runs once per step, so its instruction-cache footprint is paid as cold misses every time.
constexpr int kSmpUnroll = 11;
static_assert(kVPerThread % kSmpUnroll == 0, "sampler passes assume 3 x 11 elements per thread");

static __device__ void sample_row(const Params& P, const Smem& sm, int seq, int step_col) {
  Ctrl* c = sm.ctrl;
  float* sc = reinterpret_cast<float*>(sm.a_rm);
  float* ev = sm.red;
  float* scratch = c->scratch;
  float* rbuf = ev + 8256;                           // 2 x [kNW][6] words above the V exps (V <= 8256, tts_ar_create)
  unsigned long long* slist = reinterpret_cast<unsigned long long*>(ev + 8352);   // 32 gathered (key, weight) candidates
  int* scnt = reinterpret_cast<int*>(ev + 8416);
  if (threadIdx.x == 0) *scnt = 0;                   // ordered before the first gather by the reductions' barriers
  const int tid = threadIdx.x;
  const int V = P.V;
  const int stop = P.stop_token;

  const int gen = ld_cg_i(P.gen_count + seq);       // tokens generated before this one
  const bool fin = ld_cg_i(P.finished + seq) != 0;
  if (fin) {   // finished rows keep emitting the pad (== stop) token and are not fed to the model (Q7)
    if (tid == 0) {
      P.out_codes[(size_t)seq * P.out_stride + step_col] = stop;
      if (P.out_choice) P.out_choice[(size_t)seq * P.out_stride + step_col] = stop;
    }
    return;
  }
  const bool from_pf = P.first_step_from_prefill && step_col == 0;
  const int lrow = from_pf ? P.seq_utt[seq] : seq;
  const float* lg = (from_pf ? P.logits_pf : P.logits) + (size_t)lrow * P.v_stride;
  const uint32_t* seen = P.seen + (size_t)seq * P.seen_words;
  const float pen = P.rep_penalty;
  const bool use_temp = P.do_sample && P.temperature != 1.0f;
  const bool typ = P.typical_mass > 0.f;            // reference TypicalLogitsWarper (unified_voice.py:47-75, :659)

  // scores = processors(logits): repetition penalty, then (sampling only) temperature; kept as keys in sk[]
  // The row is copied L2 -> shared memory asynchronously (one round trip for the whole row), then transformed in
  // place: thread t owns tokens t, t+256, ... from here on.
  uint32_t* sk = reinterpret_cast<uint32_t*>(sc);
#pragma unroll 1
  for (int v = 4 * tid; v + 3 < P.v_stride; v += 4 * kConsumerThreads) cp_async_cg16(sc + v, lg + v);   // v_stride <= 8256
  cp_async_wait_all();
  consumer_sync();
  float lmax = -INFINITY, lmin = INFINITY;
#pragma unroll 3
  for (int v = tid; v < V; v += kConsumerThreads) {
    float s = sc[v];
    if (pen != 1.0f && ((__ldcg(seen + (v >> 5)) >> (v & 31)) & 1u)) s = (s < 0.f) ? s * pen : s / pen;
    if (use_temp && !typ) s = s / P.temperature;   // typical runs on the un-tempered scores: applied after it
    sk[v] = key_of(s);
    lmax = fmaxf(lmax, s);
    lmin = fminf(lmin, s);
  }
  block_maxmin(lmax, lmin, scratch);               // also orders the sk[] writes before the reads below
  float smax = lmax, smin = lmin;
  uint32_t kmax = key_of(smax);
  prof_mark(P, c, 31);

  int token;
  uint32_t kmin = key_of(-INFINITY);               // keep key >= kmin
  if (P.do_sample || typ) {
    int par = 0;
    float Z = 0.f;
    int n_keep = V;                                 // #{key >= kmin}
    // pass -1 = typical (only when asked for), pass 0 = top-k, pass 1 = softmax numerators + top-p.  One copy of the
    // search code (the sampler runs once per step, so every instruction of it is an instruction-cache miss: loops are
    // kept rolled on purpose).
#pragma unroll 1
    for (int pass = typ ? -1 : 0; pass < 2; ++pass) {
      const bool ones = pass == 0;
      const bool tmode = pass < 0;
      float lim;
      float cT = 0.f, pmn = smin, pmx = smax;       // probe range of the keys being searched
      if (tmode) {
        // TypicalLogitsWarper: shifted_i = |-log p_i - H| = |(smax + S1/Z) - s_i| with e_i = exp(s_i - smax),
        // Z = sum e_i, S1 = sum e_i (s_i - smax) (H = log Z - S1/Z).  Token i is kept iff the probability mass of the
        // tokens with a strictly smaller shifted value is < mass: the top-p rule over the key -shifted.
        float zl[2] = {0.f, 0.f};
#pragma unroll 3
        for (int v = tid; v < V; v += kConsumerThreads) {
          const float d = key_to_float(sk[v]) - smax;
          const float ex = expf(d);
          ev[v] = ex;
          zl[0] += ex;
          zl[1] += (ex > 0.f) ? ex * d : 0.f;       // nansum: 0 * -inf is skipped
        }
        block_sum_k<2>(zl, rbuf + par * 48);
        par ^= 1;
        cT = smax + zl[1] / zl[0];
        lim = P.typical_mass * zl[0];
        pmx = 0.f;
        pmn = -fmaxf(fabsf(cT - smax), fabsf(cT - smin));
      } else if (ones) {
        if (!(P.top_k > 0 && P.top_k < V)) continue;
        lim = (float)P.top_k;                       // counts <= 8448 are exact in fp32
      } else {
        // e = exp(score - max) over the top-k survivors
        float zl[2] = {0.f, 0.f};
#pragma unroll 3
        for (int v = tid; v < V; v += kConsumerThreads) {
          const uint32_t k = sk[v];
          const bool keep = k >= kmin;
          const float ex = keep ? expf(key_to_float(k) - smax) : 0.f;
          ev[v] = ex;
          zl[0] += ex;
          zl[1] += keep ? 1.f : 0.f;
        }
        block_sum_k<2>(zl, rbuf + par * 48);
        par ^= 1;
        Z = zl[0];
        n_keep = (int)zl[1];
        prof_mark(P, c, 32);
        if (!(P.top_p < 1.0f)) break;
        lim = P.top_p * Z;
      }
      // smallest T in [lo, hi] with weight{key > T} < lim.  pred(hi) holds: nothing lies above the maximum.
      uint32_t lo = kmin, hi = tmode ? key_of(0.f) : kmax;
      int n_lo = n_keep, n_hi = 0;                  // #{key >= lo}, #{key > hi}
      float m_hi = 0.f;                             // weight{key > hi}
      bool first = true;
      while (lo < hi) {
        if (n_lo - n_hi <= 32) {
          // gather the candidates (order arbitrary: sorted next), every warp finishes redundantly
#pragma unroll 3
          for (int v = tid; v < V; v += kConsumerThreads) {
            uint32_t k = sk[v];
            if (tmode) k = key_of(-fabsf(cT - key_to_float(k)));
            if (k >= lo && k <= hi) {
              const int p = atomicAdd(scnt, 1);
              if (p < 32) slist[p] = ((unsigned long long)k << 32) | __float_as_uint(ones ? 1.f : ev[v]);
            }
          }
          consumer_sync();
          const int n = min(*scnt, 32);
          const int lane = tid & 31;
          const unsigned long long it = warp_sort_desc(lane < n ? slist[lane] : 0ull);
          // walking down from hi, the first candidate whose inclusion lifts the weight to >= lim is the edge
          float incl = __uint_as_float((uint32_t)it);
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const float tt = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += tt;
          }
          const unsigned brk = __ballot_sync(0xffffffffu, lane < n && m_hi + incl >= lim);
          if (brk) lo = (uint32_t)(__shfl_sync(0xffffffffu, it, __ffs(brk) - 1) >> 32);
          consumer_sync();
          if (tid == 0) *scnt = 0;
          break;
        }
        // probes lo <= t0 <= t1 <= t2 < hi: quarters of the score interval (first round: also just below the
        // maximum, where the edge almost always is); quarters of the key interval when those degenerate
        uint32_t t[3];
        {
          const float fl = fmaxf(key_to_float(lo), pmn), fh = key_to_float(hi);
          bool ok = true;
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            float pj = fl + (fh - fl) * (0.25f * (j + 1));
            if (first) pj = fmaxf(pj, pmx - (j == 0 ? 16.f : j == 1 ? 4.f : 1.f));
            t[j] = key_of(pj);
            ok = ok && (fabsf(pj) <= 3.0e38f);
          }
          ok = ok && lo < t[0] && t[0] < t[1] && t[1] < t[2] && t[2] < hi;
          if (!ok) {
            const uint32_t q = (hi - lo) >> 2;
#pragma unroll
            for (int j = 0; j < 3; ++j) t[j] = q ? lo + q * (j + 1) - 1u : min(lo + j, hi);
          }
          first = false;
        }
        float ms[3] = {0.f, 0.f, 0.f};              // weight{key > t[j]}
        int cn[3] = {0, 0, 0};                      // #{key > t[j]}
#pragma unroll 3
        for (int v = tid; v < V; v += kConsumerThreads) {
          uint32_t k = sk[v];
          if (tmode) k = key_of(-fabsf(cT - key_to_float(k)));
          const float wgt = ones ? 1.f : ev[v];
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            const bool gt = k > t[j];
            ms[j] += gt ? wgt : 0.f;
            cn[j] += gt ? 1 : 0;
          }
        }
        {   // block sums with one barrier: float tree as block_sum_k, integer counts through redux
          const int lane = tid & 31, w = tid >> 5;
          float* buf = rbuf + par * 48;
          par ^= 1;
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            ms[j] = warp_sum(ms[j]);
            cn[j] = __reduce_add_sync(0xffffffffu, cn[j]);
          }
          if (lane == 0) {
#pragma unroll
            for (int j = 0; j < 3; ++j) {
              buf[w * 6 + j] = ms[j];
              buf[w * 6 + 3 + j] = __int_as_float(cn[j]);
            }
          }
          consumer_sync();
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            ms[j] = ((buf[0 * 6 + j] + buf[4 * 6 + j]) + (buf[2 * 6 + j] + buf[6 * 6 + j])) +
                    ((buf[1 * 6 + j] + buf[5 * 6 + j]) + (buf[3 * 6 + j] + buf[7 * 6 + j]));
            int tot = 0;
#pragma unroll
            for (int ww = 0; ww < kNW; ++ww) tot += __float_as_int(buf[ww * 6 + 3 + j]);
            cn[j] = tot;
          }
        }
        // pred(T) = weight{key > T} < lim is false..false true..true in T; answer = smallest T with pred
        if (ms[0] < lim) { hi = t[0]; m_hi = ms[0]; n_hi = cn[0]; }
        else if (ms[1] < lim) { lo = t[0] + 1u; n_lo = cn[0]; hi = t[1]; m_hi = ms[1]; n_hi = cn[1]; }
        else if (ms[2] < lim) { lo = t[1] + 1u; n_lo = cn[1]; hi = t[2]; m_hi = ms[2]; n_hi = cn[2]; }
        else { lo = t[2] + 1u; n_lo = cn[2]; }
      }
      if (tmode) {
        // apply the typical filter (removed tokens get the key of -inf), then the deferred temperature; the maximum,
        // minimum and count of the survivors restart the top-k / top-p searches (and the greedy argmax)
        float mx = -INFINITY, mn = INFINITY;
        float zl[2] = {0.f, 0.f};
#pragma unroll 3
        for (int v = tid; v < V; v += kConsumerThreads) {
          float sv = key_to_float(sk[v]);
          if (key_of(-fabsf(cT - sv)) >= lo) {
            if (use_temp) sv = sv / P.temperature;
            sk[v] = key_of(sv);
            mx = fmaxf(mx, sv);
            mn = fminf(mn, sv);
            zl[1] += 1.f;
          } else {
            sk[v] = key_of(-INFINITY);
          }
        }
        block_sum_k<2>(zl, rbuf + par * 48);
        par ^= 1;
        block_maxmin(mx, mn, scratch);
        smax = mx; smin = mn;
        kmax = key_of(smax);
        kmin = key_of(smin);
        n_keep = (int)zl[1];
        if (!P.do_sample) break;
        continue;
      }
      kmin = lo;
      n_keep = n_lo;                                // upper bound is all the next pass needs
    }
  }
  if (!P.do_sample) {
    // HF greedy: argmax of the processed scores; first index on ties (torch.argmax)
    int best = 0x7fffffff;
#pragma unroll kSmpUnroll
    for (int v = tid; v < V; v += kConsumerThreads)
      if (sk[v] == kmax) best = min(best, v);
    best = __reduce_min_sync(0xffffffffu, best);
    consumer_sync();
    if ((tid & 31) == 0) reinterpret_cast<int*>(scratch)[tid >> 5] = best;
    consumer_sync();
    int b2 = ((tid & 31) < kNW) ? reinterpret_cast<int*>(scratch)[tid & 31] : 0x7fffffff;
    token = __reduce_min_sync(0xffffffffu, b2);
  } else {
    consumer_sync();                               // ev[] complete before the vocabulary-ordered reads
    prof_mark(P, c, 33);
    // ---- multinomial: inverse CDF in vocabulary order (thread t owns tokens 33t .. 33t+32) ----
    const int v0 = tid * kVPerThread;
    const int vend = min(v0 + kVPerThread, V);
    float loc = 0.f;
#pragma unroll kSmpUnroll
    for (int v = v0; v < vend; ++v) {
      const float e = (sk[v] >= kmin) ? ev[v] : 0.f;
      ev[v] = e;                                   // own range only: kept mass, re-read below
      loc += e;
    }
    // exclusive prefix over threads (fixed tree)
    const int lane = tid & 31, w = tid >> 5;
    float incl = loc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    consumer_sync();
    if (lane == 31) scratch[w] = incl;
    consumer_sync();
    float wbase = 0.f, total = 0.f;
#pragma unroll
    for (int i = 0; i < kNW; ++i) {
      const float t = scratch[i];
      if (i < w) wbase += t;
      total += t;
    }
    const long long gseq = P.seq_indices ? P.seq_indices[seq] : P.seq_index_base + seq;
    const float u01 = philox_uniform(P.seed, (uint32_t)gseq, (uint32_t)gen);
    const float target = u01 * total;
    float run = wbase + incl - loc;
    int found = 0x7fffffff, lastkept = -1;
#pragma unroll kSmpUnroll
    for (int v = v0; v < vend; ++v) {
      const float e = ev[v];
      run += e;
      if (e > 0.f) {
        lastkept = v;
        if (run > target && found == 0x7fffffff) found = v;
      }
    }
    found = __reduce_min_sync(0xffffffffu, found);
    lastkept = __reduce_max_sync(0xffffffffu, lastkept);
    consumer_sync();
    if (lane == 0) {
      reinterpret_cast<int*>(scratch)[w] = found;
      reinterpret_cast<int*>(scratch)[kNW + w] = lastkept;
    }
    consumer_sync();
    int f2 = (lane < kNW) ? reinterpret_cast<int*>(scratch)[lane] : 0x7fffffff;
    int l2 = (lane < kNW) ? reinterpret_cast<int*>(scratch)[kNW + lane] : -1;
    f2 = __reduce_min_sync(0xffffffffu, f2);
    l2 = __reduce_max_sync(0xffffffffu, l2);
    token = (f2 != 0x7fffffff) ? f2 : l2;
  }

  prof_mark(P, c, 34);
  // ---- bookkeeping (sample_stream :1160-1190) + next-step embedding (unified_voice.py:212-214) ----
  const size_t oidx = (size_t)seq * P.out_stride + step_col;
  int tok = token;
  if (P.forced) tok = P.forced[oidx];
  if (tid == 0) {
    if (P.out_choice) P.out_choice[oidx] = token;
    P.out_codes[oidx] = tok;
    atomicOr(P.seen + (size_t)seq * P.seen_words + (tok >> 5), 1u << (tok & 31));
    P.gen_count[seq] = gen + 1;
    if (tok == stop && !P.ignore_eos) {
      P.finished[seq] = 1;
      P.row_pos[seq] = -1;
      atomicSub(P.n_active, 1);
    } else {
      P.row_pos[seq] = P.utt_plen[P.seq_utt[seq]] + gen;     // position of this token in its sequence
    }
  }
  if (P.out_latents) {   // final_norm(hidden) of the forward pass that produced these logits (:1172)
    const float* src = (from_pf ? P.cur_latent_pf : P.cur_latent) + (size_t)lrow * kD;
    float* dst = P.out_latents + oidx * kD;
    for (int i = tid; i < kD / 4; i += kConsumerThreads)
      reinterpret_cast<float4*>(dst)[i] = ld_cg_f4(src + 4 * i);
  }
  // k-th generated token (k = gen + 1) is embedded with mel_pos[k + 1] (quirk Q2: index 1 never used)
  int pidx = gen + 2;
  if (pidx >= P.max_pos_rows) pidx = P.max_pos_rows - 1;
  const float4* e4 = reinterpret_cast<const float4*>(P.mel_emb + (size_t)tok * kD);
  const float4* p4 = reinterpret_cast<const float4*>(P.mel_pos + (size_t)pidx * kD);
  float4* x4 = reinterpret_cast<float4*>(P.x + (size_t)seq * kD);
  // head-cluster kernels (P.y != nullptr): the attention block of layer 0 accumulates onto y = x + attn.c_proj bias
  const float4* b4 = P.y ? reinterpret_cast<const float4*>(layer_f(P, 0) + kFOffBproj) : nullptr;
  for (int i = tid; i < kD / 4; i += kConsumerThreads) {
    const float4 a = e4[i], b = p4[i];
    const float4 xv = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
    x4[i] = xv;
    if (b4) {
      const float4 bb = b4[i];
      longlong2* y2 = reinterpret_cast<longlong2*>(P.y + (size_t)seq * kD + 4 * i);
      y2[0] = make_longlong2(cnt_word(xv.x + bb.x, 1), cnt_word(xv.y + bb.y, 1));   // counting accumulator: 1 contribution so far
      y2[1] = make_longlong2(cnt_word(xv.z + bb.z, 1), cnt_word(xv.w + bb.w, 1));
    }
  }
  consumer_sync();   // sc / ev / scratch are reused by the next row
  prof_mark(P, c, 35);
}

static __device__ void sample_stage(const Params& P, const Instr& ins, const Smem& sm, int step_col) {
  for (int seq = blockIdx.x; seq < P.n_seqs; seq += P.G) {
    sample_row(P, sm, seq, step_col);
    if (P.stream_flags) {        // streaming: publish "this sequence's (token, latent) of step step_col is written" to the host
      consumer_sync();           // every thread's latent stores precede the fence below (CTA barrier + cumulativity)
      if (threadIdx.x == 0) {
        __threadfence_system();
        atomicAdd_system(P.stream_flags + step_col, 1);
      }
    }
  }
}

}  // namespace tts
