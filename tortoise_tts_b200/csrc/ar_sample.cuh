// ar_sample.cuh -- logits processors, warpers, multinomial and the sample-loop bookkeeping, one CTA per
// sequence.  Restates, without a sort:
//   HF RepetitionPenaltyLogitsProcessor (over ALL input ids incl. the dummy prompt ids 1 and 8192 -- quirk Q4,
//   reference unified_voice.py:647-649), TemperatureLogitsWarper, TopKLogitsWarper, TopPLogitsWarper in the
//   order of the reference's _get_logits_warper (stream_generator.py:56-101), softmax + multinomial and the
//   pad-after-stop / unfinished bookkeeping of sample_stream (stream_generator.py:1160-1190).
// Top-k / top-p thresholds come from a bisection over the monotone uint image of the fp32 scores:
//   top-k keeps score >= (k-th largest);   top-p keeps token i iff mass{score > score_i} < p
// which is HF's "ascending cumsum <= 1-p -> remove" rule (tests/test_oracle_golden.py proves the identity).
// The multinomial is an inverse CDF in vocabulary order driven by Philox4x32-10 (oracle/philox.py).
#pragma once
#include "ar_core.cuh"

namespace tts {

constexpr int kVPerThread = (8448 + kConsumerThreads - 1) / kConsumerThreads;   // covers n_mel_codes <= 8448

__device__ __forceinline__ uint32_t key_of(float v) { return float_key(__float_as_uint(v)); }

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
  float* rbuf = ev + (8448 - 64);                    // 2 x [kNW][4] floats above the V exps (V <= 8256, tts_ar_create)
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
  const int lrow = (P.first_step_from_prefill && step_col == 0) ? P.seq_utt[seq] : seq;
  const float* lg = P.logits + (size_t)lrow * P.v_stride;
  const uint32_t* seen = P.seen + (size_t)seq * P.seen_words;
  const float pen = P.rep_penalty;
  const bool use_temp = P.do_sample && P.temperature != 1.0f;

  // scores = processors(logits): repetition penalty, then (sampling only) temperature
  float lmax = -INFINITY;
#pragma unroll 1
  for (int i0 = 0; i0 < kVPerThread; i0 += kSmpUnroll) {
    float s[kSmpUnroll];
#pragma unroll
    for (int i = 0; i < kSmpUnroll; ++i) {
      const int v = tid + (i0 + i) * kConsumerThreads;
      s[i] = (v < V) ? ld_cg_f(lg + v) : -INFINITY;
    }
#pragma unroll
    for (int i = 0; i < kSmpUnroll; ++i) {
      const int v = tid + (i0 + i) * kConsumerThreads;
      if (v < V) {
        if (pen != 1.0f && ((__ldcg(seen + (v >> 5)) >> (v & 31)) & 1u)) s[i] = (s[i] < 0.f) ? s[i] * pen : s[i] / pen;
        if (use_temp) s[i] = s[i] / P.temperature;
        sc[v] = s[i];
        lmax = fmaxf(lmax, s[i]);
      }
    }
  }
  const float smax = block_max(lmax, scratch);   // also orders the sc[] writes before the reads below
  prof_mark(P, c, 31);

  int token;
  if (!P.do_sample) {
    // HF greedy: argmax of the processed scores; first index on ties (torch.argmax)
    int best = 0x7fffffff;
#pragma unroll kSmpUnroll
    for (int v = tid; v < V; v += kConsumerThreads)
      if (sc[v] == smax) best = min(best, v);
    best = __reduce_min_sync(0xffffffffu, best);
    consumer_sync();
    if ((tid & 31) == 0) reinterpret_cast<int*>(scratch)[tid >> 5] = best;
    consumer_sync();
    int b2 = ((tid & 31) < kNW) ? reinterpret_cast<int*>(scratch)[tid & 31] : 0x7fffffff;
    token = __reduce_min_sync(0xffffffffu, b2);
  } else {
    uint32_t kmin = key_of(-INFINITY);             // keep key >= kmin
    int par = 0;
    // ---- top-k: largest T with #{key >= T} >= k  (4-way search: 3 probes per round) ----
    if (P.top_k > 0 && P.top_k < V) {
      uint32_t lo = key_of(-INFINITY), hi = key_of(smax);   // count(lo) = V >= k; answer in [lo, hi]
      const float kf = (float)P.top_k;                      // counts <= 8448 are exact in fp32
      while (lo < hi) {
        const uint32_t q = (hi - lo) >> 2;
        uint32_t t[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) t[j] = q ? lo + q * (j + 1) : min(lo + 1u + j, hi);   // lo < t0 <= t1 <= t2 <= hi
        float cnt[3] = {0.f, 0.f, 0.f};
#pragma unroll kSmpUnroll
        for (int v = tid; v < V; v += kConsumerThreads) {
          const uint32_t k = key_of(sc[v]);
#pragma unroll
          for (int j = 0; j < 3; ++j) cnt[j] += (k >= t[j]) ? 1.f : 0.f;
        }
        block_sum_k<3>(cnt, rbuf + par * 32);
        par ^= 1;
        // pred(T) = count(T) >= k is true..true false..false in T; answer = largest T with pred
        if (cnt[2] >= kf) lo = t[2];
        else if (cnt[1] >= kf) { lo = t[1]; hi = t[2] - 1u; }
        else if (cnt[0] >= kf) { lo = t[0]; hi = t[1] - 1u; }
        else hi = t[0] - 1u;
      }
      kmin = lo;
    }
    // ---- e = exp(score - max) over the top-k survivors (compact loop: once-per-step code, see above) ----
    float zl[1] = {0.f};
#pragma unroll 3
    for (int v = tid; v < V; v += kConsumerThreads) {
      const float sv = sc[v];
      const float ex = (key_of(sv) >= kmin) ? expf(sv - smax) : 0.f;
      ev[v] = ex;
      zl[0] += ex;
    }
    block_sum_k<1>(zl, rbuf + par * 32);
    par ^= 1;
    const float Z = zl[0];
    prof_mark(P, c, 32);
    // ---- top-p: smallest T with mass{key > T} < p * Z ; keep key >= T  (4-way search over registers) ----
    if (P.top_p < 1.0f) {
      // (key, e) of the thread's 33 tokens in registers: the search rounds below touch no memory
      uint32_t key[kVPerThread];
      float e[kVPerThread];
#pragma unroll
      for (int i = 0; i < kVPerThread; ++i) {
        const int v = tid + i * kConsumerThreads;
        key[i] = (v < V) ? key_of(sc[v]) : 0u;                    // 0 < key(-inf): padding never passes a probe
        e[i] = (v < V) ? ev[v] : 0.f;
      }
      const float lim = P.top_p * Z;
      uint32_t lo = kmin, hi = key_of(smax);        // pred(hi): mass above the max is 0 < lim
      bool first = true;
      while (lo < hi) {
        const uint32_t q = (hi - lo) >> 2;
        uint32_t t[3];
        if (first) {   // the nucleus edge is almost always within a few units of the maximum score: probe there first
          t[0] = min(max(key_of(smax - 16.f), lo), hi);
          t[1] = min(max(key_of(smax - 4.f), t[0]), hi);
          t[2] = min(max(key_of(smax - 1.f), t[1]), hi);
          first = false;
        } else {
#pragma unroll
          for (int j = 0; j < 3; ++j) t[j] = q ? lo + q * (j + 1) - 1u : min(lo + j, hi);   // lo <= t0 <= t1 <= t2 <= hi
        }
        float ms[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int i = 0; i < kVPerThread; ++i) {
#pragma unroll
          for (int j = 0; j < 3; ++j) ms[j] += (key[i] > t[j]) ? e[i] : 0.f;
        }
        block_sum_k<3>(ms, rbuf + par * 32);
        par ^= 1;
        // pred(T) = mass{key > T} < lim is false..false true..true in T; answer = smallest T with pred
        if (ms[0] < lim) hi = t[0];
        else if (ms[1] < lim) { lo = t[0] + 1u; hi = t[1]; }
        else if (ms[2] < lim) { lo = t[1] + 1u; hi = t[2]; }
        else lo = t[2] + 1u;
      }
      kmin = lo;
    }
    consumer_sync();                               // ev[] complete before the vocabulary-ordered reads
    prof_mark(P, c, 33);
    // ---- multinomial: inverse CDF in vocabulary order (thread t owns tokens 33t .. 33t+32) ----
    const int v0 = tid * kVPerThread;
    const int vend = min(v0 + kVPerThread, V);
    float loc = 0.f;
#pragma unroll kSmpUnroll
    for (int v = v0; v < vend; ++v) {
      const float e = (key_of(sc[v]) >= kmin) ? ev[v] : 0.f;
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
    const float u01 = philox_uniform(P.seed, (uint32_t)(P.seq_index_base + seq), (uint32_t)gen);
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
    const float* src = P.cur_latent + (size_t)lrow * kD;
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
  for (int i = tid; i < kD / 4; i += kConsumerThreads) {
    const float4 a = e4[i], b = p4[i];
    x4[i] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
  }
  consumer_sync();   // sc / ev / scratch are reused by the next row
  prof_mark(P, c, 35);
}

static __device__ void sample_stage(const Params& P, const Instr& ins, const Smem& sm, int step_col) {
  for (int seq = blockIdx.x; seq < P.n_seqs; seq += P.G) sample_row(P, sm, seq, step_col);
}

}  // namespace tts
