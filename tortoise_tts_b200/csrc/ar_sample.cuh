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

// Smallest key T in [lo, hi] such that pred(T) holds; pred monotone (false..false true..true), pred(hi) true.
// count mode: pred(T) = #{key >= T} ... (see callers).  Implemented inline at the call sites.

// One sequence.  `sc` = V floats (scores), `ev` = V floats scratch, both shared.  All 512 consumer threads call.
static __device__ void sample_row(const Params& P, const Smem& sm, int seq, int step_col) {
  Ctrl* c = sm.ctrl;
  float* sc = reinterpret_cast<float*>(sm.a_rm);
  float* ev = sm.red;
  float* scratch = c->scratch;
  const int tid = threadIdx.x;
  const int V = P.V;
  const int stop = P.stop_token;

  const bool fin = ld_cg_i(P.finished + seq) != 0;
  if (fin) {   // finished rows keep emitting the pad (== stop) token and are not fed to the model (Q7)
    if (tid == 0) {
      P.out_codes[(size_t)seq * P.out_stride + step_col] = stop;
      if (P.out_choice) P.out_choice[(size_t)seq * P.out_stride + step_col] = stop;
    }
    return;
  }
  const int gen = ld_cg_i(P.gen_count + seq);       // tokens generated before this one
  const int lrow = (P.first_step_from_prefill && step_col == 0) ? P.seq_utt[seq] : seq;
  const float* lg = P.logits + (size_t)lrow * P.v_stride;
  const uint32_t* seen = P.seen + (size_t)seq * P.seen_words;
  const float pen = P.rep_penalty;

  // scores = processors(logits): repetition penalty, then (sampling only) temperature
  float lmax = -INFINITY;
  for (int v = tid; v < V; v += kConsumerThreads) {
    float s = ld_cg_f(lg + v);
    if (pen != 1.0f && ((__ldcg(seen + (v >> 5)) >> (v & 31)) & 1u)) s = (s < 0.f) ? s * pen : s / pen;
    if (P.do_sample && P.temperature != 1.0f) s = s / P.temperature;
    sc[v] = s;
    lmax = fmaxf(lmax, s);
  }
  const float smax = block_max(lmax, scratch);   // also orders the sc[] writes before the reads below

  int token;
  if (!P.do_sample) {
    // HF greedy: argmax of the processed scores; first index on ties (torch.argmax)
    int best = 0x7fffffff;
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
    // ---- top-k: largest T with #{key >= T} >= k  ----
    if (P.top_k > 0 && P.top_k < V) {
      uint32_t lo = key_of(-INFINITY), hi = key_of(smax);   // count(lo) = V >= k; answer in [lo, hi]
      while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo + 1u) >> 1);
        uint32_t cnt = 0;
        for (int v = tid; v < V; v += kConsumerThreads) cnt += (key_of(sc[v]) >= mid) ? 1u : 0u;
        cnt = block_sum_u32(cnt, scratch);
        if (cnt >= (uint32_t)P.top_k) lo = mid; else hi = mid - 1u;
      }
      kmin = lo;
    }
    // ---- e = exp(score - max) over the top-k survivors ----
    float zloc = 0.f;
    for (int v = tid; v < V; v += kConsumerThreads) {
      const float e = (key_of(sc[v]) >= kmin) ? expf(sc[v] - smax) : 0.f;
      ev[v] = e;
      zloc += e;
    }
    const float Z = block_sum(zloc, scratch);
    // ---- top-p: smallest T with mass{key > T} < p * Z ; keep key >= T ----
    if (P.top_p < 1.0f) {
      const float lim = P.top_p * Z;
      uint32_t lo = kmin, hi = key_of(smax);        // pred(hi): mass above the max is 0 < lim
      while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        float ms = 0.f;
        for (int v = tid; v < V; v += kConsumerThreads) ms += (key_of(sc[v]) > mid) ? ev[v] : 0.f;
        ms = block_sum(ms, scratch);
        if (ms < lim) hi = mid; else lo = mid + 1u;
      }
      kmin = lo;
    }
    // ---- multinomial: inverse CDF in vocabulary order ----
    const int v0 = tid * kVPerThread;
    float loc = 0.f;
#pragma unroll
    for (int i = 0; i < kVPerThread; ++i) {
      const int v = v0 + i;
      if (v < V) {
        const float e = (key_of(sc[v]) >= kmin) ? ev[v] : 0.f;
        loc += e;
      }
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
#pragma unroll
    for (int i = 0; i < kVPerThread; ++i) {
      const int v = v0 + i;
      if (v < V) {
        const float e = (key_of(sc[v]) >= kmin) ? ev[v] : 0.f;
        run += e;
        if (e > 0.f) {
          lastkept = v;
          if (run > target && found == 0x7fffffff) found = v;
        }
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
}

static __device__ void sample_stage(const Params& P, const Instr& ins, const Smem& sm, int step_col) {
  for (int seq = blockIdx.x; seq < P.n_seqs; seq += P.G) sample_row(P, sm, seq, step_col);
}

}  // namespace tts
