// tts_ar_run.cu -- C ABI call path: prompt pass, sample loop, latent re-forward, host-buffer entry.
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <algorithm>
#include <chrono>
#include "tts_ar_host.h"

using namespace tts;

namespace tts {

void fill_common_params(TtsArHandle* h, Params& P) {
  memset(&P, 0, sizeof(P));
  const TtsArConfig& c = h->cfg;
  P.G = h->G;
  P.n_layers = c.n_layers;
  P.V = c.n_mel_codes;
  P.v_tiles = h->v_tiles;
  P.v_stride = h->v_stride;
  P.layers = h->d_layers;
  P.warena = h->d_warena; P.fparams = h->d_fparams; P.w_layer_u4 = h->w_layer_u4; P.fp_layer = h->fp_layer;
  P.w_head = h->d_w_head;
  P.b_head = h->d_b_head;
  P.lnf_g = h->d_lnf_g; P.lnf_b = h->d_lnf_b; P.fn_g = h->d_fn_g; P.fn_b = h->d_fn_b;
  P.mel_emb = h->d_mel_emb; P.mel_pos = h->d_mel_pos;
  P.x = h->d_x; P.q = h->d_q; P.ctx = h->d_ctx; P.hbuf = h->d_h;
  P.xn = h->d_xn;
  P.attn_tiles_mode = h->opt_attn_tiles;
  P.logits = h->d_logits; P.cur_latent = h->d_cur_latent;
  P.logits_pf = h->d_logits; P.cur_latent_pf = h->d_cur_latent;
  P.kv = h->d_kv; P.n_pages = h->n_pages;
  P.prefix_pt = h->d_prefix_pt; P.priv_pt = h->d_priv_pt; P.utt_plen = h->d_utt_plen;
  P.head_rows = h->d_head_rows;
  P.part = h->d_part; P.attn_cnt = h->d_attn_cnt; P.plan = h->d_plan; P.plan_cap_units = h->plan_cap_units;
  P.finished = h->d_finished; P.gen_count = h->d_gen_count; P.seen = h->d_seen; P.seen_words = h->seen_words;
  P.n_active = h->d_n_active; P.seq_utt = h->d_seq_utt;
  P.stop_token = c.stop_mel_token;
  P.max_pos_rows = c.n_mel_pos;
  P.latent_dst = h->d_latent_dst;
  P.bar_counter = h->d_bar; P.watchdog = h->d_watchdog; P.steps_done = h->d_steps_done;
}

static int mt_max_for(int n_rows) {
  const int m = (std::min(n_rows, kRowGroup) + 15) / 16;
  return m <= 1 ? 1 : (m == 2 ? 2 : 4);
}

static int launch(TtsArHandle* h, Params& P, cudaStream_t st, bool cluster = false) {
  P.n_slots = smem_max_slots(P.mt_max, h->smem_budget);
  if (P.n_slots < 2) return fail(h, TTS_AR_ERR_DEVICE, "not enough shared memory for the weight ring");
  TTS_CUDA(h, cudaMemsetAsync(h->d_bar, 0, sizeof(unsigned int), st));
  cudaError_t e;
  if (cluster) {
    const int mode = decode_mode_for(P.n_rows);
    P.G = kDecCtas;
    P.y = h->d_y;
    P.mt_max = decode_mt_smem(mode);
    P.n_slots = smem_max_slots(P.mt_max, h->smem_budget);
    // y: counting accumulators, one contribution (the residual) deemed present in rows the sampler does not re-initialise
    // (finished sequences are still carried through the layers); xt: tags of a previous launch must not validate
    launch_fill_i64(h->d_y, (size_t)kDecRows * mode * kD, 1ll, st);
    TTS_CUDA(h, cudaMemsetAsync(h->d_xt, 0, sizeof(uint2) * kDecRows * mode * kD, st));
    P.xt = h->d_xt;
    h->launches += 1;
    e = launch_decode(P, mode, smem_total_bytes(P.mt_max, P.n_slots), h->cfg.device, decode_cooperative(h), st);
    if (e != cudaSuccess) return cuda_fail(h, e, "cudaLaunchKernelEx(ar_decode_kernel)");
    h->launches += 1;
    return 0;
  }
  e = launch_program(P, smem_total_bytes(P.mt_max, P.n_slots), h->cfg.device, st);
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaLaunchCooperativeKernel(ar_program_kernel)");
  h->launches += 1;
  return 0;
}

// Host -> device upload of small index arrays through the handle's pinned staging area (no pageable copies, so no
// synchronisation on the call path).  The area is recycled per call: before the first write of a call the previous
// call's copies must have been consumed (an event wait that is normally already satisfied).
struct Stager {
  TtsArHandle* h;
  cudaStream_t st;
  size_t used = 0;
  int rc = 0;
  Stager(TtsArHandle* h_, cudaStream_t st_) : h(h_), st(st_) {
    if (h->staging_busy) {
      cudaError_t e = cudaEventSynchronize(h->ev_staging);
      if (e != cudaSuccess) rc = cuda_fail(h, e, "cudaEventSynchronize(staging)");
      h->staging_busy = false;
    }
  }
  int put(void* dst, const void* src, size_t bytes) {
    if (rc) return rc;
    const size_t a = (bytes + 15) & ~(size_t)15;
    if (used + a > h->staging_bytes) return rc = fail(h, TTS_AR_ERR_CAPACITY, "host staging area too small");
    memcpy(h->h_staging + used, src, bytes);
    cudaError_t e = cudaMemcpyAsync(dst, h->h_staging + used, bytes, cudaMemcpyHostToDevice, st);
    used += a;
    if (e != cudaSuccess) rc = cuda_fail(h, e, "cudaMemcpyAsync(staging)");
    return rc;
  }
  int done() {
    if (rc) return rc;
    cudaError_t e = cudaEventRecord(h->ev_staging, st);
    if (e != cudaSuccess) return rc = cuda_fail(h, e, "cudaEventRecord(staging)");
    h->staging_busy = true;
    return 0;
  }
};

// Prompt-shaped causal pass: sequences laid out back to back as rows; every sequence is its own "utterance"
// whose KV lives in prefix pages.  Used by tts_ar_prefill (with_head) and tts_ar_forward_latents (latent_out).
int run_prompt_pass(TtsArHandle* h, Stager& sg, const std::vector<int>& seq_len, const std::vector<int>& row_utt,
                    const std::vector<int>& row_pos, const std::vector<int>& head_rows, const std::vector<int>& latent_dst,
                    bool with_head, cudaStream_t st) {
  const int n_seq = (int)seq_len.size();
  const int n_rows = (int)row_utt.size();
  if (n_rows > h->rows_cap) return fail(h, TTS_AR_ERR_CAPACITY, "prompt rows exceed capacity");
  // page tables: sequence i owns pages [base_i, base_i + ceil(len/32))
  std::vector<int> pt((size_t)n_seq * kMaxPagesPerSeg, 0);
  int next_page = 0;
  for (int i = 0; i < n_seq; ++i) {
    const int np = (seq_len[i] + kPage - 1) / kPage;
    if (np > kMaxPagesPerSeg) return fail(h, TTS_AR_ERR_CAPACITY, "sequence longer than 1024 positions");
    for (int p = 0; p < np; ++p) pt[(size_t)i * kMaxPagesPerSeg + p] = next_page++;
  }
  if (next_page > h->n_pages) return fail(h, TTS_AR_ERR_CAPACITY, "KV page pool exhausted");
  sg.put(h->d_prefix_pt, pt.data(), sizeof(int) * pt.size());
  sg.put(h->d_utt_plen, seq_len.data(), sizeof(int) * n_seq);
  sg.put(h->d_pf_row_utt, row_utt.data(), sizeof(int) * n_rows);
  sg.put(h->d_pf_row_pos, row_pos.data(), sizeof(int) * n_rows);
  if (with_head)
    sg.put(h->d_head_rows, head_rows.data(), sizeof(int) * head_rows.size());
  else
    sg.put(h->d_latent_dst, latent_dst.data(), sizeof(int) * n_rows);
  if (sg.rc) return sg.rc;
  TTS_CUDA(h, cudaMemsetAsync(h->d_plan, 0, sizeof(int) * 4, st));   // plan[0] = 0: direct (row, head) units
  return 0;
}

// Launch-attribute choice for the head-cluster kernel: cooperative (co-residency checked by the driver) unless the
// caller opted out or a profiler that cannot replay cooperative launches is injected into this process (Nsight Compute
// sets these variables in the target's environment); the 128 CTAs of a plain cluster launch are still co-resident on an
// otherwise idle B200 and every device-side wait is watchdogged.
bool decode_cooperative(TtsArHandle* h) {
  if (h->opt_no_coop) return false;
  static const bool profiled = getenv("NV_COMPUTE_PROFILER_PERFWORKS_DIR") != nullptr || getenv("CUDA_INJECTION64_PATH") != nullptr ||
                               getenv("NV_NSIGHT_INJECTION_TRANSPORT_TYPE") != nullptr;
  return !profiled;
}

}  // namespace tts

static int check_ready(TtsArHandle* h, bool need_prefill) {
  if (!h) return TTS_AR_ERR_BAD_ARG;
  if (!h->weights_loaded) return fail(h, TTS_AR_ERR_STATE, "tts_ar_load_weights has not been called");
  if (need_prefill && !h->prefilled) return fail(h, TTS_AR_ERR_STATE, "tts_ar_prefill has not been called");
  cudaError_t e = cudaSetDevice(h->cfg.device);
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaSetDevice");
  return 0;
}

extern "C" int tts_ar_prefill(TtsArHandle* h, const float* latent, const int32_t* text, int32_t text_stride,
                              const int32_t* text_len, int32_t n_utt, int32_t n_cand, void* stream) {
  int rc = check_ready(h, false);
  if (rc) return rc;
  if (!latent || !text || !text_len || n_utt < 1 || n_cand < 1) return fail(h, TTS_AR_ERR_BAD_ARG, "bad prefill argument");
  const TtsArConfig& c = h->cfg;
  if (n_utt > c.max_utts || n_utt * n_cand > c.max_seqs) return fail(h, TTS_AR_ERR_CAPACITY, "batch exceeds max_utts / max_seqs");
  cudaStream_t st = (cudaStream_t)stream;
  std::vector<int> seq_len(n_utt), row_utt, row_pos, head_rows(n_utt), none;
  for (int u = 0; u < n_utt; ++u) {
    if (text_len[u] < 0 || text_len[u] > c.max_text_len || text_len[u] > text_stride)
      return fail(h, TTS_AR_ERR_CAPACITY, "text length exceeds max_text_len / text_stride");
    seq_len[u] = text_len[u] + 4;          // latent + start_text + text + stop_text + start_mel
    for (int p = 0; p < seq_len[u]; ++p) { row_utt.push_back(u); row_pos.push_back(p); }
    head_rows[u] = (int)row_utt.size() - 1;
  }
  const int n_rows = (int)row_utt.size();
  h->prefilled = false;
  TTS_CUDA(h, cudaEventRecord(h->ev0, st));
  TTS_CUDA(h, cudaMemsetAsync(h->d_watchdog, 0, 2 * sizeof(int), st));   // watchdog + bad-token flags of the previous call
  Stager sg(h, st);
  if (sg.rc) return sg.rc;
  sg.put(h->d_text_len, text_len, sizeof(int) * n_utt);
  rc = run_prompt_pass(h, sg, seq_len, row_utt, row_pos, head_rows, none, true, st);
  if (rc) return rc;
  EmbedArgs ea{};
  ea.x = h->d_x; ea.row_utt = h->d_pf_row_utt; ea.row_pos_in = h->d_pf_row_pos; ea.text_len = h->d_text_len;
  ea.text = text; ea.text_stride = text_stride; ea.codes = nullptr; ea.codes_len = h->d_codes_len; ea.codes_stride = 0;
  ea.latent = latent; ea.text_emb = h->d_text_emb; ea.text_pos = h->d_text_pos; ea.mel_emb = h->d_mel_emb; ea.mel_pos = h->d_mel_pos;
  ea.start_text = c.start_text_token; ea.stop_text = c.stop_text_token; ea.start_mel = c.start_mel_token; ea.stop_mel = c.stop_mel_token;
  ea.n_text_rows = c.n_text_rows; ea.n_mel_codes = c.n_mel_codes; ea.err_flag = h->d_watchdog + 1;
  launch_embed_rows(ea, n_rows, st);
  h->launches += 1;
  Params P;
  fill_common_params(h, P);
  P.decode_mode = 0;
  P.row_utt = h->d_pf_row_utt; P.row_pos = h->d_pf_row_pos;
  P.mt_max = mt_max_for(std::max(n_rows, n_utt));
  P.has_sample = 0; P.n_rows = n_rows; P.tail = 1; P.n_head_rows = n_utt;
  P.n_steps = 1;
  P.n_seqs = 0;
  rc = h->opt_dense_interp ? launch(h, P, st) : run_dense_pass(h, P, st);
  if (rc) return rc;
  // sequences: utterance-major, candidates adjacent (repeat_interleave, stream_generator.py:618-623)
  h->n_utt = n_utt; h->n_cand = n_cand; h->n_seqs = n_utt * n_cand;
  h->plen = seq_len;
  std::vector<int> seq_utt(h->n_seqs), priv((size_t)h->n_seqs * kMaxPagesPerSeg, 0);
  int next_page = 0;
  for (int u = 0; u < n_utt; ++u) next_page += (seq_len[u] + kPage - 1) / kPage;
  for (int s = 0; s < h->n_seqs; ++s) {
    seq_utt[s] = s / n_cand;
    for (int p = 0; p < h->nqp; ++p) priv[(size_t)s * kMaxPagesPerSeg + p] = next_page++;
  }
  if (next_page > h->n_pages) return fail(h, TTS_AR_ERR_CAPACITY, "KV page pool exhausted");
  sg.put(h->d_seq_utt, seq_utt.data(), sizeof(int) * h->n_seqs);
  sg.put(h->d_priv_pt, priv.data(), sizeof(int) * priv.size());
  rc = sg.done();
  if (rc) return rc;
  launch_reset_sampler(h->n_seqs, h->seen_words, h->d_seen, h->d_finished, h->d_gen_count, h->d_seq_pos, h->d_n_active,
                       kDecMaxRows, c.start_mel_token, h->d_attn_cnt, h->n_seqs * kH, st);
  h->launches += 1;
  h->dec_mt_max = mt_max_for(h->n_seqs);
  TTS_CUDA(h, cudaEventRecord(h->ev1, st));
  TTS_CUDA(h, cudaGetLastError());
  h->tokens_done = 0;
  h->first_step_pending = true;
  h->prefilled = true;
  h->timing_valid = false;
  return 0;
}

static int run_steps(TtsArHandle* h, const TtsArSampling* s, int n_steps, bool skip_last_model, long long* out_codes,
                     float* out_latents, const int32_t* forced, int32_t* out_choice, cudaStream_t st, bool streaming = false) {
  const TtsArConfig& c = h->cfg;
  if (!s || s->struct_size != (int)sizeof(TtsArSampling)) return fail(h, TTS_AR_ERR_BAD_ARG, "TtsArSampling.struct_size mismatch");
  if (!out_codes || n_steps < 1) return fail(h, TTS_AR_ERR_BAD_ARG, "bad generate argument");
  if (s->do_sample && !(s->temperature > 0.f)) return fail(h, TTS_AR_ERR_BAD_ARG, "temperature must be > 0");
  if (s->do_sample && !(s->top_p > 0.f)) return fail(h, TTS_AR_ERR_BAD_ARG, "top_p must be > 0");
  if (!(s->repetition_penalty > 0.f)) return fail(h, TTS_AR_ERR_BAD_ARG, "repetition_penalty must be > 0");
  // the model is run (n_steps - skip) more times: positions / mel_pos rows must exist for every fed token
  const int fed = h->tokens_done + n_steps - (skip_last_model ? 1 : 0);
  if (fed > c.max_new_tokens || fed + 1 >= c.n_mel_pos) return fail(h, TTS_AR_ERR_CAPACITY, "generation length exceeds max_new_tokens");
  Params P;
  fill_common_params(h, P);
  P.decode_mode = 1;
  P.row_utt = h->d_seq_utt; P.row_pos = h->d_seq_pos;
  P.mt_max = h->dec_mt_max;
  P.has_sample = 1; P.n_rows = h->n_seqs; P.tail = 1; P.n_head_rows = h->n_seqs;
  P.n_steps = n_steps; P.skip_last_model = skip_last_model ? 1 : 0;
  P.n_seqs = h->n_seqs;
  P.prof = h->d_prof; P.prof_step = h->prof_step;
  P.first_step_from_prefill = h->first_step_pending ? 1 : 0;
  P.do_sample = s->do_sample; P.top_k = s->top_k; P.ignore_eos = s->ignore_eos;
  P.temperature = s->temperature; P.top_p = s->top_p; P.rep_penalty = s->repetition_penalty;
  P.typical_mass = (s->typical_mass > 0.f && s->typical_mass < 1.f) ? s->typical_mass : 0.f;
  P.seed = s->seed; P.seq_index_base = s->seq_index_base; P.seq_indices = (const long long*)s->seq_indices;
  P.out_codes = out_codes; P.out_stride = n_steps; P.out_latents = out_latents; P.forced = forced; P.out_choice = out_choice;
  // Up to 64 sequences are ONE launch of a head-cluster kernel (MODE 1 / 2 / 4: every streamed weight chunk is multiplied
  // with all of the launch's rows).  Larger batches run as consecutive launches over groups of 64: every per-sequence array
  // is a pointer offset; sequences are independent and the Philox key is the global sequence index, so the result is
  // bit-identical to separate calls on the groups.
  const bool cluster = h->use_cluster && !h->opt_decode_interp;
  h->n_groups = cluster ? (h->n_seqs + kDecMaxRows - 1) / kDecMaxRows : 0;
  int rc = 0;
  if (streaming) {
    if (h->n_groups != 1) return fail(h, TTS_AR_ERR_CAPACITY, "streaming needs the head-cluster kernel and at most 64 sequences");
    if (n_steps > h->stream_cap) return fail(h, TTS_AR_ERR_CAPACITY, "stream longer than max_new_tokens");
    memset(h->h_stream_flags, 0, sizeof(int) * (n_steps + 1));   // the previous stream has ended (tts_ar_stream_begin checks)
    P.stream_done = h->d_stream_flags;
    P.stream_flags = h->d_stream_flags + 1;
    h->stream_steps = n_steps;
  }
  if (h->n_groups > 1) {
    const int stash = h->logit_rows_cap - (h->cfg.max_utts + 63) / 64 * 64;   // the prompt-pass rows, kept for every group
    if (h->first_step_pending) {
      TTS_CUDA(h, cudaMemcpyAsync(h->d_logits + (size_t)stash * h->v_stride, h->d_logits, sizeof(float) * h->n_utt * h->v_stride,
                                  cudaMemcpyDeviceToDevice, st));
      TTS_CUDA(h, cudaMemcpyAsync(h->d_cur_latent + (size_t)stash * kD, h->d_cur_latent, sizeof(float) * h->n_utt * kD,
                                  cudaMemcpyDeviceToDevice, st));
    }
    for (int g = 0; g < h->n_groups && !rc; ++g) {
      const int s0 = g * kDecMaxRows, cnt = std::min(kDecMaxRows, h->n_seqs - s0);
      Params Q = P;
      Q.n_rows = Q.n_head_rows = Q.n_seqs = cnt;
      Q.x += (size_t)s0 * kD; Q.q += (size_t)s0 * kD;
      Q.hbuf += (size_t)s0 * kD4;
      Q.logits += (size_t)s0 * h->v_stride; Q.cur_latent += (size_t)s0 * kD;
      Q.logits_pf = h->d_logits + (size_t)stash * h->v_stride; Q.cur_latent_pf = h->d_cur_latent + (size_t)stash * kD;
      Q.row_utt += s0; Q.row_pos += s0; Q.seq_utt += s0; Q.priv_pt += (size_t)s0 * kMaxPagesPerSeg;
      Q.finished += s0; Q.gen_count += s0; Q.seen += (size_t)s0 * h->seen_words;
      Q.n_active = h->d_n_active + 1 + g; Q.steps_done = h->d_steps_done + 1 + g;
      Q.seq_index_base += s0;
      if (Q.seq_indices) Q.seq_indices += s0;
      Q.out_codes += (size_t)s0 * n_steps;
      if (Q.out_latents) Q.out_latents += (size_t)s0 * n_steps * kD;
      if (Q.forced) Q.forced += (size_t)s0 * n_steps;
      if (Q.out_choice) Q.out_choice += (size_t)s0 * n_steps;
      rc = launch(h, Q, st, true);
    }
  } else {
    rc = launch(h, P, st, cluster);
  }
  if (rc) return rc;
  // SURVEY.md 8(d) algorithmic bytes, assuming every sequence stays live (exact under ignore_eos)
  const double W = 2.0 * ((double)c.n_layers * 12596224.0 + 4096.0 + (double)c.n_mel_codes * (kD + 1));
  const double kappa = (double)c.n_layers * 2 * kD * 2;
  double prompt = 0;
  for (int u = 0; u < h->n_utt; ++u) prompt += h->plen[u] * kappa;
  double bytes = 0;
  const int passes = n_steps - (skip_last_model ? 1 : 0);
  for (int k = 0; k < passes; ++k) bytes += W + 4096.0 * h->n_seqs + prompt + (double)h->n_seqs * (h->tokens_done + k + 1) * kappa;
  h->last_bytes = bytes;
  h->last_model_passes = passes;
  h->tokens_done += n_steps;
  h->first_step_pending = false;
  TTS_CUDA(h, cudaEventRecord(h->ev2, st));
  h->timing_valid = true;
  return 0;
}

extern "C" int tts_ar_generate(TtsArHandle* h, const TtsArSampling* s, int32_t max_new, int64_t* out_codes,
                               float* out_latents, const int32_t* forced, int32_t* out_choice, void* stream) {
  int rc = check_ready(h, true);
  if (rc) return rc;
  if (!h->first_step_pending) return fail(h, TTS_AR_ERR_STATE, "tts_ar_generate needs a fresh tts_ar_prefill");
  return run_steps(h, s, max_new, true, (long long*)out_codes, out_latents, forced, out_choice, (cudaStream_t)stream);
}

extern "C" int tts_ar_step(TtsArHandle* h, const TtsArSampling* s, int32_t n_steps, int64_t* out_codes,
                           float* out_latents, const int32_t* forced, int32_t* out_choice, void* stream) {
  int rc = check_ready(h, true);
  if (rc) return rc;
  return run_steps(h, s, n_steps, false, (long long*)out_codes, out_latents, forced, out_choice, (cudaStream_t)stream);
}

extern "C" int tts_ar_logits_debug(TtsArHandle* h, float* out, int32_t* rows_out, void* stream) {
  int rc = check_ready(h, true);
  if (rc) return rc;
  if (!out) return fail(h, TTS_AR_ERR_BAD_ARG, "null argument");
  const int rows = h->first_step_pending ? h->n_utt : h->n_seqs;
  TTS_CUDA(h, cudaMemcpy2DAsync(out, sizeof(float) * h->cfg.n_mel_codes, h->d_logits, sizeof(float) * h->v_stride,
                                sizeof(float) * h->cfg.n_mel_codes, rows, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  if (rows_out) *rows_out = rows;
  return 0;
}

extern "C" int tts_ar_sync_result(TtsArHandle* h, TtsArStats* out, void* stream) {
  if (!h || !out) return TTS_AR_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  TTS_CUDA(h, cudaSetDevice(h->cfg.device));
  TTS_CUDA(h, cudaMemcpyAsync(&h->h_pinned_i[0], h->d_steps_done, sizeof(int), cudaMemcpyDeviceToHost, st));
  TTS_CUDA(h, cudaMemcpyAsync(&h->h_pinned_i[1], h->d_watchdog, sizeof(int), cudaMemcpyDeviceToHost, st));
  TTS_CUDA(h, cudaMemcpyAsync(&h->h_pinned_i[3], h->d_watchdog + 1, sizeof(int), cudaMemcpyDeviceToHost, st));
  TTS_CUDA(h, cudaMemcpyAsync(&h->h_pinned_i[2], h->d_n_active, sizeof(int), cudaMemcpyDeviceToHost, st));
  const int ng = h->n_groups > 1 ? h->n_groups : 0;
  if (ng) {   // per-group counters: the call ran until its slowest group finished
    TTS_CUDA(h, cudaMemcpyAsync(&h->h_pinned_i[8], h->d_steps_done + 1, ng * sizeof(int), cudaMemcpyDeviceToHost, st));
    TTS_CUDA(h, cudaMemcpyAsync(&h->h_pinned_i[8 + kMaxGroups], h->d_n_active + 1, ng * sizeof(int), cudaMemcpyDeviceToHost, st));
  }
  TTS_CUDA(h, cudaStreamSynchronize(st));
  if (ng) {
    int steps = 0, active = 0;
    for (int g = 0; g < ng; ++g) {
      steps = std::max(steps, h->h_pinned_i[8 + g]);
      active += h->h_pinned_i[8 + kMaxGroups + g];
    }
    h->h_pinned_i[0] = steps;
    h->h_pinned_i[2] = active;
  }
  memset(out, 0, sizeof(*out));
  out->struct_size = (int)sizeof(TtsArStats);
  out->steps_run = h->h_pinned_i[0];
  out->watchdog = h->h_pinned_i[1];
  out->all_finished = h->h_pinned_i[2] == 0;
  out->kernel_launches = h->launches;
  out->last_algorithmic_bytes = h->last_bytes;
  out->kv_pages = h->n_pages;
  out->ctas = h->G;
  if (h->timing_valid) {
    float a = 0.f, b = 0.f;
    if (cudaEventElapsedTime(&a, h->ev0, h->ev1) == cudaSuccess) out->last_prefill_ms = a;
    if (cudaEventElapsedTime(&b, h->ev0, h->ev2) == cudaSuccess) out->last_device_ms = b;
    cudaGetLastError();
  }
  if (out->watchdog || h->h_pinned_i[3]) {
    // reported once: the flags are cleared so that the handle stays usable (they are also cleared by the next prefill)
    cudaMemsetAsync(h->d_watchdog, 0, 2 * sizeof(int), st);
    if (out->watchdog) return fail(h, TTS_AR_ERR_WATCHDOG, "device-side wait timed out (watchdog); results are invalid");
    // the reference's nn.Embedding raises IndexError on such an id (unified_voice.py:642, 205)
    return fail(h, TTS_AR_ERR_BAD_ARG, "a text or mel token id is outside its embedding table");
  }
  return 0;
}

extern "C" int tts_ar_forward_latents(TtsArHandle* h, const float* latent, const int32_t* text, int32_t text_stride,
                                      const int32_t* text_len, const int32_t* codes, int32_t codes_stride,
                                      const int32_t* codes_len, int32_t n, float* out, void* stream) {
  int rc = check_ready(h, false);
  if (rc) return rc;
  if (!latent || !text || !text_len || !codes || !codes_len || !out || n < 1) return fail(h, TTS_AR_ERR_BAD_ARG, "bad forward_latents argument");
  const TtsArConfig& c = h->cfg;
  cudaStream_t st = (cudaStream_t)stream;
  h->prefilled = false;   // the KV pool and activations are reused: generation state is gone
  TTS_CUDA(h, cudaMemsetAsync(h->d_watchdog, 0, 2 * sizeof(int), st));
  int i0 = 0;
  while (i0 < n) {
    // sub-batch that fits the row capacity
    int i1 = i0, rows = 0, pages = 0;
    while (i1 < n) {
      if (text_len[i1] < 0 || text_len[i1] > c.max_text_len || codes_len[i1] < 0 || codes_len[i1] + 2 > c.n_mel_pos ||
          text_len[i1] > text_stride || codes_len[i1] > codes_stride)
        return fail(h, TTS_AR_ERR_CAPACITY, "text / codes length out of range");
      const int len = text_len[i1] + 3 + codes_len[i1] + 2;
      if (len > kMaxPagesPerSeg * kPage) return fail(h, TTS_AR_ERR_CAPACITY, "sequence longer than 1024 positions");
      const int np = (len + kPage - 1) / kPage;
      if (rows + len > h->rf_rows_cap || pages + np > h->n_pages || (i1 - i0) >= h->pt_rows) {
        if (i1 == i0) return fail(h, TTS_AR_ERR_CAPACITY, "sequence exceeds re-forward row capacity");
        break;
      }
      rows += len;
      pages += np;
      ++i1;
    }
    const int nb = i1 - i0;
    std::vector<int> seq_len(nb), row_utt, row_pos, latent_dst, none;
    for (int b = 0; b < nb; ++b) {
      const int T = text_len[i0 + b], M = codes_len[i0 + b];
      seq_len[b] = T + 3 + M + 2;
      for (int p = 0; p < seq_len[b]; ++p) {
        row_utt.push_back(b);
        row_pos.push_back(p);
        const int j = p - (T + 3);                          // mel position; keep j = 0 .. M-1 (:272: enc[:, -(M+2):][:, :-2])
        latent_dst.push_back((j >= 0 && j < M) ? (i0 + b) * codes_stride + j : -1);
      }
    }
    Stager sg(h, st);
    if (sg.rc) return sg.rc;
    sg.put(h->d_text_len, text_len + i0, sizeof(int) * nb);
    sg.put(h->d_codes_len, codes_len + i0, sizeof(int) * nb);
    rc = run_prompt_pass(h, sg, seq_len, row_utt, row_pos, none, latent_dst, false, st);
    if (rc) return rc;
    rc = sg.done();
    if (rc) return rc;
    const int n_rows = (int)row_utt.size();
    EmbedArgs ea{};
    ea.x = h->d_x; ea.row_utt = h->d_pf_row_utt; ea.row_pos_in = h->d_pf_row_pos; ea.text_len = h->d_text_len;
    ea.text = text + (size_t)i0 * text_stride; ea.text_stride = text_stride;
    ea.codes = codes + (size_t)i0 * codes_stride; ea.codes_len = h->d_codes_len; ea.codes_stride = codes_stride;
    ea.latent = latent + (size_t)i0 * kD;
    ea.text_emb = h->d_text_emb; ea.text_pos = h->d_text_pos; ea.mel_emb = h->d_mel_emb; ea.mel_pos = h->d_mel_pos;
    ea.start_text = c.start_text_token; ea.stop_text = c.stop_text_token; ea.start_mel = c.start_mel_token; ea.stop_mel = c.stop_mel_token;
    ea.n_text_rows = c.n_text_rows; ea.n_mel_codes = c.n_mel_codes; ea.err_flag = h->d_watchdog + 1;
    launch_embed_rows(ea, n_rows, st);
    h->launches += 1;
    Params P;
    fill_common_params(h, P);
    P.decode_mode = 0;
    P.row_utt = h->d_pf_row_utt; P.row_pos = h->d_pf_row_pos;
    P.mt_max = mt_max_for(n_rows);
    P.has_sample = 0; P.n_rows = n_rows; P.tail = 2; P.n_head_rows = 0;
    P.n_steps = 1;
    P.latent_out = out;
    rc = h->opt_dense_interp ? launch(h, P, st) : run_dense_pass(h, P, st);
    if (rc) return rc;
    i0 = i1;
  }
  TTS_CUDA(h, cudaGetLastError());
  return 0;
}

extern "C" int tts_ar_inference_speech_host(TtsArHandle* h, const float* latent_host, const int32_t* text_host,
                                            int32_t text_stride, const int32_t* text_len, int32_t n_utt, int32_t n_cand,
                                            const TtsArSampling* s, int32_t max_new, int64_t* out_codes_host,
                                            int32_t* steps_run, void* stream) {
  int rc = check_ready(h, false);
  if (rc) return rc;
  const TtsArConfig& c = h->cfg;
  if (!latent_host || !text_host || !text_len || !out_codes_host || n_utt < 1 || n_cand < 1 || max_new < 1)
    return fail(h, TTS_AR_ERR_BAD_ARG, "bad inference_speech argument");
  if (n_utt > c.max_utts || n_utt * n_cand > c.max_seqs || text_stride > std::max(c.max_text_len, 1) || max_new > c.max_new_tokens)
    return fail(h, TTS_AR_ERR_CAPACITY, "batch exceeds capacity");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t nb_lat = sizeof(float) * (size_t)n_utt * kD, nb_txt = sizeof(int) * (size_t)n_utt * text_stride;
  memcpy(h->h_pinned_latent, latent_host, nb_lat);
  memcpy(h->h_pinned_text, text_host, nb_txt);
  TTS_CUDA(h, cudaMemcpyAsync(h->d_in_latent, h->h_pinned_latent, nb_lat, cudaMemcpyHostToDevice, st));
  TTS_CUDA(h, cudaMemcpyAsync(h->d_in_text, h->h_pinned_text, nb_txt, cudaMemcpyHostToDevice, st));
  rc = tts_ar_prefill(h, h->d_in_latent, h->d_in_text, text_stride, text_len, n_utt, n_cand, st);
  if (rc) return rc;
  launch_fill_i64((long long*)h->d_out_codes, (size_t)n_utt * n_cand * max_new, c.stop_mel_token, st);   // columns a group that finished early never writes
  rc = tts_ar_generate(h, s, max_new, (int64_t*)h->d_out_codes, nullptr, nullptr, nullptr, st);
  if (rc) return rc;
  const size_t nb_out = sizeof(long long) * (size_t)h->n_seqs * max_new;
  TTS_CUDA(h, cudaMemcpyAsync(h->h_pinned_codes, h->d_out_codes, nb_out, cudaMemcpyDeviceToHost, st));
  TtsArStats stt;
  rc = tts_ar_sync_result(h, &stt, st);
  if (rc) return rc;
  memcpy(out_codes_host, h->h_pinned_codes, nb_out);
  if (steps_run) *steps_run = stt.steps_run;
  return 0;
}

extern "C" int tts_ar_fixup_codes(TtsArHandle* h, int64_t* codes, int32_t rows, int32_t len, int32_t stride, int32_t calm_token,
                                  int32_t* first_stop, int32_t* trim_len, void* stream) {
  if (!h || !codes) return TTS_AR_ERR_BAD_ARG;
  if (rows < 0 || len < 3 || stride < len) return fail(h, TTS_AR_ERR_BAD_ARG, "fixup_codes: need rows >= 0, len >= 3, stride >= len");
  if (rows == 0) return 0;
  launch_fixup_codes(reinterpret_cast<long long*>(codes), rows, len, stride, h->cfg.stop_mel_token, calm_token, first_stop,
                     trim_len, (cudaStream_t)stream);
  TTS_CUDA(h, cudaGetLastError());
  h->launches += 1;
  return 0;
}

extern "C" int tts_ar_debug_copy(TtsArHandle* h, const char* name, void* dst, uint64_t bytes, void* stream) {
  if (!h || !name || !dst) return TTS_AR_ERR_BAD_ARG;
  const void* src = nullptr;
  const std::string n(name);
  if (n == "x") src = h->d_x;
  else if (n == "q") src = h->d_q;
  else if (n == "logits") src = h->d_logits;
  else if (n == "cur_latent") src = h->d_cur_latent;
  else if (n == "kv") src = h->d_kv;
  else if (n == "ctx") src = h->d_ctx;
  else if (n == "h") src = h->d_h;
  else if (n == "prof") src = h->d_prof;
  else return fail(h, TTS_AR_ERR_BAD_ARG, "unknown debug buffer");
  TTS_CUDA(h, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return 0;
}

extern "C" int tts_ar_set_option(TtsArHandle* h, const char* name, int32_t value) {
  if (!h || !name) return TTS_AR_ERR_BAD_ARG;
  const std::string n(name);
  if (n == "decode_kernel") {            // 0: head-cluster kernels when co-resident (default); 1: the interpreter kernel
    if (value != 0 && value != 1) return fail(h, TTS_AR_ERR_BAD_ARG, "decode_kernel: 0 (auto) or 1 (interpreter)");
    h->opt_decode_interp = value;
  } else if (n == "dense_kernel") {      // 0: tcgen05 pipeline (default); 1: the interpreter kernel in prompt mode
    if (value != 0 && value != 1) return fail(h, TTS_AR_ERR_BAD_ARG, "dense_kernel: 0 (tcgen05) or 1 (interpreter)");
    h->opt_dense_interp = value;
  } else if (n == "attn_tiles") {        // prompt-pass attention: 0 auto, 1 one unit per (row, head), 2 always 16-row tiles
    if (value < 0 || value > 2) return fail(h, TTS_AR_ERR_BAD_ARG, "attn_tiles: 0, 1 or 2");
    h->opt_attn_tiles = value;
  } else if (n == "cooperative_launch") {   // 1 (default): cooperative cluster launch; 0: plain cluster launch
    h->opt_no_coop = value ? 0 : 1;
  } else if (n == "prof_step") {         // per-stage timestamps of one decode step (scripts/stage_prof.py); -1 off
    h->prof_step = value;
    if (value >= 0 && !h->d_prof) {
      void* p = nullptr;
      cudaError_t e = cudaMalloc(&p, 16384 * sizeof(unsigned long long));
      if (e != cudaSuccess) return cuda_fail(h, e, "cudaMalloc(prof)");
      h->allocs.push_back(p);
      h->d_prof = reinterpret_cast<unsigned long long*>(p);
    }
  } else {
    return fail(h, TTS_AR_ERR_BAD_ARG, "unknown option");
  }
  return 0;
}

// ---- streaming hand-off (reference get_generator / sample_stream yield per step, stream_generator.py:1172) ----
extern "C" int tts_ar_stream_begin(TtsArHandle* h, const TtsArSampling* s, int32_t max_new, int64_t* out_codes, float* out_latents,
                                   void* stream) {
  int rc = check_ready(h, true);
  if (rc) return rc;
  if (!h->first_step_pending) return fail(h, TTS_AR_ERR_STATE, "tts_ar_stream_begin needs a fresh tts_ar_prefill");
  if (h->streaming) return fail(h, TTS_AR_ERR_STATE, "a stream is already open: call tts_ar_stream_end first");
  if (!out_latents) return fail(h, TTS_AR_ERR_BAD_ARG, "streaming publishes (token, latent) pairs: out_latents is required");
  rc = run_steps(h, s, max_new, true, (long long*)out_codes, out_latents, nullptr, nullptr, (cudaStream_t)stream, true);
  if (rc) return rc;
  h->streaming = true;
  return 0;
}

extern "C" int tts_ar_stream_poll(TtsArHandle* h, int32_t step, int32_t timeout_us) {
  if (!h || !h->streaming) return TTS_AR_ERR_STATE;
  if (step < 0 || step >= h->stream_steps) return 2;
  volatile int* f = h->h_stream_flags;
  const auto t0 = std::chrono::steady_clock::now();
  for (;;) {
    if (f[1 + step] >= h->n_seqs) return 1;                         // every sequence's (token, latent) of this step is in the device buffers
    const int done = f[0];
    if (done && step >= done - 1) return 2;                          // the launch ended (all sequences finished) before this step
    if (timeout_us >= 0 &&
        std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t0).count() > timeout_us)
      return 0;
  }
}

extern "C" int tts_ar_stream_end(TtsArHandle* h, TtsArStats* out, void* stream) {
  if (!h || !h->streaming) return TTS_AR_ERR_STATE;
  h->streaming = false;
  TtsArStats tmp;
  return tts_ar_sync_result(h, out ? out : &tmp, stream);
}
