// tts_ar.cu -- C ABI: handle lifetime, capacity planning, weight snapshot/repack.  (The call path --
// prompt pass, sample loop, latent re-forward -- is in tts_ar_run.cu.)
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <algorithm>
#include "tts_ar_host.h"

using namespace tts;

static std::string g_create_error;

namespace tts {
int fail(TtsArHandle* h, int code, const std::string& msg) {
  if (h) h->err = msg; else g_create_error = msg;
  return code;
}
int cuda_fail(TtsArHandle* h, cudaError_t e, const char* what) {
  std::string m = std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what;
  cudaGetLastError();
  return fail(h, TTS_AR_ERR_CUDA, m);
}
}  // namespace tts

template <typename T>
static int dalloc(TtsArHandle* h, T** p, size_t count) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, std::max<size_t>(count * sizeof(T), 256));
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaMalloc");
  h->allocs.push_back(q);
  *p = reinterpret_cast<T*>(q);
  return 0;
}
#define DALLOC(h, p, n)                     \
  do {                                      \
    int rc__ = dalloc((h), &(p), (size_t)(n)); \
    if (rc__) return rc__;                  \
  } while (0)

extern "C" int tts_ar_abi_version(void) { return TTS_AR_ABI_VERSION; }

extern "C" const char* tts_ar_last_error(TtsArHandle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

static int create_impl(TtsArHandle* h) {
  const TtsArConfig& c = h->cfg;
  TTS_CUDA(h, cudaSetDevice(c.device));
  cudaDeviceProp prop;
  TTS_CUDA(h, cudaGetDeviceProperties(&prop, c.device));
  if (prop.major != 10) return fail(h, TTS_AR_ERR_DEVICE, "tts_ar needs an sm_100 (Blackwell B200) device; kernels are built for sm_100a only");
  int coop = 0;
  TTS_CUDA(h, cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c.device));
  if (!coop) return fail(h, TTS_AR_ERR_DEVICE, "device does not support cooperative launch");
  h->G = prop.multiProcessorCount;
  h->smem_budget = (int)prop.sharedMemPerBlockOptin;
  if (smem_max_slots(4, h->smem_budget) < 2) return fail(h, TTS_AR_ERR_DEVICE, "not enough shared memory per block");

  h->plen_max = c.max_text_len + 4;                      // [latent, start_text, text.., stop_text, start_mel]
  h->npp = (h->plen_max + kPage - 1) / kPage;
  h->nqp = (c.max_new_tokens + kPage - 1) / kPage;
  const int rf_len = c.max_text_len + c.max_new_tokens + 5;   // latent + (T+2) + (M+2)
  if (rf_len > kMaxPagesPerSeg * kPage) return fail(h, TTS_AR_ERR_BAD_ARG, "max_text_len + max_new_tokens + 5 exceeds 1024 positions");
  const int rf_batch = std::max(1, std::min(c.max_seqs, 65536 / rf_len));
  h->rf_rows_cap = rf_batch * rf_len;
  const int rf_pages = rf_batch * ((rf_len + kPage - 1) / kPage);
  h->n_pages = std::max(c.max_utts * h->npp + c.max_seqs * h->nqp, rf_pages);
  h->rows_cap = std::max({c.max_seqs, c.max_utts * h->plen_max, h->rf_rows_cap});
  h->rows_cap = (h->rows_cap + 127) / 128 * 128;    // 128-row GEMM instructions read whole m-tiles
  h->logit_rows_cap = (std::max(c.max_seqs, c.max_utts) + 63) / 64 * 64 + (c.max_utts + 63) / 64 * 64;   // + the prompt-pass rows kept for later groups
  h->plan_cap_units = std::max(h->G * kNW, 1);
  h->seen_words = (c.n_mel_codes + 31) / 32;
  h->v_tiles = (c.n_mel_codes + 7) / 8;
  h->v_stride = h->v_tiles * 8;

  const size_t L = (size_t)c.n_layers;
  // packed weight arena: per layer 384 + 128 + 512 + 512 chunks; head v_tiles chunks
  const size_t chunks_per_layer = kD3 / 8 + kD / 8 + kD4 / 8 + (kD / 8) * 4 + kProjkChunks;
  const size_t u4_per_chunk = kChunkBytes / 16;
  DALLOC(h, h->d_warena, (L * chunks_per_layer + h->v_tiles) * u4_per_chunk);
  h->w_layer_u4 = (long long)(chunks_per_layer * u4_per_chunk);
  h->d_w_head = h->d_warena + L * chunks_per_layer * u4_per_chunk;
  // fp32 parameter arena: per layer ln1 g/b, ln2 g/b (4*1024), b_qkv 3072, b_proj 1024, b_fc 4096, b_proj2 1024
  const size_t fp_per_layer = 4 * kD + kD3 + kD + kD4 + kD;
  DALLOC(h, h->d_fparams, L * fp_per_layer + 4 * kD + h->v_stride);
  h->fp_layer = (int)fp_per_layer;
  DALLOC(h, h->d_layers, L);
  h->layers_host.resize(L);
  for (size_t i = 0; i < L; ++i) {
    LayerDev& ld = h->layers_host[i];
    uint4* wb = h->d_warena + i * chunks_per_layer * u4_per_chunk;
    ld.w_qkv = wb;
    ld.w_proj = wb + (size_t)(kD3 / 8) * u4_per_chunk;
    ld.w_fc = ld.w_proj + (size_t)(kD / 8) * u4_per_chunk;
    ld.w_proj2 = ld.w_fc + (size_t)(kD4 / 8) * u4_per_chunk;
    ld.w_projk = ld.w_proj2 + (size_t)(kD / 8) * 4 * u4_per_chunk;
    float* fb = h->d_fparams + i * fp_per_layer;
    ld.ln1_g = fb; ld.ln1_b = fb + kD; ld.ln2_g = fb + 2 * kD; ld.ln2_b = fb + 3 * kD;
    ld.b_qkv = fb + 4 * kD; ld.b_proj = fb + 4 * kD + kD3; ld.b_fc = ld.b_proj + kD; ld.b_proj2 = ld.b_fc + kD4;
  }
  {
    float* fb = h->d_fparams + L * fp_per_layer;
    h->d_lnf_g = fb; h->d_lnf_b = fb + kD; h->d_fn_g = fb + 2 * kD; h->d_fn_b = fb + 3 * kD; h->d_b_head = fb + 4 * kD;
  }
  DALLOC(h, h->d_mel_emb, (size_t)c.n_mel_codes * kD);
  DALLOC(h, h->d_mel_pos, (size_t)c.n_mel_pos * kD);
  DALLOC(h, h->d_text_emb, (size_t)c.n_text_rows * kD);
  DALLOC(h, h->d_text_pos, (size_t)c.n_text_pos * kD);

  DALLOC(h, h->d_x, (size_t)h->rows_cap * kD);
  DALLOC(h, h->d_q, (size_t)h->rows_cap * kD);
  DALLOC(h, h->d_y, (size_t)kDecMaxRows * kD);
  DALLOC(h, h->d_xt, (size_t)kDecMaxRows * kD);
  DALLOC(h, h->d_ctx, (size_t)h->rows_cap * kD);
  DALLOC(h, h->d_xn, (size_t)h->rows_cap * kD);
  DALLOC(h, h->d_h, (size_t)h->rows_cap * kD4);
  DALLOC(h, h->d_logits, (size_t)h->logit_rows_cap * h->v_stride);
  DALLOC(h, h->d_cur_latent, (size_t)h->logit_rows_cap * kD);
  DALLOC(h, h->d_part, (size_t)h->plan_cap_units * kPartStride);
  DALLOC(h, h->d_kv, (size_t)c.n_layers * h->n_pages * 2 * kH * kPage * kHD);
  const int pt_rows = std::max(c.max_utts, rf_batch);
  h->pt_rows = pt_rows;
  DALLOC(h, h->d_prefix_pt, (size_t)pt_rows * kMaxPagesPerSeg);
  DALLOC(h, h->d_priv_pt, (size_t)c.max_seqs * kMaxPagesPerSeg);
  DALLOC(h, h->d_utt_plen, pt_rows);
  DALLOC(h, h->d_text_len, pt_rows);
  DALLOC(h, h->d_codes_len, pt_rows);
  DALLOC(h, h->d_pf_row_utt, h->rows_cap);
  DALLOC(h, h->d_pf_row_pos, h->rows_cap);
  DALLOC(h, h->d_latent_dst, h->rows_cap);
  DALLOC(h, h->d_head_rows, h->logit_rows_cap);
  DALLOC(h, h->d_seq_utt, c.max_seqs);
  DALLOC(h, h->d_seq_pos, c.max_seqs);
  DALLOC(h, h->d_attn_cnt, (size_t)c.max_seqs * kH);
  DALLOC(h, h->d_plan, 8 + c.max_seqs + 1);
  DALLOC(h, h->d_finished, c.max_seqs);
  DALLOC(h, h->d_gen_count, c.max_seqs);
  DALLOC(h, h->d_n_active, 4 + kMaxGroups);
  DALLOC(h, h->d_seen, (size_t)c.max_seqs * h->seen_words);
  DALLOC(h, h->d_bar, 4);
  DALLOC(h, h->d_watchdog, 4);
  DALLOC(h, h->d_steps_done, 4 + kMaxGroups);
  DALLOC(h, h->d_in_latent, (size_t)c.max_utts * kD);
  DALLOC(h, h->d_in_text, (size_t)c.max_utts * std::max(c.max_text_len, 1));
  DALLOC(h, h->d_out_codes, (size_t)c.max_seqs * std::max(c.max_new_tokens, 1));
  TTS_CUDA(h, cudaMallocHost(&h->h_pinned_i, sizeof(int) * (8 + 2 * kMaxGroups)));
  // staging for the per-call index uploads: page tables, lengths, row maps (all int32)
  h->staging_bytes = sizeof(int) * ((size_t)pt_rows * (kMaxPagesPerSeg + 4) + 3 * (size_t)h->rows_cap + (size_t)h->logit_rows_cap +
                                    (size_t)c.max_seqs * (kMaxPagesPerSeg + 2)) + 4096;
  TTS_CUDA(h, cudaMallocHost(&h->h_staging, h->staging_bytes));
  TTS_CUDA(h, cudaEventCreateWithFlags(&h->ev_staging, cudaEventDisableTiming));
  h->stream_cap = c.max_new_tokens + 8;     // [0]: done flag, [1 + step]: per-step counters
  TTS_CUDA(h, cudaHostAlloc(&h->h_stream_flags, sizeof(int) * (h->stream_cap + 1), cudaHostAllocMapped));
  TTS_CUDA(h, cudaHostGetDevicePointer(&h->d_stream_flags, h->h_stream_flags, 0));
  memset(h->h_stream_flags, 0, sizeof(int) * (h->stream_cap + 1));
  TTS_CUDA(h, cudaMallocHost(&h->h_pinned_codes, sizeof(long long) * (size_t)c.max_seqs * std::max(c.max_new_tokens, 1)));
  TTS_CUDA(h, cudaMallocHost(&h->h_pinned_latent, sizeof(float) * (size_t)c.max_utts * kD));
  TTS_CUDA(h, cudaMallocHost(&h->h_pinned_text, sizeof(int) * (size_t)c.max_utts * std::max(c.max_text_len, 1)));
  TTS_CUDA(h, cudaMemset(h->d_attn_cnt, 0, sizeof(int) * (size_t)c.max_seqs * kH));
  TTS_CUDA(h, cudaMemset(h->d_watchdog, 0, 16));
  TTS_CUDA(h, cudaMemset(h->d_kv, 0, sizeof(__nv_bfloat16) * (size_t)c.n_layers * h->n_pages * 2 * kH * kPage * kHD));
  TTS_CUDA(h, cudaEventCreate(&h->ev0));
  TTS_CUDA(h, cudaEventCreate(&h->ev1));
  TTS_CUDA(h, cudaEventCreate(&h->ev2));
  const int smem1 = smem_total_bytes(1, smem_max_slots(1, h->smem_budget));
  if (program_max_ctas_per_sm(smem1) < 1) return fail(h, TTS_AR_ERR_DEVICE, "interpreter kernel does not fit on an SM");
  {
    int ncl = 1 << 30;
    for (int mode = 1; mode <= 4; mode *= 2) {
      const int mt = decode_mt_smem(mode);
      ncl = std::min(ncl, decode_max_clusters(mode, smem_total_bytes(mt, smem_max_slots(mt, h->smem_budget)), c.device));
    }
    const int smemd = smem_total_bytes(1, smem_max_slots(1, h->smem_budget));
    h->use_cluster = ncl >= kDecClusters;
    if (getenv("TTS_AR_VERBOSE")) fprintf(stderr, "[tts_ar] co-resident clusters of %d CTAs: %d (smem %d B) -> decode kernel: %s\n",
                                          kClusterSize, ncl, smemd, h->use_cluster ? "head-cluster" : "interpreter");
  }
  {
    int rc = dense_create(h);   // K-major weight image + TMA tensor maps of the tcgen05 passes
    if (rc) return rc;
  }
  TTS_CUDA(h, cudaDeviceSynchronize());
  return 0;
}

extern "C" int tts_ar_create(const TtsArConfig* cfg, TtsArHandle** out) {
  if (!cfg || !out) return fail(nullptr, TTS_AR_ERR_BAD_ARG, "null argument");
  if (cfg->struct_size != (int)sizeof(TtsArConfig)) return fail(nullptr, TTS_AR_ERR_BAD_ARG, "TtsArConfig.struct_size mismatch");
  if (cfg->model_dim != kD || cfg->n_heads != kH) return fail(nullptr, TTS_AR_ERR_BAD_ARG, "model_dim must be 1024 and n_heads 16");
  if (cfg->n_layers < 1 || cfg->n_mel_codes < 2 || cfg->n_mel_codes > 8256 || cfg->max_utts < 1 || cfg->max_seqs < cfg->max_utts ||
      cfg->max_text_len < 1 || cfg->max_text_len + 2 > cfg->n_text_pos || cfg->max_new_tokens < 1 ||
      cfg->max_new_tokens + 2 > cfg->n_mel_pos || cfg->stop_mel_token >= cfg->n_mel_codes || cfg->start_mel_token >= cfg->n_mel_codes ||
      cfg->start_text_token >= cfg->n_text_rows || cfg->max_seqs > kDecMaxRows * kMaxGroups)
    return fail(nullptr, TTS_AR_ERR_BAD_ARG, "TtsArConfig out of range");
  TtsArHandle* h = new TtsArHandle();
  h->cfg = *cfg;
  int rc = create_impl(h);
  if (rc) {
    g_create_error = h->err;
    tts_ar_destroy(h);
    return rc;
  }
  *out = h;
  return 0;
}

extern "C" void tts_ar_destroy(TtsArHandle* h) {
  if (!h) return;
  cudaSetDevice(h->cfg.device);
  cudaDeviceSynchronize();
  for (void* p : h->allocs) cudaFree(p);
  if (h->h_pinned_i) cudaFreeHost(h->h_pinned_i);
  if (h->h_pinned_codes) cudaFreeHost(h->h_pinned_codes);
  if (h->h_pinned_latent) cudaFreeHost(h->h_pinned_latent);
  if (h->h_pinned_text) cudaFreeHost(h->h_pinned_text);
  if (h->h_staging) cudaFreeHost(h->h_staging);
  if (h->h_stream_flags) cudaFreeHost(h->h_stream_flags);
  if (h->ev_staging) cudaEventDestroy(h->ev_staging);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->ev2) cudaEventDestroy(h->ev2);
  delete h;
}

extern "C" int tts_ar_load_weights(TtsArHandle* h, const TtsArWeights* w, void* stream) {
  if (!h || !w) return fail(h, TTS_AR_ERR_BAD_ARG, "null argument");
  if (w->struct_size != (int)sizeof(TtsArWeights)) return fail(h, TTS_AR_ERR_BAD_ARG, "TtsArWeights.struct_size mismatch");
  if (w->n_layers != h->cfg.n_layers || !w->layers) return fail(h, TTS_AR_ERR_BAD_ARG, "layer count mismatch");
  if (w->dtype != 0 && w->dtype != 1) return fail(h, TTS_AR_ERR_BAD_ARG, "dtype must be 0 (float32) or 1 (bfloat16)");
  cudaStream_t st = (cudaStream_t)stream;
  TTS_CUDA(h, cudaSetDevice(h->cfg.device));
  const int dt = w->dtype;
  const TtsArConfig& c = h->cfg;
  for (int i = 0; i < c.n_layers; ++i) {
    const TtsArLayerWeights& lw = w->layers[i];
    const LayerDev& ld = h->layers_host[i];
    const void* ptrs[] = {lw.ln_1_w, lw.ln_1_b, lw.c_attn_w, lw.c_attn_b, lw.attn_proj_w, lw.attn_proj_b,
                          lw.ln_2_w, lw.ln_2_b, lw.c_fc_w, lw.c_fc_b, lw.mlp_proj_w, lw.mlp_proj_b};
    for (const void* p : ptrs)
      if (!p) return fail(h, TTS_AR_ERR_BAD_ARG, "null weight pointer in layer");
    launch_pack_weight(lw.c_attn_w, dt, const_cast<uint4*>(ld.w_qkv), kD, kD3, kD3 / 8, 0, st);
    launch_pack_weight(lw.attn_proj_w, dt, const_cast<uint4*>(ld.w_proj), kD, kD, kD / 8, 0, st);
    launch_pack_weight(lw.c_fc_w, dt, const_cast<uint4*>(ld.w_fc), kD, kD4, kD4 / 8, 0, st);
    launch_pack_weight(lw.mlp_proj_w, dt, const_cast<uint4*>(ld.w_proj2), kD4, kD, kD / 8, 0, st);
    launch_pack_projk(lw.attn_proj_w, dt, const_cast<uint4*>(ld.w_projk), st);
    dense_pack_layer(h, i, lw, dt, st);
    launch_to_f32(lw.ln_1_w, dt, const_cast<float*>(ld.ln1_g), kD, kD, st);
    launch_to_f32(lw.ln_1_b, dt, const_cast<float*>(ld.ln1_b), kD, kD, st);
    launch_to_f32(lw.ln_2_w, dt, const_cast<float*>(ld.ln2_g), kD, kD, st);
    launch_to_f32(lw.ln_2_b, dt, const_cast<float*>(ld.ln2_b), kD, kD, st);
    launch_to_f32(lw.c_attn_b, dt, const_cast<float*>(ld.b_qkv), kD3, kD3, st);
    launch_to_f32(lw.attn_proj_b, dt, const_cast<float*>(ld.b_proj), kD, kD, st);
    launch_to_f32(lw.c_fc_b, dt, const_cast<float*>(ld.b_fc), kD4, kD4, st);
    launch_to_f32(lw.mlp_proj_b, dt, const_cast<float*>(ld.b_proj2), kD, kD, st);
    h->launches += 13;
  }
  const void* tops[] = {w->ln_f_w, w->ln_f_b, w->final_norm_w, w->final_norm_b, w->mel_head_w, w->mel_head_b,
                        w->mel_embedding, w->mel_pos_embedding, w->text_embedding, w->text_pos_embedding};
  for (const void* p : tops)
    if (!p) return fail(h, TTS_AR_ERR_BAD_ARG, "null weight pointer");
  launch_pack_weight(w->mel_head_w, dt, h->d_w_head, kD, c.n_mel_codes, h->v_tiles, 1, st);
  dense_pack_head(h, w->mel_head_w, dt, st);
  launch_to_f32(w->mel_head_b, dt, h->d_b_head, c.n_mel_codes, h->v_stride, st);
  launch_to_f32(w->ln_f_w, dt, h->d_lnf_g, kD, kD, st);
  launch_to_f32(w->ln_f_b, dt, h->d_lnf_b, kD, kD, st);
  launch_to_f32(w->final_norm_w, dt, h->d_fn_g, kD, kD, st);
  launch_to_f32(w->final_norm_b, dt, h->d_fn_b, kD, kD, st);
  launch_to_f32(w->mel_embedding, dt, h->d_mel_emb, (size_t)c.n_mel_codes * kD, (size_t)c.n_mel_codes * kD, st);
  launch_to_f32(w->mel_pos_embedding, dt, h->d_mel_pos, (size_t)c.n_mel_pos * kD, (size_t)c.n_mel_pos * kD, st);
  launch_to_f32(w->text_embedding, dt, h->d_text_emb, (size_t)c.n_text_rows * kD, (size_t)c.n_text_rows * kD, st);
  launch_to_f32(w->text_pos_embedding, dt, h->d_text_pos, (size_t)c.n_text_pos * kD, (size_t)c.n_text_pos * kD, st);
  h->launches += 10;
  TTS_CUDA(h, cudaMemcpyAsync(h->d_layers, h->layers_host.data(), sizeof(LayerDev) * c.n_layers, cudaMemcpyHostToDevice, st));
  TTS_CUDA(h, cudaGetLastError());
  TTS_CUDA(h, cudaStreamSynchronize(st));   // layers_host is pageable; also surfaces pack-kernel faults here
  h->weights_loaded = true;
  h->prefilled = false;
  return 0;
}
