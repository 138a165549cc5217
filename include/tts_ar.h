/* tts_ar.h -- C ABI of the B200-native UnifiedVoice autoregressive (AR) mel-code sampler.
 *
 * Drop-in boundary for ONE path of e-c-k-e-r/tortoise-tts: everything below
 * `UnifiedVoice.inference_speech / compute_embeddings+get_generator / forward(return_latent=True)`
 * (reference tortoise_tts/models/unified_voice.py:614-679, 544-600), i.e. the GPT-2 prefill, the
 * KV-cached per-step decode, the mel head and the HF logits processors + sample loop
 * (reference tortoise_tts/models/stream_generator.py:56-101, 1106-1190).  The reference has no FFI of
 * its own (it is 100% Python over torch/HF); the Python adapter `tortoise_tts_b200.UnifiedVoiceB200`
 * keeps the reference's method signatures and binds these entry points with ctypes
 * (see INTEGRATION.md).  Plain C: opaque handle, raw pointers and sizes, no torch / C++ types.
 *
 * Conventions
 *  - every call returns 0 on success or a negative TtsArStatus; nothing aborts or throws.
 *  - `stream` is a cudaStream_t passed as void*; all device work is enqueued on it; calls do not
 *    synchronise unless documented ("SYNC").  One handle == one device; calls on a handle must be
 *    serialised by the caller (the reference model object is not re-entrant either:
 *    GPT2InferenceModel.cached_mel_emb, unified_voice.py:109,143-144).
 *  - "dev" pointers are device memory owned by the caller (e.g. torch tensors); "host" pointers are
 *    ordinary host memory.
 */
#ifndef TTS_AR_H_
#define TTS_AR_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TTS_AR_ABI_VERSION 2
#define TTS_AR_MODEL_DIM 1024   /* compiled-in: d_model, 16 heads x 64 (unified_voice.py:336-338) */
#define TTS_AR_HEADS 16
#define TTS_AR_HEAD_DIM 64

typedef enum {
  TTS_AR_OK = 0,
  TTS_AR_ERR_BAD_ARG = -1,      /* reference raises AssertionError / ValueError here */
  TTS_AR_ERR_CAPACITY = -2,     /* batch / length exceeds what tts_ar_create was sized for */
  TTS_AR_ERR_CUDA = -3,         /* a CUDA call failed; tts_ar_last_error() has the text */
  TTS_AR_ERR_STATE = -4,        /* call order: weights not loaded / prefill not done */
  TTS_AR_ERR_DEVICE = -5,       /* not an sm_100 device, or cooperative launch unsupported */
  TTS_AR_ERR_WATCHDOG = -6      /* a device-side wait timed out (reported by tts_ar_sync_result) */
} TtsArStatus;

typedef struct TtsArHandle TtsArHandle;

/* Model geometry (defaults = UnifiedVoice.__init__, unified_voice.py:335-354) + capacity. */
typedef struct {
  int32_t struct_size;          /* sizeof(TtsArConfig), for ABI checks */
  int32_t device;               /* CUDA device ordinal */
  int32_t n_layers;             /* 30 */
  int32_t model_dim;            /* must equal TTS_AR_MODEL_DIM */
  int32_t n_heads;              /* must equal TTS_AR_HEADS */
  int32_t n_mel_codes;          /* 8194 (rows of mel_embedding / mel_head) */
  int32_t n_text_rows;          /* 256 (rows of text_embedding) */
  int32_t n_mel_pos;            /* 608 rows of mel_pos_embedding */
  int32_t n_text_pos;           /* 404 rows of text_pos_embedding */
  int32_t start_text_token;     /* 255 */
  int32_t stop_text_token;      /* 0 */
  int32_t start_mel_token;      /* 8192 */
  int32_t stop_mel_token;       /* 8193 (eos == pad, unified_voice.py:664) */
  int32_t max_utts;             /* capacity: utterances per prefill */
  int32_t max_seqs;             /* capacity: sequences (utterances x candidates) */
  int32_t max_text_len;         /* capacity: text tokens per utterance (<= n_text_pos - 2) */
  int32_t max_new_tokens;       /* capacity: generated tokens per sequence (<= n_mel_pos - 2) */
} TtsArConfig;

/* One GPT-2 block; names are the reference state_dict keys under gpt.h.{i}..
 * Conv1D weights are [in, out] row-major exactly as stored by HF. */
typedef struct {
  const void *ln_1_w, *ln_1_b;          /* [1024] */
  const void *c_attn_w, *c_attn_b;      /* [1024,3072], [3072] */
  const void *attn_proj_w, *attn_proj_b;/* [1024,1024], [1024] */
  const void *ln_2_w, *ln_2_b;          /* [1024] */
  const void *c_fc_w, *c_fc_b;          /* [1024,4096], [4096] */
  const void *mlp_proj_w, *mlp_proj_b;  /* [4096,1024], [1024] */
} TtsArLayerWeights;

typedef struct {
  int32_t struct_size;
  int32_t dtype;                        /* 0 = float32, 1 = bfloat16 (all tensors, dev pointers) */
  int32_t n_layers;
  const TtsArLayerWeights* layers;      /* host array of n_layers entries (dev pointers inside) */
  const void *ln_f_w, *ln_f_b;          /* gpt.ln_f */
  const void *final_norm_w, *final_norm_b;
  const void *mel_head_w, *mel_head_b;  /* [n_mel_codes,1024] (nn.Linear [out,in]), [n_mel_codes] */
  const void *mel_embedding;            /* [n_mel_codes,1024] */
  const void *mel_pos_embedding;        /* [n_mel_pos,1024] */
  const void *text_embedding;           /* [n_text_rows,1024] */
  const void *text_pos_embedding;       /* [n_text_pos,1024] */
} TtsArWeights;

/* hf_generate_kwargs the reference's callers pass (inference.py:334-346, train.py:119-129). */
typedef struct {
  int32_t struct_size;
  int32_t do_sample;            /* 0: argmax of processed scores (HF greedy) */
  float temperature;            /* TemperatureLogitsWarper, skipped when == 1 */
  int32_t top_k;                /* TopKLogitsWarper, 0 = off */
  float top_p;                  /* TopPLogitsWarper, >= 1 = off */
  float repetition_penalty;     /* RepetitionPenaltyLogitsProcessor over ALL ids incl. dummy prompt */
  uint64_t seed;                /* Philox key */
  int64_t seq_index_base;       /* global index of local sequence 0 (multi-GPU invariance) */
  int32_t ignore_eos;           /* benchmarking: never finish, run exactly max_new steps */
  float typical_mass;           /* TypicalLogitsWarper(mass) of unified_voice.py:47-75, run as a logits PROCESSOR (:659): after
                                   the repetition penalty, before temperature / top-k / top-p, also in greedy mode;
                                   outside (0,1) = off (0 in zero-initialised structs) */
  const int64_t* seq_indices;   /* dev int64 [n_seq] or NULL: global index of every local sequence, for shards that are not
                                   contiguous in the global order (length-balanced utterance shards); overrides seq_index_base */
} TtsArSampling;

typedef struct {
  int32_t struct_size;
  int32_t steps_run;            /* columns of out_codes that are valid */
  int32_t all_finished;
  int32_t watchdog;             /* != 0: a device-side wait timed out, results invalid */
  int64_t kernel_launches;      /* launches of this library's kernels since create */
  double last_algorithmic_bytes;/* SURVEY.md 8(d) bytes of the last generate/step call */
  float last_device_ms;         /* CUDA-event time of the last prefill+generate on its stream */
  float last_prefill_ms;
  int32_t kv_pages;             /* pages in the KV pool (32 positions x 16 heads x K,V per layer each) */
  int32_t ctas;                 /* CTAs of the persistent kernel (== SMs) */
} TtsArStats;

int tts_ar_abi_version(void);

/* Allocates device state (KV page pool, activations, packed-weight arena). SYNC. */
int tts_ar_create(const TtsArConfig* cfg, TtsArHandle** out);
void tts_ar_destroy(TtsArHandle* h);
const char* tts_ar_last_error(TtsArHandle* h);   /* h may be NULL: error of the last failed create */

/* Snapshot + repack the module's weights (replaces the state_dict the reference model owns;
 * call again after load_state_dict / LoRA merge, reference lora.py:197-205). */
int tts_ar_load_weights(TtsArHandle* h, const TtsArWeights* w, void* stream);

/* Prefix build + prompt pass.  Replaces inference_speech lines unified_voice.py:639-650 and the
 * prefill branch of GPT2InferenceModel.forward (:203-211) for `n_utt` utterances, each expanded to
 * `n_cand` candidates (num_return_sequences; stream_generator.py:618-623).
 *   latent   dev float32 [n_utt, 1024]
 *   text     dev int32   [n_utt, text_stride]  (un-padded ids; start/stop are added here)
 *   text_len host int32  [n_utt]               (each utterance runs at its true length) */
int tts_ar_prefill(TtsArHandle* h, const float* latent, const int32_t* text, int32_t text_stride,
                   const int32_t* text_len, int32_t n_utt, int32_t n_cand, void* stream);

/* The sample loop (stream_generator.py:1106-1190 / HF _sample): up to `max_new` tokens per sequence.
 *   out_codes   dev int64 [n_seq, max_new]   finished rows are padded with stop_mel_token (Q7)
 *   out_latents dev float32 [n_seq, max_new, 1024] or NULL: final_norm(hidden) per step (:1172)
 *   forced      dev int32 [n_seq, max_new] or NULL: teacher forcing (parity hook)
 *   out_choice  dev int32 [n_seq, max_new] or NULL: the sampler's own pick per step */
int tts_ar_generate(TtsArHandle* h, const TtsArSampling* s, int32_t max_new, int64_t* out_codes,
                    float* out_latents, const int32_t* forced, int32_t* out_choice, void* stream);

/* Streaming variant for get_generator (unified_voice.py:670-679): `n_steps` more tokens, state kept.
 * Same buffers as above with max_new := n_steps. */
int tts_ar_step(TtsArHandle* h, const TtsArSampling* s, int32_t n_steps, int64_t* out_codes,
                float* out_latents, const int32_t* forced, int32_t* out_choice, void* stream);

/* Streaming hand-off (reference get_generator -> sample_stream yields (next_tokens, final_norm(hidden)) per
 * step, stream_generator.py:1172, consumed by the HiFiGAN branch of inference.py:266-329): ONE persistent launch runs the whole
 * sample loop on `stream` and publishes every step through a flag ring in mapped pinned host memory; the host polls the flag and
 * then reads column `step` of the DEVICE buffers -- no per-chunk launches, allocations or stream synchronisation.
 *   tts_ar_stream_begin  after tts_ar_prefill, <= 64 sequences; buffers as tts_ar_generate (out_latents required).  ASYNC.
 *   tts_ar_stream_poll   waits (busy-poll, timeout_us < 0: forever) until (token, latent) of `step` of EVERY sequence are written:
 *                        1 = ready, 0 = time-out, 2 = the loop ended before this step (all sequences finished / max_new reached).
 *   tts_ar_stream_end    waits for the launch and reports like tts_ar_sync_result.  SYNC.
 * Pass a stream of your own (not the one the consumer of the tokens enqueues work on): the launch occupies it until the loop ends. */
int tts_ar_stream_begin(TtsArHandle* h, const TtsArSampling* s, int32_t max_new, int64_t* out_codes, float* out_latents, void* stream);
int tts_ar_stream_poll(TtsArHandle* h, int32_t step, int32_t timeout_us);
int tts_ar_stream_end(TtsArHandle* h, TtsArStats* out, void* stream);

/* Parity hook: the logits the NEXT sample would use.  out dev float32 [rows, n_mel_codes];
 * rows = n_utt right after prefill, n_seq after any step.  *rows_out is written on return. */
int tts_ar_logits_debug(TtsArHandle* h, float* out, int32_t* rows_out, void* stream);

/* UnifiedVoice.forward(..., return_latent=True, clip_inputs=False) (unified_voice.py:544-600).
 *   latent dev float32 [n, 1024]; text dev int32 [n, text_stride] + host text_len[n];
 *   codes dev int32 [n, codes_stride] + host codes_len[n];
 *   out dev float32 [n, codes_stride, 1024] (rows >= codes_len[i] are left untouched) */
int tts_ar_forward_latents(TtsArHandle* h, const float* latent, const int32_t* text, int32_t text_stride,
                           const int32_t* text_len, const int32_t* codes, int32_t codes_stride,
                           const int32_t* codes_len, int32_t n, float* out, void* stream);

/* Whole inference_speech call on HOST buffers: H2D of latent/text, prefill, generate, D2H of the
 * codes, all on `stream` (the caller's, like every other entry point), then waits for it.  SYNC.
 * This is the end-to-end entry bench.py times. */
int tts_ar_inference_speech_host(TtsArHandle* h, const float* latent_host, const int32_t* text_host,
                                 int32_t text_stride, const int32_t* text_len, int32_t n_utt,
                                 int32_t n_cand, const TtsArSampling* s, int32_t max_new,
                                 int64_t* out_codes_host, int32_t* steps_run, void* stream);

/* Epilogue between the sampler and its callers, in place on device codes [rows][stride] (int64):
 * reference tortoise_tts/inference.py:353-366 (and train.py:135-147): in a row that contains the stop token, it and
 * everything after it become `calm_token` (83) and the row's last three codes become 45, 45, 248; a row without a
 * stop token is left alone and reported as first_stop = -1 (the reference raises on it: `.min()` of an empty index
 * tensor, inference.py:355 -- the caller decides).  trim_len[r] = first index at which more than 8 consecutive calm
 * tokens have been seen in the fixed row, `len` if never (inference.py:381-389 cuts the latents of every row at
 * trim_len[0]).  first_stop / trim_len: device int32[rows], nullable.  ASYNC on `stream`. */
int tts_ar_fixup_codes(TtsArHandle* h, int64_t* codes, int32_t rows, int32_t len, int32_t stride, int32_t calm_token,
                       int32_t* first_stop, int32_t* trim_len, void* stream);

/* Parity hook: copies `bytes` of an internal device buffer to dst (dev).  name: "x" (residual stream, fp32
 * [rows,1024]), "q", "logits", "cur_latent", "kv" (page pool, bf16), "ctx", "h" (bf16 fragment order). */
int tts_ar_debug_copy(TtsArHandle* h, const char* name, void* dst, uint64_t bytes, void* stream);

/* Parity hook for the tcgen05 / TMEM GEMM of the dense passes alone (csrc/tc_gemm.cuh):
 * out[M][N] (dev fp32) = a[M][K] . wt[N][K]^T + bias[N]; a, wt dev bf16 row-major, K % 64 == 0, 16-byte aligned;
 * bias nullable; bn = 128 or 256 selects the tile width.  ASYNC on `stream`. */
int tts_ar_debug_gemm(TtsArHandle* h, const void* a_bf16, const void* wt_bf16, const float* bias, float* out,
                      int32_t M, int32_t N, int32_t K, int32_t bn, void* stream);

/* Allocation-free GEMM.  No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  No handle:
 * out[M][N] (dev fp32) = bf16(a)[M][K] . wt[N][K]^T + bias[N] on the calling thread's CURRENT device, with the tcgen05
 * kernel of tts_ar_debug_gemm; `a` is dev fp32 row-major (rounded to bf16, nearest-even, as the dense passes round their
 * LayerNorm output), wt dev bf16 row-major, both 16-byte aligned, K % 64 == 0, bias nullable, bn = 128 or 256.
 * The bf16 copy of `a` and the kernel's watchdog word are staged in a static __device__ array of 4 MiB + 256 B that every
 * GPU loading the library holds until exit: M * K <= 2,097,152 (e.g. 2048 x 1024, 512 x 4096), else TTS_AR_ERR_CAPACITY
 * (text via tts_ar_last_error(NULL)).  Calls no cudaMalloc / cudaFree / allocating routine.  All calls share the array:
 * issue them on ONE stream -- calls that overlap on the GPU race on the staged operand.
 *   status_dev  dev int32 or NULL: receives 0, or 1 if a device-side wait timed out (out is then invalid).  ASYNC. */
int tts_ar_gemm_f32_noalloc(const float* a_f32, const void* wt_bf16, const float* bias, float* out, int32_t M, int32_t N,
                            int32_t K, int32_t bn, int32_t* status_dev, void* stream);

/* Explicit per-handle options (all default 0 unless noted); returns TTS_AR_ERR_BAD_ARG for unknown names/values.
 *   "decode_kernel"      0 head-cluster kernels whenever their clusters are co-resident, 1 the interpreter kernel
 *   "dense_kernel"       0 tcgen05 pipeline for the prompt pass / latent re-forward, 1 the interpreter kernel
 *   "attn_tiles"         prompt-shaped attention: 0 auto, 1 one unit per (row, head), 2 always 16-row mma tiles
 *   "cooperative_launch" 1 (default) cooperative cluster launch, 0 plain cluster launch (profilers)
 *   "prof_step"          decode step whose per-stage timestamps are recorded (scripts/stage_prof.py); -1 off */
int tts_ar_set_option(TtsArHandle* h, const char* name, int32_t value);

/* ---- conditioning-latent encoder ---------------------------------------------------------------------
 * Replaces UnifiedVoice.get_conditioning (reference tortoise_tts/models/unified_voice.py:535-542, called once per voice by
 * tortoise_tts/emb/mel.py:97): ConditioningEncoder (:269-293) = 1x1 conv 80 -> 1024, then 6 AttentionBlocks
 * (tortoise_tts/models/arch_utils.py:136-190: GroupNorm32(32 groups), qkv 1x1 conv, QKVAttentionLegacy :59-95 with 16 heads,
 * proj_out 1x1 conv, residual), column 0 of the result, mean over the clips.  The 1x1 convolutions of the blocks run on the
 * tcgen05 GEMM (bf16 operands, fp32 accumulation); the init conv, GroupNorm statistics, softmax and the residual are fp32.
 * Weights are `conditioning_encoder.*` of the same state_dict, so the entry points live on the AR handle and use its
 * activation buffers: frames <= the handle's row capacity (>= max_text_len + max_new_tokens + 5), and a call invalidates a
 * pending tts_ar_prefill (prefill again before generating -- the reference computes the latent before the AR call too). */
typedef struct {
  const void *norm_w, *norm_b;          /* attn.{i}.norm      [1024] */
  const void *qkv_w, *qkv_b;            /* attn.{i}.qkv       [3072,1024,1], [3072]; channel = head * 192 + {q,k,v} * 64 + c */
  const void *proj_w, *proj_b;          /* attn.{i}.proj_out  [1024,1024,1], [1024] */
} TtsArCondBlockWeights;

typedef struct {
  int32_t struct_size;
  int32_t dtype;                        /* 0 = float32, 1 = bfloat16 (all tensors, dev pointers) */
  int32_t n_blocks;                     /* 6 */
  int32_t spec_dim;                     /* 80 mel bins */
  const void *init_w, *init_b;          /* init [1024,spec_dim,1], [1024] */
  const TtsArCondBlockWeights* blocks;  /* host array of n_blocks entries */
} TtsArCondWeights;

int tts_ar_load_conditioning_weights(TtsArHandle* h, const TtsArCondWeights* w, void* stream);
/* The mel-spectrogram front end of the same row: TorchMelSpectrogram.forward (reference tortoise_tts/models/arch_utils.py:361-395,
 * called by tortoise_tts/emb/mel.py:50-65) = torchaudio MelSpectrogram(n_fft 1024, hop 256, Hann, power 2, centre / reflect padding,
 * 80 HTK-scale slaney-normalised filters over 0..8000 Hz at 22050 Hz), log(clamp(., 1e-5)), / mel_norms.  One CTA per frame: a
 * 1024-point DFT against a sine / cosine table in shared memory with fp64 accumulation, the filter bank and the logarithm fused.
 *   wav    dev float32 [n_wav, n_samples] (n_samples > 512);  mel_norms dev float32 [80] or NULL
 *   mel    dev float32 [n_wav, 80, 1 + n_samples / 256]
 * ASYNC on `stream`; the first call on a handle allocates and fills the 0.2 MB of tables (SYNC once). */
int tts_ar_mel_spectrogram(TtsArHandle* h, const float* wav, int32_t n_wav, int32_t n_samples, const float* mel_norms, float* mel,
                           void* stream);
/*   mel     dev float32 [n_clips, spec_dim, frames]  (speech_conditioning_input[b] of the reference)
 *   latent  dev float32 [1024]: mean over the clips of column 0 of the encoder output.  ASYNC on `stream`. */
int tts_ar_get_conditioning(TtsArHandle* h, const float* mel, int32_t n_clips, int32_t frames, float* latent, void* stream);

/* ---- CLVP candidate scoring -------------------------------------------------------------------------
 * Replaces `clvp(text_tokens.repeat(n, 1), codes, return_loss=False)` of the reference (tortoise_tts/inference.py:392-396):
 * CLVP.forward (tortoise_tts/models/clvp.py:100-133) in eval mode with use_xformers=True -- two pre-norm x-transformers
 * encoders (RMSNorm, no-bias q/k/v, rotary embedding on the first 32 dims of q, k and v, full softmax attention, GLU
 * feed-forward with exact GELU; xtransformers.py:337-346, 431-479, 583-731, 912-1016), final LayerNorm, mean over positions,
 * to_*_latent, L2 normalisation, sim = <text, speech> * exp(temperature).  The caller keeps `torch.topk(scores, k)`.
 * The GEMMs run on the tcgen05 kernel of the dense passes (csrc/tc_gemm.cuh) in bf16 with fp32 accumulation; the residual
 * stream, norms, softmax and the latents are fp32.  Own handle: CLVP is a separate module with its own weights. */
typedef struct TtsClvpHandle TtsClvpHandle;

typedef struct {
  int32_t struct_size;
  int32_t device;
  int32_t dim;                  /* 768: dim_text == dim_speech == dim_latent (clvp.py:31-33) */
  int32_t heads;                /* 12 heads x 64 */
  int32_t depth;                /* 20: text_enc_depth == speech_enc_depth */
  int32_t n_text_tokens;        /* 256 */
  int32_t n_speech_tokens;      /* 8192 */
  int32_t max_seqs;             /* capacity: candidates per call */
  int32_t max_text_len;         /* capacity: text tokens per row */
  int32_t max_speech_len;       /* capacity: codes per candidate */
} TtsClvpConfig;

/* One (attention, feed-forward) pair of an encoder: reference keys
 * {text,speech}_transformer.transformer.attn_layers.layers.{2i}.* and .{2i+1}.*; nn.Linear weights are [out, in]. */
typedef struct {
  const void* attn_norm_g;                   /* layers.{2i}.0.0.g          [768] */
  const void *to_q_w, *to_k_w, *to_v_w;      /* layers.{2i}.1.wrap.to_*    [768,768], no bias */
  const void *to_out_w, *to_out_b;           /* layers.{2i}.1.wrap.to_out  [768,768], [768] */
  const void* ff_norm_g;                     /* layers.{2i+1}.0.0.g        [768] */
  const void *ff_proj_w, *ff_proj_b;         /* layers.{2i+1}.1.wrap.net.0.proj  [3072,768], [3072] (value rows, then gate rows) */
  const void *ff_out_w, *ff_out_b;           /* layers.{2i+1}.1.wrap.net.3       [768,1536], [768] */
} TtsClvpLayerWeights;

typedef struct {
  const void* emb;                           /* {text,speech}_emb.weight [n_tokens, 768] */
  const TtsClvpLayerWeights* layers;         /* host array of `depth` entries (dev pointers inside) */
  const void *norm_w, *norm_b;               /* transformer.norm (final LayerNorm) */
  const void* to_latent_w;                   /* to_{text,speech}_latent.weight [768,768], no bias */
} TtsClvpEncoderWeights;

typedef struct {
  int32_t struct_size;
  int32_t dtype;                             /* 0 = float32, 1 = bfloat16 (all tensors, dev pointers) */
  int32_t depth;
  TtsClvpEncoderWeights text, speech;
  const void* temperature;                   /* scalar parameter (sim is multiplied by exp of it) */
} TtsClvpWeights;

int tts_ar_clvp_create(const TtsClvpConfig* cfg, TtsClvpHandle** out);     /* allocates; SYNC */
void tts_ar_clvp_destroy(TtsClvpHandle* h);
const char* tts_ar_clvp_last_error(TtsClvpHandle* h);                      /* h may be NULL: error of the last failed create */
int tts_ar_clvp_load_weights(TtsClvpHandle* h, const TtsClvpWeights* w, void* stream);
/* scores[i] = sim(text row (n_text == 1 ? 0 : i), codes row i), i < n.
 *   text   dev int32 [n_text, text_stride], n_text == 1 (one text for every candidate) or n_text == n, text_len tokens each
 *   codes  dev int64 [n, codes_stride] (the layout tts_ar_generate / tts_ar_fixup_codes leave), codes_len codes each
 *   scores dev float32 [n];  speech_latents dev float32 [n, 768] or NULL: the L2-normalised speech latents (parity hook)
 * ASYNC on `stream`; a token id outside its table is reported by tts_ar_clvp_sync (the reference's nn.Embedding raises). */
int tts_ar_clvp_score(TtsClvpHandle* h, const int32_t* text, int32_t n_text, int32_t text_len, int32_t text_stride,
                      const int64_t* codes, int32_t n, int32_t codes_len, int32_t codes_stride, float* scores,
                      float* speech_latents, void* stream);
/* indices[r] (dev int64 [k]) = the candidate with the r-th largest score: `torch.topk(scores, k=candidates).indices` of
 * reference inference.py:395 without leaving the device (ties: lower index first; scores must not be NaN).  ASYNC. */
int tts_ar_clvp_topk(TtsClvpHandle* h, const float* scores, int32_t n, int32_t k, int64_t* indices, void* stream);
/* Waits for `stream`; TTS_AR_ERR_BAD_ARG after an out-of-range token id, TTS_AR_ERR_WATCHDOG after a device-side time-out
 * (each reported once and cleared).  *launches (nullable) = this library's kernel launches on the handle so far.  SYNC. */
int tts_ar_clvp_sync(TtsClvpHandle* h, int64_t* launches, void* stream);

/* Waits for `stream`, then reports the last call's outcome (TTS_AR_ERR_WATCHDOG after a device-side time-out,
 * TTS_AR_ERR_BAD_ARG when a token id was outside its embedding table -- the reference's nn.Embedding raises).
 * Either condition is reported once and cleared. SYNC. */
int tts_ar_sync_result(TtsArHandle* h, TtsArStats* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TTS_AR_H_ */
