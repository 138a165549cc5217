// tts_ar_host.h -- host-side state behind the opaque TtsArHandle (tts_ar.cu, tts_ar_run.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <string>
#include <vector>
#include "../../include/tts_ar.h"
#include "ar_launch.h"

// K-major bf16 weight image + TMA tensor maps of the dense (tcgen05) passes: dense_pass.cu
struct DenseW {
  bool ok = false;
  __nv_bfloat16* wt = nullptr;        // per layer: c_attn^T [3072][1024] | attn.c_proj^T [1024][1024] | c_fc^T [4096][1024] | mlp.c_proj^T [1024][4096]
  __nv_bfloat16* wt_head = nullptr;   // mel_head [n_mel_codes][1024]
  size_t per_layer = 0;
  struct Layer {
    __nv_bfloat16 *qkv, *proj, *fc, *proj2;
    CUtensorMap m_qkv[2], m_proj[2], m_fc[2], m_proj2[2];   // [0]: 128-row boxes, [1]: 256-row boxes
  };
  std::vector<Layer> layers;
  CUtensorMap m_head[2];
  CUtensorMap a_xn, a_ctx, a_h;       // activation buffers as A operands (row-major bf16 in this path)
};

// conditioning-latent encoder (cond_enc.cu): fp32 vectors + K-major bf16 images of the 1x1 convolutions of the attention blocks
struct CondW {
  bool loaded = false;
  int n_blocks = 0, spec_dim = 0;
  float *init_wt = nullptr, *init_b = nullptr;      // init conv, transposed [spec_dim][1024]; bias
  float* stats = nullptr;                           // GroupNorm (mean, rstd) per group [32][2]
  float *mel_win = nullptr, *mel_fb = nullptr;      // mel front end: Hann window [1024], filters [513][80]
  double* mel_tw = nullptr;                         // cos | sin of 2 pi i / 1024, [2][1024]
  struct Block {
    float *gn_w, *gn_b, *b_qkv, *b_proj;
    __nv_bfloat16 *qkv, *proj;                      // [3072][1024] rows reordered to q | k | v, [1024][1024]
    CUtensorMap m_qkv[2], m_proj[2];
  };
  std::vector<Block> blocks;
};

constexpr int kMaxGroups = 16;   // a batch runs as at most this many consecutive head-cluster launches of <= 64 sequences

struct TtsArHandle {
  TtsArConfig cfg{};
  int G = 0;                 // CTAs == SMs
  int smem_budget = 0;       // opt-in dynamic shared memory per CTA
  std::string err;
  std::vector<void*> allocs; // everything cudaMalloc'ed, freed in destroy

  // capacity
  int plen_max = 0;          // max prompt length P+1 = text + 4
  int rows_cap = 0;          // activation rows
  int npp = 0, nqp = 0;      // pages per prompt / per private segment
  int n_pages = 0;
  int rf_rows_cap = 0;       // rows per latent re-forward sub-batch
  int logit_rows_cap = 0;
  int plan_cap_units = 0;
  int seen_words = 0;
  int v_tiles = 0, v_stride = 0;

  // weights (device)
  bool weights_loaded = false;
  std::vector<tts::LayerDev> layers_host;
  tts::LayerDev* d_layers = nullptr;
  uint4* d_warena = nullptr;
  long long w_layer_u4 = 0;         // uint4 per layer in the packed-weight arena
  int fp_layer = 0;                 // floats per layer in the fp32 parameter arena
  float* d_fparams = nullptr;       // LN gains/biases, linear biases (fp32)
  uint4* d_w_head = nullptr;
  float *d_b_head = nullptr, *d_lnf_g = nullptr, *d_lnf_b = nullptr, *d_fn_g = nullptr, *d_fn_b = nullptr;
  float *d_mel_emb = nullptr, *d_mel_pos = nullptr, *d_text_emb = nullptr, *d_text_pos = nullptr;

  DenseW dense;
  CondW cond;

  // activations / cache (device)
  long long* d_y = nullptr;            // cluster decode kernels: attention-block counting accumulator [64][1024]
  uint2* d_xt = nullptr;               // cluster decode kernels: tagged residual stream between blocks [64][1024]
  bool use_cluster = false;         // 16 clusters of 8 CTAs can be co-resident (checked at create)
  float *d_x = nullptr, *d_q = nullptr, *d_logits = nullptr, *d_cur_latent = nullptr, *d_part = nullptr;
  __nv_bfloat16 *d_ctx = nullptr, *d_h = nullptr, *d_kv = nullptr, *d_xn = nullptr;
  int *d_prefix_pt = nullptr, *d_priv_pt = nullptr, *d_utt_plen = nullptr;
  int *d_pf_row_utt = nullptr, *d_pf_row_pos = nullptr, *d_head_rows = nullptr, *d_latent_dst = nullptr;
  int *d_seq_utt = nullptr, *d_seq_pos = nullptr;
  int *d_attn_cnt = nullptr, *d_plan = nullptr;
  int *d_finished = nullptr, *d_gen_count = nullptr, *d_n_active = nullptr;
  uint32_t* d_seen = nullptr;
  int *d_text_len = nullptr, *d_codes_len = nullptr;
  unsigned int* d_bar = nullptr;
  int *d_watchdog = nullptr, *d_steps_done = nullptr;
  unsigned long long* d_prof = nullptr;
  int prof_step = -1;
  // staging for the host-buffer entry point (pinned host + device)
  float* d_in_latent = nullptr;
  int* d_in_text = nullptr;
  long long* d_out_codes = nullptr;
  int n_groups = 1;              // head-cluster launches (groups of <= 64 sequences) the last run_steps used; 0: interpreter
  int *h_pinned_i = nullptr;     // pinned int scratch: [0]=steps_done [1]=watchdog
  long long* h_pinned_codes = nullptr;
  float* h_pinned_latent = nullptr;
  int* h_pinned_text = nullptr;

  // pinned staging for the small host -> device index uploads of a call (tts_ar_run.cu Stager)
  uint8_t* h_staging = nullptr;
  size_t staging_bytes = 0;
  cudaEvent_t ev_staging = nullptr;
  bool staging_busy = false;
  // streaming hand-off: mapped pinned flags (host pointer / device alias), tts_ar_stream_begin
  int* h_stream_flags = nullptr;
  int* d_stream_flags = nullptr;
  int stream_cap = 0, stream_steps = 0;
  bool streaming = false;
  // explicit options (tts_ar_set_option)
  int opt_decode_interp = 0, opt_dense_interp = 0, opt_attn_tiles = 0, opt_no_coop = 0;

  // per-call state
  bool prefilled = false;
  int n_utt = 0, n_cand = 0, n_seqs = 0;
  std::vector<int> plen;        // per utterance P+1
  int tokens_done = 0;          // tokens sampled so far for the current prompt
  bool first_step_pending = false;
  int dec_n_instr = 0, dec_mt_max = 1, pf_n_instr = 0, pt_rows = 0;
  long long launches = 0;
  double last_bytes = 0.0;
  int last_model_passes = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
  bool timing_valid = false;
};

namespace tts {
int fail(TtsArHandle* h, int code, const std::string& msg);
int cuda_fail(TtsArHandle* h, cudaError_t e, const char* what);
void fill_common_params(TtsArHandle* h, Params& P);
bool decode_cooperative(TtsArHandle* h);
// dense (tcgen05) passes: prompt pass and latent re-forward (dense_pass.cu)
int dense_create(TtsArHandle* h);
void dense_pack_layer(TtsArHandle* h, int l, const TtsArLayerWeights& lw, int dtype, cudaStream_t st);
void dense_pack_head(TtsArHandle* h, const void* mel_head_w, int dtype, cudaStream_t st);
int run_dense_pass(TtsArHandle* h, const Params& P, cudaStream_t st);
}  // namespace tts

#define TTS_CUDA(h, call)                                          \
  do {                                                             \
    cudaError_t e__ = (call);                                      \
    if (e__ != cudaSuccess) return tts::cuda_fail((h), e__, #call); \
  } while (0)
