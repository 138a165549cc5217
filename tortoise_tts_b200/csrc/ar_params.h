// ar_params.h -- structures shared between the host runtime (tts_ar.cu) and the device program
// interpreter (ar_kernels.cu).  One persistent cooperative kernel executes a small "program" of
// instructions (one per fused stage of the GPT-2 block) with grid-wide barriers between dependent
// stages; the same interpreter runs the prompt pass (rows = prompt positions) and the decode loop
// (rows = live sequences).
#pragma once
#include <stdint.h>
#include <cuda_bf16.h>
#include "layout.h"

namespace tts {

enum Op : int {
  OP_LN_QKV = 0,   // x -> ln_1 -> c_attn (+bias) -> q scratch (pre-scaled 1/8) and K/V pages
  OP_ATTN = 1,     // split-KV attention over the paged cache -> ctx (A-fragment layout)
  OP_PROJ = 2,     // ctx -> attn.c_proj (+bias) -> x += .
  OP_LN_FC = 3,    // x -> ln_2 -> c_fc (+bias) -> gelu_new -> h (A-fragment layout)
  OP_PROJ2 = 4,    // h -> mlp.c_proj (+bias) -> x += .
  OP_HEAD = 5,     // x -> ln_f -> final_norm -> latent ; -> mel_head (+bias) -> logits
  OP_SAMPLE = 6,   // logits -> processors/warpers -> token ; next-step embedding
  OP_LATENT = 7    // x -> ln_f -> final_norm -> out rows (latent re-forward; no GEMM)
};

enum InstrFlags : int {
  F_BARRIER_BEFORE = 1,  // grid barrier before executing
  F_CHECK_EXIT = 2,      // after the barrier: leave the launch if no sequence is unfinished
  F_ROW_MAP = 4,         // OP_HEAD / OP_LATENT: read x rows through head_rows[]
  F_PLAN = 8,            // OP_LN_QKV of layer 0 in decode: last CTA also builds the attention split plan
  F_FIXED_SRC = 32,      // LayerNorm source is the fixed-point accumulator P.y
  F_ROW_RANGE = 16       // epilogue writes rows [pad0, pad1) only (cluster decode kernel: the other half is the sibling's)
};

struct Instr {
  int op;
  int layer;
  int row0;     // first row (multiple of 16)
  int nrows;    // rows in this instruction (<= 128 for GEMM ops, ProgIter::row_group; all rows for ATTN/SAMPLE/LATENT)
  int flags;
  int rot;      // tile-schedule rotation (layout.h tile_range)
  int pad0, pad1;
};

struct LayerDev {
  const uint4 *w_qkv, *w_proj, *w_fc, *w_proj2;            // packed chunks (layout.h)
  const uint4* w_projk;                                    // attn.c_proj again, split-K chunks (layout.h projk_elem_offset)
  const float *ln1_g, *ln1_b, *ln2_g, *ln2_b;
  const float *b_qkv, *b_proj, *b_fc, *b_proj2;
};

struct Params {
  // geometry
  int G;                 // CTAs in the grid (== SMs)
  int n_layers;
  int V;                 // mel codes (8194)
  int v_tiles;           // ceil(V / 8)
  int v_stride;          // logits row stride (floats)
  int n_slots;           // ring slots of kChunkBytes
  int mt_max;            // largest M-tile count any instruction uses (sizes the A staging area)
  int decode_mode;       // 1: rows are sequences (private pages); 0: rows are prompt positions
  // weights
  const LayerDev* layers;
  // the same pointers as arithmetic (no dependent pointer load on the critical path): layer l's packed weights start at
  // warena + l * w_layer_u4 (c_attn | attn.c_proj | c_fc | mlp.c_proj | split-K attn.c_proj), its fp32 parameters at
  // fparams + l * fp_layer (ln_1 g,b | ln_2 g,b | b_attn | b_proj | b_fc | b_proj2)   (tts_ar.cu create_impl)
  const uint4* warena;
  const float* fparams;
  long long w_layer_u4;
  int fp_layer;
  const uint4* w_head;
  const float *b_head, *lnf_g, *lnf_b, *fn_g, *fn_b;
  const float *mel_emb, *mel_pos;
  // activations (row-indexed)
  float* x;              // [rows][1024] residual stream, fp32
  long long* y;          // cluster decode kernel: [16][1024] residual stream between the attention and MLP blocks,
                         //   accumulated by the 16 heads with integer atomics as counting fixed-point words (dev_utils.cuh
                         //   cnt_*: order-independent => deterministic, and self-validating); nullptr in the interpreter kernel
  uint2* xt;             // cluster decode kernels: [rows][1024] (fp32 bits, tag) -- the residual stream between blocks, self-validating
  float* q;              // [rows][1024] query, pre-scaled by 1/sqrt(64)
  __nv_bfloat16* ctx;    // A-fragment layout, K = 1024
  __nv_bfloat16* hbuf;   // A-fragment layout, K = 4096
  __nv_bfloat16* xn;     // A-fragment layout, K = 1024: LayerNorm'd rows of the current ln_1 / ln_2 stage when more than 16
                         //   rows run through the interpreter (normalised ONCE, grid-wide, instead of by every CTA); nullable
  float* logits;         // [logit rows][v_stride]
  float* cur_latent;     // [logit rows][1024] final_norm(ln_f(x)) of the rows fed to the head
  const float* logits_pf;      // prompt-pass logits / latents, row = utterance: what the first SAMPLE after a prefill reads
  const float* cur_latent_pf;  //   (== logits / cur_latent unless the launch covers a 16-sequence group of a larger batch)
  // KV cache
  __nv_bfloat16* kv;     // page pool (layout.h kv_row_offset)
  int n_pages;
  const int* prefix_pt;  // [utts][kMaxPagesPerSeg]
  const int* priv_pt;    // [seqs][kMaxPagesPerSeg]
  const int* utt_plen;   // [utts] prompt length P+1
  int* row_utt;          // [rows] utterance of each row
  int* row_pos;          // [rows] position of the row's token in its sequence; -1 = inactive row
  const int* head_rows;  // [logit rows] row map for F_ROW_MAP
  // attention scratch
  float* part;           // [units][kPartStride]
  int* attn_cnt;         // [rows * 16] arrival counters (zero between uses)
  int* plan;             // [0]=mode(0 direct,1 split) [1]=n_units [2]=tp ; [4 + r]=unit_off[r] (rows+1 entries)
  int plan_cap_units;
  // sampler state (per sequence)
  int* finished;
  int* gen_count;        // tokens generated so far
  uint32_t* seen;        // [seqs][seen_words] repetition-penalty id set
  int seen_words;
  int* n_active;         // unfinished sequences
  const int* seq_utt;
  int first_step_from_prefill;  // 1: the first SAMPLE of this launch reads logits row seq_utt[s]
  int n_seqs;
  // sampler config
  int do_sample, top_k, ignore_eos;
  float temperature, top_p, rep_penalty;
  float typical_mass;            // TypicalLogitsWarper mass in (0,1); 0 = off
  unsigned long long seed;
  long long seq_index_base;
  const long long* seq_indices;   // nullable [seqs]: global sequence index per local sequence (overrides seq_index_base + seq)
  int stop_token;
  int max_pos_rows;      // rows of mel_pos (bounds guard)
  // outputs
  long long* out_codes;  // [seqs][out_stride], column = step index within this launch
  int out_stride;
  float* out_latents;    // nullable [seqs][out_stride][1024]
  const int* forced;     // nullable [seqs][out_stride]
  int* out_choice;       // nullable [seqs][out_stride]
  float* latent_out;     // OP_LATENT destination rows [n][1024] via latent_dst
  const int* latent_dst; // [rows] destination row index in latent_out or -1
  // streaming hand-off (tts_ar_stream_begin): mapped pinned host memory; flags[step] counts the sequences whose (token, latent) of
  // that step are in out_codes / out_latents, done = steps run + 1 once the launch has ended
  int* stream_flags;
  int* stream_done;
  // control
  unsigned int* bar_counter;
  int* watchdog;
  int* steps_done;
  // program shape (generated procedurally by ProgIter, identically in the producer and the consumers)
  int has_sample;        // decode: every step starts with OP_SAMPLE
  int n_rows;            // rows run through the layers (sequences in decode, prompt positions otherwise)
  int tail;              // 1: OP_HEAD over n_head_rows (F_ROW_MAP in prompt mode), 2: OP_LATENT over n_rows
  int n_head_rows;
  int n_steps;
  unsigned long long* prof;   // nullable: CTA 0 stage timestamps (ns) of step prof_step: [2*pc] before / [2*pc+1] after the barrier
  int prof_step;
  int attn_tiles_mode;   // interpreter prompt-mode attention (tts_ar_set_option "attn_tiles"): 0 auto, 1 per-(row, head) units, 2 always 16-row mma tiles
  int skip_last_model;   // generate(): the model is not run after the last sampled token
  int ctx_row_major;     // dense passes (dense_pass.cu): attention context is bf16 [rows][1024], not mma A-fragment order
};

constexpr int kU4PerChunkP = kChunkBytes / 16;
struct LayerPtr {          // layer-l views computed from Params (device side)
  const uint4 *w_qkv, *w_proj, *w_fc, *w_proj2, *w_projk;
  const float *ln1_g, *ln1_b, *ln2_g, *ln2_b, *b_qkv, *b_proj, *b_fc, *b_proj2;
};
TTS_HD const uint4* layer_w(const Params& P, int l) { return P.warena + (long long)l * P.w_layer_u4; }
TTS_HD const float* layer_f(const Params& P, int l) { return P.fparams + (long long)l * P.fp_layer; }
constexpr long long kOffProj = (long long)(kD3 / 8) * kU4PerChunkP;
constexpr long long kOffFc = kOffProj + (long long)(kD / 8) * kU4PerChunkP;
constexpr long long kOffProj2 = kOffFc + (long long)(kD4 / 8) * kU4PerChunkP;
constexpr long long kOffProjk = kOffProj2 + (long long)(kD / 8) * 4 * kU4PerChunkP;
constexpr int kFOffLn1g = 0, kFOffLn1b = kD, kFOffLn2g = 2 * kD, kFOffLn2b = 3 * kD, kFOffBqkv = 4 * kD,
              kFOffBproj = 4 * kD + kD3, kFOffBfc = 5 * kD + kD3, kFOffBproj2 = 5 * kD + kD3 + kD4;
static_assert(kFOffLn1b == kFOffLn1g + kD && kFOffLn2b == kFOffLn2g + kD, "ar_decode.cu stages (gain, bias) pairs as one block");

// dynamic shared memory layout of the interpreter kernel
constexpr int kSmemCtrlBytes = 1024;
constexpr int kSmemRedBytes = 33792;                     // >= 16 warps * 4 tiles * 512 B, >= V floats
TTS_HD int smem_a_bytes(int mt_max) { return mt_max * 16 * kARowBytes; }
TTS_HD int smem_total_bytes(int mt_max, int n_slots) {
  return kSmemCtrlBytes + smem_a_bytes(mt_max) + kSmemRedBytes + n_slots * kChunkBytes;
}
TTS_HD int smem_max_slots(int mt_max, int budget) {
  int s = (budget - kSmemCtrlBytes - smem_a_bytes(mt_max) - kSmemRedBytes) / kChunkBytes;
  return s > 12 ? 12 : s;
}

}  // namespace tts
