// ar_launch.h -- host-callable launchers implemented in ar_kernels.cu.
#pragma once
#include <cuda_runtime.h>
#include "ar_params.h"

namespace tts {

struct EmbedArgs {
  float* x;
  const int* row_utt;      // [rows]
  const int* row_pos_in;   // [rows] position inside the utterance's sequence
  const int* text_len;     // [utts] (device)
  const int* text;         // [utts][text_stride]
  int text_stride;
  const int* codes;        // nullable [utts][codes_stride] (latent re-forward)
  const int* codes_len;    // [utts] (device)
  int codes_stride;
  const float* latent;     // [utts][1024]
  const float *text_emb, *text_pos, *mel_emb, *mel_pos;
  int start_text, stop_text, start_mel, stop_mel, n_text_rows, n_mel_codes;
  int* err_flag;           // set to 1 when a token id is outside its embedding table (the reference raises IndexError)
};

void launch_pack_weight(const void* src, int dtype, uint4* dst, int K, int N, int n_tiles, int transposed, cudaStream_t st);
void launch_to_f32(const void* src, int dtype, float* dst, size_t n, size_t n_pad, cudaStream_t st);
void launch_embed_rows(const EmbedArgs& a, int n_rows, cudaStream_t st);
void launch_fixup_codes(long long* codes, int rows, int len, int stride, int stop, int calm, int* first_stop, int* trim,
                        cudaStream_t st);
void launch_reset_sampler(int n_seqs, int seen_words, uint32_t* seen, int* finished, int* gen_count, int* row_pos,
                          int* n_active, int group_size, int start_mel, int* attn_cnt, int n_cnt, cudaStream_t st);
void launch_fill_i64(long long* p, size_t n, long long v, cudaStream_t st);
cudaError_t launch_program(const Params& P, int smem_bytes, int device, cudaStream_t st);
int program_max_ctas_per_sm(int smem_bytes);
// head-cluster decode kernel (ar_decode.cu)
void launch_pack_projk(const void* src, int dtype, uint4* dst, cudaStream_t st);
// mode = 16-row m-tiles per launch (1, 2, 4: <= 16, 32, 64 sequences)
int decode_mode_for(int n_rows);
int decode_mt_smem(int mode);
cudaError_t launch_decode(const Params& P, int mode, int smem_bytes, int device, bool cooperative, cudaStream_t st);
int decode_max_clusters(int mode, int smem_bytes, int device);

}  // namespace tts
