// ar_kernels.cu -- the persistent interpreter kernel (one CTA per SM, cooperative launch) and the small
// setup kernels (weight repacking, prompt embedding, sampler-state reset).
#include <algorithm>
#include "ar_gemm.cuh"
#include "ar_attn.cuh"
#include "ar_sample.cuh"
#include "ar_launch.h"

namespace tts {

// ---- the interpreter -------------------------------------------------------------------------------------
// Warps 0..15: consumers (compute, grid barriers).  Warp 16: producer (weight stream).  The producer is not
// gated by the grid barriers -- weights do not depend on activations -- so HBM keeps streaming the next
// stage's (and next layer's, next step's) weights while the consumers wait on a barrier or do attention.
__global__ void __launch_bounds__(kThreads, 1) ar_program_kernel(const __grid_constant__ Params P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const Smem sm = carve_smem(smem_raw, P.mt_max);
  Ctrl* c = sm.ctrl;
  if (threadIdx.x == 0) {
    for (int i = 0; i < P.n_slots; ++i) {
      mbar_init(&c->full[i], 1);
      mbar_init(&c->empty[i], kNW);
    }
    c->abort_flag = 0;
    c->ibuf[0] = 0;
    c->ibuf[1] = 0;
    c->stop_flag = 0;
    c->stop_ci = 0;
    fence_mbar_init();
  }
  __syncthreads();

  if (threadIdx.x >= kConsumerThreads) {
    producer_loop(P, c, sm.ring);
    return;
  }

  Pipe pipe;
  pipe.ci = 0; pipe.slot = 0; pipe.phase = 0; pipe.n_slots = P.n_slots;
  uint32_t epoch = 0;
  int steps_done = 0;
  bool alive = true;
  for (int step = 0; step < P.n_steps && alive; ++step) {
    ProgIter it;
    it.reset(P);
    Instr ins;
    for (int pc = 0; alive && it.next(P, ins); ++pc) {
      if (!instr_live(P, step, ins)) continue;
      const bool prof = P.prof && step == P.prof_step && blockIdx.x == 0 && threadIdx.x == 0;
      if (prof) P.prof[2 * pc] = globaltimer_ns();
      if (blockIdx.x == 0 && threadIdx.x == 0) c->ibuf[1] = (P.prof && step == P.prof_step) ? 1 : 0;
      prof_mark(P, c, 100 + ins.op);
      if (ins.flags & F_BARRIER_BEFORE) alive = grid_barrier(P, c, epoch);
      if (prof) P.prof[2 * pc + 1] = globaltimer_ns();
      prof_mark(P, c, 200 + ins.op);
      if (!alive) break;
      if ((ins.flags & F_CHECK_EXIT) && ld_cg_i(P.n_active) == 0) { alive = false; break; }
      switch (ins.op) {
        case OP_LN_QKV:
          if ((ins.flags & F_PLAN) && blockIdx.x == P.G - 1 && (threadIdx.x >> 5) == 0) build_attn_plan(P, P.n_seqs);
          // fallthrough
        case OP_PROJ:
        case OP_LN_FC:
        case OP_PROJ2:
        case OP_HEAD: {
          if ((ins.op == OP_LN_QKV || ins.op == OP_LN_FC) && ins.row0 == 0 && ln_split_active(P)) {
            const float* fp = layer_f(P, ins.layer);
            if (ins.op == OP_LN_QKV) ln_rows_to_global(P, ins, fp + kFOffLn1g, fp + kFOffLn1b);
            else ln_rows_to_global(P, ins, fp + kFOffLn2g, fp + kFOffLn2b);
            alive = grid_barrier(P, c, epoch);
            if (!alive) break;
          }
          const int mt = (ins.nrows + 15) >> 4;
          if (mt <= 1) gemm_stage<1>(P, ins, sm, pipe);
          else if (mt == 2) gemm_stage<2>(P, ins, sm, pipe);
          else if (mt <= 4) gemm_stage<4>(P, ins, sm, pipe);
          else gemm_stage<8>(P, ins, sm, pipe);
          break;
        }
        case OP_ATTN:
          attn_stage(P, ins, sm);
          break;
        case OP_SAMPLE:
          sample_stage(P, ins, sm, step);
          steps_done = step + 1;
          break;
        case OP_LATENT: {
          for (int rg = blockIdx.x; rg * 16 < ins.nrows; rg += P.G) {
            Instr sub = ins;
            sub.row0 = ins.row0 + rg * 16;
            sub.nrows = min(16, ins.nrows - rg * 16);
            ln_rows_to_smem<true>(P, sub, sm.a_rm, P.x, P.lnf_g, P.lnf_b, P.fn_g, P.fn_b, 1);
          }
          break;
        }
        default: break;
      }
      if (c->abort_flag) alive = false;
    }
  }
  consumer_sync();
  if (threadIdx.x == 0) {
    if (blockIdx.x == 0) *P.steps_done = steps_done;
    c->stop_ci = pipe.ci;
    __threadfence_block();
    c->stop_flag = 1;
  }
}

// ---- weight repacking: HF [K, N] (Conv1D, row-major) or [N, K] (nn.Linear) -> streamed chunk order ----------
template <typename T>
__global__ void pack_weight_kernel(const T* __restrict__ src, uint4* __restrict__ dst, int K, int N, int n_tiles,
                                   int transposed /* 1: src is [N, K] */) {
  // one thread per 32-bit word (two consecutive k) of the packed image
  const size_t total_words = (size_t)n_tiles * (K / kChunkK) * (kChunkBytes / 4);
  uint32_t* out = reinterpret_cast<uint32_t*>(dst);
  for (size_t wi = (size_t)blockIdx.x * blockDim.x + threadIdx.x; wi < total_words; wi += (size_t)gridDim.x * blockDim.x) {
    const size_t chunk = wi / (kChunkBytes / 4);
    const int in_chunk = (int)(wi % (kChunkBytes / 4));
    const int kq_n = K / kChunkK;
    const int tile = (int)(chunk / kq_n), kq = (int)(chunk % kq_n);
    // invert chunk_elem_offset for word `in_chunk` (half = 0)
    const int word = in_chunk & 3, lane = (in_chunk >> 2) & 31, s = in_chunk >> 7;
    const int odd = word >> 1, hi = word & 1;
    const int t = lane & 3, n = lane >> 2;
    const int kb = 2 * s + odd;
    const int k = kq * kChunkK + kb * 16 + hi * 8 + 2 * t;
    const int col = tile * 8 + n;
    float v0 = 0.f, v1 = 0.f;
    if (col < N) {
      if (transposed) {
        v0 = (float)src[(size_t)col * K + k];
        v1 = (float)src[(size_t)col * K + k + 1];
      } else {
        v0 = (float)src[(size_t)k * N + col];
        v1 = (float)src[(size_t)(k + 1) * N + col];
      }
    }
    out[wi] = pack_bf16x2(v0, v1);
  }
}

template <typename T>
__global__ void to_f32_kernel(const T* __restrict__ src, float* __restrict__ dst, size_t n, size_t n_pad) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pad; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = (i < n) ? (float)src[i] : 0.f;
}

void launch_pack_weight(const void* src, int dtype, uint4* dst, int K, int N, int n_tiles, int transposed, cudaStream_t st) {
  const size_t words = (size_t)n_tiles * (K / kChunkK) * (kChunkBytes / 4);
  const int blocks = (int)((words + 255) / 256 > 4096 ? 4096 : (words + 255) / 256);
  if (dtype == 0)
    pack_weight_kernel<float><<<blocks, 256, 0, st>>>((const float*)src, dst, K, N, n_tiles, transposed);
  else
    pack_weight_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>((const __nv_bfloat16*)src, dst, K, N, n_tiles, transposed);
}

void launch_to_f32(const void* src, int dtype, float* dst, size_t n, size_t n_pad, cudaStream_t st) {
  const int blocks = (int)((n_pad + 255) / 256 > 2048 ? 2048 : (n_pad + 255) / 256);
  if (dtype == 0)
    to_f32_kernel<float><<<blocks, 256, 0, st>>>((const float*)src, dst, n, n_pad);
  else
    to_f32_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>((const __nv_bfloat16*)src, dst, n, n_pad);
}

// ---- prompt embedding: rows of every utterance (reference unified_voice.py:639-645 + :203-211) ---------------
//   row 0: conditioning latent; rows 1..T+2: text_embedding([start, text.., stop]) + text_pos[0..T+1];
//   mode 0 (inference): last row = mel_embedding(start_mel) + mel_pos[0]
//   mode 1 (latent re-forward, :577-590): then mel_embedding([start, codes.., stop]) + mel_pos[0..M+1]
__global__ void embed_rows_kernel(EmbedArgs a) {
  const int row = blockIdx.x;
  const int u = a.row_utt[row];
  const int pos = a.row_pos_in[row];
  const int T = a.text_len[u];
  float* xp = a.x + (size_t)row * kD;
  const float* e0;
  const float* e1 = nullptr;
  if (pos == 0) {
    e0 = a.latent + (size_t)u * kD;
  } else if (pos <= T + 2) {
    const int j = pos - 1;
    int tok = (j == 0) ? a.start_text : (j == T + 1 ? a.stop_text : a.text[(size_t)u * a.text_stride + (j - 1)]);
    if ((tok < 0 || tok >= a.n_text_rows) && threadIdx.x == 0 && a.err_flag) *a.err_flag = 1;
    tok = min(max(tok, 0), a.n_text_rows - 1);
    e0 = a.text_emb + (size_t)tok * kD;
    e1 = a.text_pos + (size_t)j * kD;
  } else {
    const int j = pos - (T + 3);                       // mel position index
    int tok;
    if (j == 0) tok = a.start_mel;
    else if (a.codes && j <= a.codes_len[u]) tok = a.codes[(size_t)u * a.codes_stride + (j - 1)];
    else tok = a.stop_mel;
    if ((tok < 0 || tok >= a.n_mel_codes) && threadIdx.x == 0 && a.err_flag) *a.err_flag = 1;
    tok = min(max(tok, 0), a.n_mel_codes - 1);
    e0 = a.mel_emb + (size_t)tok * kD;
    e1 = a.mel_pos + (size_t)j * kD;
  }
  for (int i = threadIdx.x; i < kD; i += blockDim.x) xp[i] = e0[i] + (e1 ? e1[i] : 0.f);
}

void launch_embed_rows(const EmbedArgs& a, int n_rows, cudaStream_t st) { embed_rows_kernel<<<n_rows, 256, 0, st>>>(a); }

// ---- sampler state reset after a prompt pass (fake_inputs = [1]*P + [start_mel], unified_voice.py:647-649) ----
__global__ void reset_sampler_kernel(int n_seqs, int seen_words, uint32_t* seen, int* finished, int* gen_count,
                                     int* row_pos, int* n_active, int group_size, int start_mel, int* attn_cnt, int n_cnt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_seqs * seen_words) {
    const int wd = i % seen_words;
    uint32_t v = 0;
    if (wd == 0) v |= 1u << 1;                         // dummy prompt id 1
    if (wd == (start_mel >> 5)) v |= 1u << (start_mel & 31);
    seen[i] = v;
  }
  if (i < n_seqs) {
    finished[i] = 0;
    gen_count[i] = 0;
    row_pos[i] = -1;
  }
  if (i < n_cnt) attn_cnt[i] = 0;
  if (i == 0) n_active[0] = n_seqs;
  // groups of a batch larger than one head-cluster launch run as consecutive launches (tts_ar_run.cu run_steps)
  if (i * group_size < n_seqs) n_active[1 + i] = min(group_size, n_seqs - i * group_size);
}

void launch_reset_sampler(int n_seqs, int seen_words, uint32_t* seen, int* finished, int* gen_count, int* row_pos,
                          int* n_active, int group_size, int start_mel, int* attn_cnt, int n_cnt, cudaStream_t st) {
  int n = n_seqs * seen_words;
  if (n_cnt > n) n = n_cnt;
  reset_sampler_kernel<<<(n + 255) / 256, 256, 0, st>>>(n_seqs, seen_words, seen, finished, gen_count, row_pos, n_active,
                                                        group_size, start_mel, attn_cnt, n_cnt);
}

// ---- stop-token fix-up + calm-token trim of the generated codes (reference inference.py:353-366, 381-389; the same
// loop in train.py:135-147, 161-169): one CTA per row.
//   row has a stop token at first index s:  codes[s:] = calm (83); then the last three codes = 45, 45, 248
//   row has none: untouched, first_stop = -1 (the reference raises on such a row, quirk Q10: the adapter decides)
//   trim[r] = first index k at which more than 8 consecutive calm tokens have been seen (row length if never):
//   the reference cuts the latents of ALL rows at trim[0].
__global__ void fixup_codes_kernel(long long* __restrict__ codes, int len, int stride, int stop, int calm, int* first_stop,
                                   int* trim) {
  __shared__ int s_first, s_trim;
  long long* row = codes + (size_t)blockIdx.x * stride;
  if (threadIdx.x == 0) { s_first = len; s_trim = len; }
  __syncthreads();
  int mine = len;
  for (int k = threadIdx.x; k < len; k += blockDim.x)
    if (row[k] == (long long)stop) { mine = k; break; }          // ascending per thread: the first hit is its minimum
  if (mine < len) atomicMin(&s_first, mine);
  __syncthreads();
  const int s = s_first;
  if (s < len) {
    for (int k = s + threadIdx.x; k < len; k += blockDim.x) row[k] = calm;
    __syncthreads();
    if (threadIdx.x < 3 && len >= 3) row[len - 3 + threadIdx.x] = threadIdx.x == 2 ? 248 : 45;
    __syncthreads();
  }
  // more than 8 consecutive calm tokens ending at k  <=>  row[k-8 .. k] are all calm
  int t = len;
  for (int k = 8 + threadIdx.x; k < len; k += blockDim.x) {
    bool all = true;
#pragma unroll
    for (int j = 0; j < 9; ++j) all = all && row[k - j] == (long long)calm;
    if (all) { t = k; break; }
  }
  if (t < len) atomicMin(&s_trim, t);
  __syncthreads();
  if (threadIdx.x == 0) {
    if (first_stop) first_stop[blockIdx.x] = s < len ? s : -1;
    if (trim) trim[blockIdx.x] = s_trim;
  }
}

void launch_fixup_codes(long long* codes, int rows, int len, int stride, int stop, int calm, int* first_stop, int* trim,
                        cudaStream_t st) {
  fixup_codes_kernel<<<rows, 128, 0, st>>>(codes, len, stride, stop, calm, first_stop, trim);
}

__global__ void fill_i64_kernel(long long* p, size_t n, long long v) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

void launch_fill_i64(long long* p, size_t n, long long v, cudaStream_t st) {
  const int blocks = (int)std::min<size_t>((n + 255) / 256, 1024);
  if (n) fill_i64_kernel<<<blocks, 256, 0, st>>>(p, n, v);
}

cudaError_t launch_program(const Params& P, int smem_bytes, int device, cudaStream_t st) {
  static unsigned long long attr_set = 0;   // one bit per device ordinal: the opt-in is per device (function handle)
  if (!(attr_set >> (device & 63) & 1ull)) {
    cudaError_t e = cudaFuncSetAttribute(ar_program_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
    if (e != cudaSuccess) return e;
    attr_set |= 1ull << (device & 63);
  }
  void* args[] = {const_cast<Params*>(&P)};
  return cudaLaunchCooperativeKernel((const void*)ar_program_kernel, dim3(P.G), dim3(kThreads), args, (size_t)smem_bytes, st);
}

int program_max_ctas_per_sm(int smem_bytes) {
  int n = 0;
  cudaFuncSetAttribute(ar_program_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, ar_program_kernel, kThreads, (size_t)smem_bytes);
  return n;
}

}  // namespace tts
