// cond_enc.cu -- the conditioning-latent encoder: mel spectrogram of the voice clips -> the 1024-d
// latent the AR prompt starts with.
// Replaces: UnifiedVoice.get_conditioning (reference tortoise_tts/models/unified_voice.py:535-542) = ConditioningEncoder
// (:269-293: 1x1 conv, 6 x AttentionBlock, h[:, :, 0]) per clip, mean over the clips; AttentionBlock =
// x + proj_out(QKVAttentionLegacy(qkv(GroupNorm32(x)))) (tortoise_tts/models/arch_utils.py:24-44, 59-95, 136-190).
// Rows = mel frames (positions), columns = channels (the transpose of the reference's [C, T]):
//     cond_init_kernel (fp32 1x1 conv) -> per block: cond_gn_stats_kernel (32 groups x all frames) -> cond_gn_apply_kernel
//     (-> bf16 A operand) -> qkv GEMM (tcgen05, rows reordered to q | k | v at load) -> full_attn_kernel<1024> ->
//     proj_out GEMM (+bias, +residual) -> cond_take_kernel (row 0, running mean over the clips)
// Uses the AR handle's activation buffers (x, xn, ctx, h) and its A-operand tensor maps (dense_pass.cu).
#include <cmath>
#include <cstdio>
#include <string>
#include <vector>
#include "tc_launch.cuh"
#include "full_attn.cuh"
#include "tts_ar_host.h"

namespace tts {

constexpr int kCondGroups = 32;                  // normalization(1024) -> GroupNorm32(32, 1024) (arch_utils.py:29-44)
constexpr int kCondGroupCh = kD / kCondGroups;   // 32 channels per group

// x[t][c] = b[c] + sum_k Wt[k][c] * mel[k][t]   (nn.Conv1d(spec_dim, 1024, 1), unified_voice.py:279, 286)
__global__ void __launch_bounds__(256) cond_init_kernel(const float* __restrict__ mel, const float* __restrict__ wt, const float* __restrict__ b,
                                                        float* __restrict__ x, int spec_dim, int frames) {
  __shared__ float col[128];
  const int t = blockIdx.x;
  for (int k = threadIdx.x; k < spec_dim; k += 256) col[k] = mel[(size_t)k * frames + t];
  __syncthreads();
  float4 acc = *reinterpret_cast<const float4*>(b + 4 * threadIdx.x);
  for (int k = 0; k < spec_dim; ++k) {
    const float4 w = *reinterpret_cast<const float4*>(wt + (size_t)k * kD + 4 * threadIdx.x);
    const float m = col[k];
    acc.x += w.x * m; acc.y += w.y * m; acc.z += w.z * m; acc.w += w.w * m;
  }
  *reinterpret_cast<float4*>(x + (size_t)t * kD + 4 * threadIdx.x) = acc;
}

// GroupNorm statistics of group blockIdx.x over (32 channels x all frames): two passes, biased variance, eps 1e-5
__global__ void __launch_bounds__(256) cond_gn_stats_kernel(const float* __restrict__ x, float* __restrict__ stats, int frames) {
  __shared__ float red[8];
  __shared__ float s_mean;
  const int grp = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = threadIdx.x & 7, r0 = threadIdx.x >> 3;          // 8 threads x float4 = the group's 32 channels of one row
  const float* base = x + grp * kCondGroupCh + 4 * sub;
  const float n = (float)frames * kCondGroupCh;
  float s = 0.f;
  for (int r = r0; r < frames; r += 32) {
    const float4 v = *reinterpret_cast<const float4*>(base + (size_t)r * kD);
    s += v.x + v.y + v.z + v.w;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = 0.f;
    for (int w = 0; w < 8; ++w) tot += red[w];
    s_mean = tot / n;
  }
  __syncthreads();
  const float mean = s_mean;
  float q = 0.f;
  for (int r = r0; r < frames; r += 32) {
    const float4 v = *reinterpret_cast<const float4*>(base + (size_t)r * kD);
    const float a = v.x - mean, b = v.y - mean, c = v.z - mean, d = v.w - mean;
    q += a * a + b * b + c * c + d * d;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  __syncthreads();
  if (lane == 0) red[warp] = q;
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = 0.f;
    for (int w = 0; w < 8; ++w) tot += red[w];
    stats[2 * grp] = mean;
    stats[2 * grp + 1] = rsqrtf(tot / n + 1e-5f);
  }
}

// xn[t][c] = bf16((x[t][c] - mean_g) * rstd_g * w[c] + b[c]); one warp per row
__global__ void __launch_bounds__(256) cond_gn_apply_kernel(const float* __restrict__ x, const float* __restrict__ stats, const float* __restrict__ w,
                                                            const float* __restrict__ b, __nv_bfloat16* __restrict__ xn, int frames) {
  const int lane = threadIdx.x & 31;
  const int gw = blockIdx.x * 8 + (threadIdx.x >> 5), nw = gridDim.x * 8;
  for (int r = gw; r < frames; r += nw) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = 128 * j + 4 * lane;
      const float2 st = *reinterpret_cast<const float2*>(stats + 2 * (c / kCondGroupCh));
      const float4 v = *reinterpret_cast<const float4*>(x + (size_t)r * kD + c);
      const float4 g = *reinterpret_cast<const float4*>(w + c), bb = *reinterpret_cast<const float4*>(b + c);
      *reinterpret_cast<uint2*>(xn + (size_t)r * kD + c) =
          make_uint2(pack_bf16x2((v.x - st.x) * st.y * g.x + bb.x, (v.y - st.x) * st.y * g.y + bb.y),
                     pack_bf16x2((v.z - st.x) * st.y * g.z + bb.z, (v.w - st.x) * st.y * g.w + bb.w));
    }
  }
}

// latent[c] = (first clip ? 0 : latent[c]) + x[0][c] * scale     (h[:, :, 0], then the mean over the clips in clip order)
__global__ void __launch_bounds__(256) cond_take_kernel(const float* __restrict__ x, float* __restrict__ latent, float scale, int first) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c < kD) latent[c] = (first ? 0.f : latent[c]) + x[c] * scale;
}

// qkv conv weight [3072][1024] (+ bias): source channel head * 192 + part * 64 + c  ->  row part * 1024 + head * 64 + c, so the
// GEMM output is q | k | v with 64-dim heads (what full_attn_kernel reads); also bf16 conversion.  One CTA per destination row.
template <typename T>
__global__ void __launch_bounds__(256) cond_pack_qkv_kernel(const T* __restrict__ w, const T* __restrict__ b, __nv_bfloat16* __restrict__ wd, float* __restrict__ bd) {
  const int dst = blockIdx.x;
  const int part = dst >> 10, head = (dst >> 6) & 15, c = dst & 63;
  const int src = head * 192 + part * 64 + c;
  for (int k = threadIdx.x; k < kD; k += 256) wd[(size_t)dst * kD + k] = __float2bfloat16((float)w[(size_t)src * kD + k]);
  if (threadIdx.x == 0) bd[dst] = (float)b[src];
}

// init conv weight [1024][spec_dim] -> fp32 transposed [spec_dim][1024]
template <typename T>
__global__ void __launch_bounds__(256) cond_pack_init_kernel(const T* __restrict__ w, float* __restrict__ wt, int spec_dim) {
  const int k = blockIdx.x;
  for (int c = threadIdx.x; c < kD; c += 256) wt[(size_t)k * kD + c] = (float)w[(size_t)c * spec_dim + k];
}

// ---- mel-spectrogram front end (TorchMelSpectrogram.forward, arch_utils.py:361-395) --------------------------------------------
constexpr int kMelFft = 1024, kMelHop = 256, kMelBins = kMelFft / 2 + 1, kMelN = 80;

// One CTA per (frame, clip): windowed frame (centre padding of 512 samples each side, reflected) -> 513-bin power spectrum by a
// direct DFT against the cos | sin table (index (k * j) mod 1024, fp64 accumulation) -> 80 mel filters -> log(clamp) / norm.
__global__ void __launch_bounds__(256) mel_frames_kernel(const float* __restrict__ wav, int n, const float* __restrict__ win,
                                                         const double* __restrict__ tw, const float* __restrict__ fb,
                                                         const float* __restrict__ norms, float* __restrict__ mel, int frames) {
  __shared__ double cs[kMelFft], sn[kMelFft];
  __shared__ float x[kMelFft], pw[kMelBins + 3];
  const int f = blockIdx.x, clip = blockIdx.y;
  const float* w = wav + (size_t)clip * n;
  for (int j = threadIdx.x; j < kMelFft; j += 256) {
    int i = f * kMelHop + j - kMelFft / 2;
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    x[j] = w[i] * win[j];
    cs[j] = tw[j];
    sn[j] = tw[kMelFft + j];
  }
  __syncthreads();
  for (int k = threadIdx.x; k < kMelBins; k += 256) {
    double re = 0.0, im = 0.0;
#pragma unroll 4
    for (int j = 0; j < kMelFft; ++j) {
      const int idx = (k * j) & (kMelFft - 1);
      re += (double)x[j] * cs[idx];
      im += (double)x[j] * sn[idx];
    }
    pw[k] = (float)(re * re + im * im);
  }
  __syncthreads();
  if (threadIdx.x < kMelN) {
    const int m = threadIdx.x;
    float acc = 0.f;
    for (int k = 0; k < kMelBins; ++k) acc = fmaf(pw[k], fb[k * kMelN + m], acc);
    float v = logf(fmaxf(acc, 1e-5f));
    if (norms) v /= norms[m];
    mel[((size_t)clip * kMelN + m) * frames + f] = v;
  }
}

}  // namespace tts

using namespace tts;

template <typename T>
static int cond_alloc(TtsArHandle* h, T** p, size_t n) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, n * sizeof(T));
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaMalloc(conditioning encoder)");
  h->allocs.push_back(q);
  *p = reinterpret_cast<T*>(q);
  return 0;
}

extern "C" int tts_ar_load_conditioning_weights(TtsArHandle* h, const TtsArCondWeights* w, void* stream) {
  if (!h || !w) return fail(h, TTS_AR_ERR_BAD_ARG, "null argument");
  if (w->struct_size != (int)sizeof(TtsArCondWeights)) return fail(h, TTS_AR_ERR_BAD_ARG, "TtsArCondWeights.struct_size mismatch");
  if (w->dtype != 0 && w->dtype != 1) return fail(h, TTS_AR_ERR_BAD_ARG, "dtype must be 0 (float32) or 1 (bfloat16)");
  if (w->n_blocks < 1 || w->n_blocks > 64 || w->spec_dim < 1 || w->spec_dim > 128 || !w->init_w || !w->init_b || !w->blocks)
    return fail(h, TTS_AR_ERR_BAD_ARG, "conditioning weights: need 1..64 blocks, spec_dim 1..128, non-null pointers");
  if (!h->dense.ok) return fail(h, TTS_AR_ERR_STATE, "dense path not initialised");
  TTS_CUDA(h, cudaSetDevice(h->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  CondW& c = h->cond;
  if (c.blocks.empty() || c.n_blocks != w->n_blocks || c.spec_dim != w->spec_dim) {
    if (!c.blocks.empty()) return fail(h, TTS_AR_ERR_BAD_ARG, "conditioning weights: geometry differs from the first load");
    int rc;
    if ((rc = cond_alloc(h, &c.init_wt, (size_t)w->spec_dim * kD)) || (rc = cond_alloc(h, &c.init_b, kD)) ||
        (rc = cond_alloc(h, &c.stats, 2 * kCondGroups)))
      return rc;
    c.blocks.resize(w->n_blocks);
    bool ok = true;
    for (auto& B : c.blocks) {
      float* f = nullptr;
      if ((rc = cond_alloc(h, &f, (size_t)2 * kD + kD3 + kD)) || (rc = cond_alloc(h, &B.qkv, (size_t)kD3 * kD)) ||
          (rc = cond_alloc(h, &B.proj, (size_t)kD * kD)))
        return rc;
      B.gn_w = f; B.gn_b = f + kD; B.b_qkv = f + 2 * kD; B.b_proj = f + 2 * kD + kD3;
      for (int k = 0; k < 2; ++k) {
        const uint32_t box = k == 0 ? 128 : 256;
        ok = ok && make_map(&B.m_qkv[k], B.qkv, kD3, kD, box) && make_map(&B.m_proj[k], B.proj, kD, kD, box);
      }
    }
    if (!ok) return fail(h, TTS_AR_ERR_CUDA, "cuTensorMapEncodeTiled failed (conditioning encoder)");
    c.n_blocks = w->n_blocks;
    c.spec_dim = w->spec_dim;
  }
  const int dt = w->dtype;
  if (dt == 0)
    cond_pack_init_kernel<float><<<w->spec_dim, 256, 0, st>>>((const float*)w->init_w, c.init_wt, w->spec_dim);
  else
    cond_pack_init_kernel<__nv_bfloat16><<<w->spec_dim, 256, 0, st>>>((const __nv_bfloat16*)w->init_w, c.init_wt, w->spec_dim);
  launch_to_f32(w->init_b, dt, c.init_b, kD, kD, st);
  h->launches += 2;
  for (int i = 0; i < w->n_blocks; ++i) {
    const TtsArCondBlockWeights& bw = w->blocks[i];
    if (!bw.norm_w || !bw.norm_b || !bw.qkv_w || !bw.qkv_b || !bw.proj_w || !bw.proj_b)
      return fail(h, TTS_AR_ERR_BAD_ARG, "null weight pointer in conditioning block");
    CondW::Block& B = c.blocks[i];
    launch_to_f32(bw.norm_w, dt, B.gn_w, kD, kD, st);
    launch_to_f32(bw.norm_b, dt, B.gn_b, kD, kD, st);
    launch_to_f32(bw.proj_b, dt, B.b_proj, kD, kD, st);
    if (dt == 0)
      cond_pack_qkv_kernel<float><<<kD3, 256, 0, st>>>((const float*)bw.qkv_w, (const float*)bw.qkv_b, B.qkv, B.b_qkv);
    else
      cond_pack_qkv_kernel<__nv_bfloat16><<<kD3, 256, 0, st>>>((const __nv_bfloat16*)bw.qkv_w, (const __nv_bfloat16*)bw.qkv_b, B.qkv, B.b_qkv);
    launch_kmajor_pack(bw.proj_w, dt, B.proj, kD, kD, 1, st);
    h->launches += 5;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(h, e, "conditioning weight packing");
  c.loaded = true;
  return 0;
}

extern "C" int tts_ar_get_conditioning(TtsArHandle* h, const float* mel, int32_t n_clips, int32_t frames, float* latent, void* stream) {
  if (!h || !mel || !latent) return fail(h, TTS_AR_ERR_BAD_ARG, "null argument");
  CondW& c = h->cond;
  if (!c.loaded) return fail(h, TTS_AR_ERR_STATE, "conditioning weights not loaded");
  if (h->streaming) return fail(h, TTS_AR_ERR_STATE, "a streaming launch is in flight");
  if (n_clips < 1 || frames < 1) return fail(h, TTS_AR_ERR_BAD_ARG, "get_conditioning: need n_clips >= 1 and frames >= 1");
  if (frames > h->rows_cap) {
    char msg[160];
    snprintf(msg, sizeof msg, "get_conditioning: %d frames exceed the handle's row capacity %d", frames, h->rows_cap);
    return fail(h, TTS_AR_ERR_CAPACITY, msg);
  }
  TTS_CUDA(h, cudaSetDevice(h->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  h->prefilled = false;                        // the activation buffers of a pending prompt pass are reused
  const int n_sm = h->G;
  const int row_grid = std::max(1, std::min((frames + 7) / 8, 4 * n_sm));
  const dim3 attn_grid((frames + 63) / 64, kH, 1);
  __nv_bfloat16* qkv = h->d_h;                 // [frames][3072] inside the [rows_cap][4096] hidden buffer
  TcGemmArgs g{};
  g.M = frames;
  g.watchdog = h->d_watchdog;
  cudaError_t e = cudaSuccess;
  for (int clip = 0; clip < n_clips && e == cudaSuccess; ++clip) {
    cond_init_kernel<<<frames, 256, 0, st>>>(mel + (size_t)clip * c.spec_dim * frames, c.init_wt, c.init_b, h->d_x, c.spec_dim, frames);
    for (int i = 0; i < c.n_blocks; ++i) {
      const CondW::Block& B = c.blocks[i];
      cond_gn_stats_kernel<<<kCondGroups, 256, 0, st>>>(h->d_x, c.stats, frames);
      cond_gn_apply_kernel<<<row_grid, 256, 0, st>>>(h->d_x, c.stats, B.gn_w, B.gn_b, h->d_xn, frames);
      g.N = kD3; g.K = kD; g.bias = B.b_qkv; g.out_bf16 = qkv; g.rot_tab = nullptr;
      e = launch_tc_auto<TC_EPI_ROT>(h->dense.a_xn, B.m_qkv[0], B.m_qkv[1], g, n_sm, st);
      if (e != cudaSuccess) break;
      full_attn_kernel<kD><<<attn_grid, 128, 0, st>>>(qkv, h->d_ctx, frames);
      g.N = kD; g.K = kD; g.bias = B.b_proj; g.out_f32 = h->d_x; g.ldo = kD;
      e = launch_tc_auto<TC_EPI_RESID>(h->dense.a_ctx, B.m_proj[0], B.m_proj[1], g, n_sm, st);
      if (e != cudaSuccess) break;
      h->launches += 5;
    }
    if (e != cudaSuccess) break;
    cond_take_kernel<<<kD / 256, 256, 0, st>>>(h->d_x, latent, 1.0f / (float)n_clips, clip == 0);
    h->launches += 2;
  }
  if (e != cudaSuccess) return cuda_fail(h, e, "conditioning encoder kernel launch");
  e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(h, e, "conditioning encoder");
  return 0;
}

static int mel_tables(TtsArHandle* h) {
  CondW& c = h->cond;
  if (c.mel_fb) return 0;
  std::vector<float> win(kMelFft), fb((size_t)kMelBins * kMelN);
  std::vector<double> tw(2 * kMelFft);
  const double two_pi = 6.283185307179586476925286766559;
  for (int j = 0; j < kMelFft; ++j) {
    win[j] = (float)(0.5 - 0.5 * cos(two_pi * j / kMelFft));          // periodic Hann (torch.hann_window default)
    tw[j] = cos(two_pi * j / kMelFft);
    tw[kMelFft + j] = sin(two_pi * j / kMelFft);
  }
  // torchaudio.functional.melscale_fbanks(513, 0, 8000, 80, 22050, norm="slaney", mel_scale="htk")
  const double sr = 22050.0, f_max = 8000.0, m_max = 2595.0 * log10(1.0 + f_max / 700.0);
  std::vector<double> fp(kMelN + 2);
  for (int i = 0; i < kMelN + 2; ++i) fp[i] = 700.0 * (pow(10.0, (m_max * i / (kMelN + 1)) / 2595.0) - 1.0);
  for (int k = 0; k < kMelBins; ++k) {
    const double fr = (double)((int)sr / 2) * k / (kMelBins - 1);
    for (int m = 0; m < kMelN; ++m) {
      const double down = (fr - fp[m]) / (fp[m + 1] - fp[m]), up = (fp[m + 2] - fr) / (fp[m + 2] - fp[m + 1]);
      const double tri = std::max(0.0, std::min(down, up));
      fb[(size_t)k * kMelN + m] = (float)(tri * 2.0 / (fp[m + 2] - fp[m]));
    }
  }
  int rc;
  if ((rc = cond_alloc(h, &c.mel_win, win.size())) || (rc = cond_alloc(h, &c.mel_tw, tw.size()))) return rc;
  float* fbp = nullptr;
  if ((rc = cond_alloc(h, &fbp, fb.size()))) return rc;
  TTS_CUDA(h, cudaMemcpy(c.mel_win, win.data(), win.size() * sizeof(float), cudaMemcpyHostToDevice));
  TTS_CUDA(h, cudaMemcpy(c.mel_tw, tw.data(), tw.size() * sizeof(double), cudaMemcpyHostToDevice));
  TTS_CUDA(h, cudaMemcpy(fbp, fb.data(), fb.size() * sizeof(float), cudaMemcpyHostToDevice));
  c.mel_fb = fbp;
  return 0;
}

extern "C" int tts_ar_mel_spectrogram(TtsArHandle* h, const float* wav, int32_t n_wav, int32_t n_samples, const float* mel_norms,
                                      float* mel, void* stream) {
  if (!h || !wav || !mel) return fail(h, TTS_AR_ERR_BAD_ARG, "null argument");
  if (n_wav < 1 || n_wav > 65535 || n_samples <= kMelFft / 2)
    return fail(h, TTS_AR_ERR_BAD_ARG, "mel_spectrogram: need 1..65535 clips of more than 512 samples (reflect padding)");
  TTS_CUDA(h, cudaSetDevice(h->cfg.device));
  const int rc = mel_tables(h);
  if (rc) return rc;
  const int frames = 1 + n_samples / kMelHop;
  mel_frames_kernel<<<dim3(frames, n_wav), 256, 0, (cudaStream_t)stream>>>(wav, n_samples, h->cond.mel_win, h->cond.mel_tw, h->cond.mel_fb,
                                                                           mel_norms, mel, frames);
  h->launches += 1;
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(h, e, "mel_frames_kernel launch");
  return 0;
}
