// clvp.cu -- CLVP candidate scoring: the two x-transformers encoders of the reference's CLVP module,
// pooled, projected and compared, behind tts_ar_clvp_* (include/tts_ar.h).
// Replaces: CLVP.forward(text, speech_tokens, return_loss=False) in eval mode (reference tortoise_tts/models/clvp.py:100-133)
// as called by tortoise_tts/inference.py:392-396.  Per encoder layer (xtransformers.py:912-1016, pre-norm):
//     rms_rows (RMSNorm, :337-346) -> q|k|v GEMM (tcgen05; rotary on the first 32 dims of q, k, v in the epilogue, :279-290,
//     660-664) -> full softmax attention per (sequence, head) (:672-716) -> to_out GEMM (+bias, +residual) ->
//     rms_rows -> GLU GEMM (value * gelu_erf(gate) in the epilogue, :431-440) -> ff out GEMM (+bias, +residual)
// tail: LayerNorm (:1217,1238) -> mean over positions (clvp.py:15-17: eval masks are all-true) -> to_latent -> L2 normalise;
// sim = <text, speech> * exp(temperature) (clvp.py:123-132).
// Rows = every position of every candidate (16 x 500 = 8000 for BASELINE config C2), so the GEMMs are dense contractions
// and run on the tensor-core kernel of the dense passes (tc_gemm.cuh); everything else is HBM-bound row work in fp32.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include "tc_launch.cuh"
#include "full_attn.cuh"
#include "ar_launch.h"
#include "../../include/tts_ar.h"

namespace tts {

constexpr int kCD = 768;                  // model width
constexpr int kCH = 12;                   // heads
constexpr int kCQKV = 3 * kCD;            // 2304
constexpr int kCFF = 1536;                // hidden units of the feed-forward (dim * ff_mult)
constexpr int kCRot = 32;                 // rotary dims per head

// ---- embedding rows ---------------------------------------------------------------------------------------------------
// x[seq * len + p][:] = emb[token]; tokens come as int32 (text) or int64 (codes, the AR sampler's output layout)
template <typename TokT>
__global__ void __launch_bounds__(192) clvp_embed_kernel(float* __restrict__ x, const float* __restrict__ emb, const TokT* __restrict__ tok,
                                                         int len, int stride, int n_tokens, int* err_flag) {
  const int row = blockIdx.x;
  const int s = row / len, p = row - s * len;
  long long id = (long long)tok[(size_t)s * stride + p];
  if (id < 0 || id >= n_tokens) {
    if (threadIdx.x == 0) *err_flag = 1;
    id = 0;
  }
  reinterpret_cast<float4*>(x + (size_t)row * kCD)[threadIdx.x] = reinterpret_cast<const float4*>(emb + (size_t)id * kCD)[threadIdx.x];
}

// ---- RMSNorm of rows -> bf16 A operand (xtransformers.py:337-346: x / max(||x|| * dim^-0.5, 1e-8) * g) -------------------
__global__ void __launch_bounds__(256) clvp_rms_kernel(const float* __restrict__ x, const float* __restrict__ gain, __nv_bfloat16* __restrict__ xn, int n) {
  const int lane = threadIdx.x & 31;
  const int gw = blockIdx.x * 8 + (threadIdx.x >> 5), nw = gridDim.x * 8;
  for (int row = gw; row < n; row += nw) {
    const float* xp = x + (size_t)row * kCD;
    float4 v[6];
    float ss = 0.f;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      v[j] = *reinterpret_cast<const float4*>(xp + 128 * j + 4 * lane);
      ss += v[j].x * v[j].x + v[j].y * v[j].y + v[j].z * v[j].z + v[j].w * v[j].w;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    const float inv = 1.0f / fmaxf(sqrtf(ss) * 0.03608439182435161f, 1e-8f);      // 768^-0.5
    __nv_bfloat16* op = xn + (size_t)row * kCD;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const float4 g = *reinterpret_cast<const float4*>(gain + 128 * j + 4 * lane);
      *reinterpret_cast<uint2*>(op + 128 * j + 4 * lane) =
          make_uint2(pack_bf16x2(v[j].x * inv * g.x, v[j].y * inv * g.y), pack_bf16x2(v[j].z * inv * g.z, v[j].w * inv * g.w));
    }
  }
}

// ---- tail: final LayerNorm of every row, mean over the sequence, to_latent, L2 normalise -------------------------------------
// Two kernels.  clvp_pool_partial_kernel: CTA (sequence, slice p of kPoolSlices): warp w normalises rows p + kPoolSlices (w + 8 i)
// (two-pass statistics, eps 1e-5, biased variance as F.layer_norm) and keeps a running sum; the 8 warp sums are added in warp
// order -> partial[seq][p][768].  clvp_pool_final_kernel: one CTA per sequence adds the slices in order (deterministic), applies
// the LayerNorm affine to the mean, lat = W . mean (fp32 weights, coalesced rows), lat / max(||lat||, 1e-12) (F.normalize).
constexpr int kPoolSlices = 8;

__global__ void __launch_bounds__(256) clvp_pool_partial_kernel(const float* __restrict__ x, float* __restrict__ partial, int len) {
  __shared__ float part[8][kCD];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int seq = blockIdx.x, slice = blockIdx.y;
  float acc[24];
#pragma unroll
  for (int i = 0; i < 24; ++i) acc[i] = 0.f;
  for (int r = slice + kPoolSlices * warp; r < len; r += kPoolSlices * 8) {
    const float* xp = x + ((size_t)seq * len + r) * kCD;
    float4 v[6];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      v[j] = *reinterpret_cast<const float4*>(xp + 128 * j + 4 * lane);
      s += v[j].x + v[j].y + v[j].z + v[j].w;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mu = s * (1.0f / kCD);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      v[j].x -= mu; v[j].y -= mu; v[j].z -= mu; v[j].w -= mu;
      q += v[j].x * v[j].x + v[j].y * v[j].y + v[j].z * v[j].z + v[j].w * v[j].w;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q * (1.0f / kCD) + 1e-5f);
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      acc[4 * j] += v[j].x * rstd; acc[4 * j + 1] += v[j].y * rstd; acc[4 * j + 2] += v[j].z * rstd; acc[4 * j + 3] += v[j].w * rstd;
    }
  }
#pragma unroll
  for (int j = 0; j < 6; ++j)
    *reinterpret_cast<float4*>(&part[warp][128 * j + 4 * lane]) = make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
  __syncthreads();
  for (int c = threadIdx.x; c < kCD; c += 256) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += part[w][c];
    partial[((size_t)seq * kPoolSlices + slice) * kCD + c] = s;
  }
}

__global__ void __launch_bounds__(256) clvp_pool_final_kernel(const float* __restrict__ partial, const float* __restrict__ nw,
                                                              const float* __restrict__ nb, const float* __restrict__ w_lat,
                                                              float* __restrict__ lat_out, int len) {
  __shared__ float mean[kCD];
  __shared__ float lat[kCD];
  __shared__ float red[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int seq = blockIdx.x;
  // mean of LN(x) = (sum of normalised rows / len) * weight + bias  (the affine part is the same for every row)
  for (int c = threadIdx.x; c < kCD; c += 256) {
    float s = 0.f;
#pragma unroll
    for (int p = 0; p < kPoolSlices; ++p) s += partial[((size_t)seq * kPoolSlices + p) * kCD + c];
    mean[c] = s / (float)len * nw[c] + nb[c];
  }
  __syncthreads();
  float sq = 0.f;
  for (int j = warp; j < kCD; j += 8) {
    const float* wr = w_lat + (size_t)j * kCD;
    float d = 0.f;
#pragma unroll
    for (int i = 0; i < kCD / 32; ++i) d += wr[lane + 32 * i] * mean[lane + 32 * i];
#pragma unroll
    for (int o = 16; o; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if (lane == 0) lat[j] = d;
    sq += d * d;                                   // every lane holds d
  }
  if (lane == 0) red[warp] = sq;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) tot += red[w];
  const float inv = 1.0f / fmaxf(sqrtf(tot), 1e-12f);
  for (int c = threadIdx.x; c < kCD; c += 256) lat_out[(size_t)seq * kCD + c] = lat[c] * inv;
}

// scores[i] = <text_lat[n_text == 1 ? 0 : i], speech_lat[i]> * exp(temperature)   (clvp.py:128-132); one warp per candidate
__global__ void __launch_bounds__(256) clvp_score_kernel(const float* __restrict__ text_lat, const float* __restrict__ speech_lat,
                                                         const float* __restrict__ temperature, int n_text, int n, float* __restrict__ scores,
                                                         float* __restrict__ speech_out) {
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (i >= n) return;
  const float* a = text_lat + (size_t)(n_text == 1 ? 0 : i) * kCD;
  const float* b = speech_lat + (size_t)i * kCD;
  float d = 0.f;
#pragma unroll
  for (int k = 0; k < kCD / 32; ++k) {
    d += a[lane + 32 * k] * b[lane + 32 * k];
    if (speech_out) speech_out[(size_t)i * kCD + lane + 32 * k] = b[lane + 32 * k];
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
  if (lane == 0) scores[i] = d * expf(temperature[0]);
}

// indices[r] = the candidate with the r-th largest score, r < k (torch.topk(scores, k).indices of reference inference.py:395; ties:
// the lower index first).  Rank by counting -- n is a candidate count (<= 4096), one CTA.
__global__ void __launch_bounds__(256) clvp_topk_kernel(const float* __restrict__ scores, int n, int k, long long* __restrict__ indices) {
  for (int i = threadIdx.x; i < n; i += 256) {
    const float si = scores[i];
    int rank = 0;
    for (int j = 0; j < n; ++j) {
      const float sj = scores[j];
      rank += (sj > si || (sj == si && j < i)) ? 1 : 0;
    }
    if (rank < k) indices[rank] = i;
  }
}

// ---- GLU weight image: rows interleaved per 32 hidden units -----------------------------------------------------------------
// src nn.Linear [3072][768]: rows [0,1536) = value, [1536,3072) = gate (GLU.forward: proj(x).chunk(2), xtransformers.py:438).
// dst row (u / 32) * 64 + (u % 32) = value row u, + 32 = gate row u, so that one 64-column slice of an output tile holds both
// halves of 32 hidden units (tc_gemm.cuh TC_EPI_GLU).  The bias is permuted the same way.
template <typename T>
__global__ void __launch_bounds__(192) clvp_pack_glu_kernel(const T* __restrict__ w, const T* __restrict__ b, __nv_bfloat16* __restrict__ wd, float* __restrict__ bd) {
  const int dst = blockIdx.x;                         // 0 .. 3071
  const int blk = dst >> 6, in = dst & 63;
  const int u = blk * 32 + (in & 31);
  const int src = (in < 32) ? u : kCFF + u;
  const T* wr = w + (size_t)src * kCD;
  for (int k = threadIdx.x; k < kCD; k += 192) wd[(size_t)dst * kCD + k] = __float2bfloat16((float)wr[k]);
  if (threadIdx.x == 0) bd[dst] = (float)b[src];
}

}  // namespace tts

using namespace tts;

// ---- handle -------------------------------------------------------------------------------------------------------------------
struct ClvpLayerDev {
  __nv_bfloat16 *qkv, *out, *glu, *ff2;               // K-major bf16 images: [2304][768], [768][768], [3072][768] (interleaved), [768][1536]
  float *g_attn, *b_out, *g_ff, *b_glu, *b_ff2;       // fp32
  CUtensorMap m_qkv[2], m_out[2], m_glu[2], m_ff2[2]; // box rows 128 / 256
};

struct ClvpEncDev {
  float* emb = nullptr;
  int n_tokens = 0;
  float *norm_w = nullptr, *norm_b = nullptr, *w_lat = nullptr;
  std::vector<ClvpLayerDev> layers;
};

struct TtsClvpHandle {
  TtsClvpConfig cfg{};
  std::string err;
  int n_sm = 0;
  int rows_cap = 0;
  std::vector<void*> allocs;
  ClvpEncDev enc[2];                                  // 0 = text, 1 = speech
  float* d_temp = nullptr;
  float *d_x = nullptr, *d_rot = nullptr, *d_lat_text = nullptr, *d_lat_speech = nullptr, *d_pool = nullptr;
  __nv_bfloat16 *d_xn = nullptr, *d_qkv = nullptr, *d_ctx = nullptr, *d_h = nullptr;
  int* d_flags = nullptr;                             // [0] bad token, [1] watchdog
  int* h_flags = nullptr;                             // pinned
  CUtensorMap a_xn, a_ctx, a_h;
  bool loaded = false;
  long long launches = 0;
};

static std::string g_clvp_create_error;

static int cfail(TtsClvpHandle* h, int code, const std::string& msg) {
  if (h) h->err = msg; else g_clvp_create_error = msg;
  return code;
}
static int ccuda(TtsClvpHandle* h, cudaError_t e, const char* what) {
  return cfail(h, TTS_AR_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

template <typename T>
static int calloc_dev(TtsClvpHandle* h, T** p, size_t n) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, n * sizeof(T));
  if (e != cudaSuccess) return ccuda(h, e, "cudaMalloc");
  h->allocs.push_back(q);
  *p = reinterpret_cast<T*>(q);
  return 0;
}

extern "C" const char* tts_ar_clvp_last_error(TtsClvpHandle* h) { return h ? h->err.c_str() : g_clvp_create_error.c_str(); }

extern "C" void tts_ar_clvp_destroy(TtsClvpHandle* h) {
  if (!h) return;
  cudaSetDevice(h->cfg.device);
  for (void* p : h->allocs) cudaFree(p);
  if (h->h_flags) cudaFreeHost(h->h_flags);
  delete h;
}

static int clvp_create_impl(TtsClvpHandle* h) {
  const TtsClvpConfig& c = h->cfg;
  cudaError_t e = cudaSetDevice(c.device);
  if (e != cudaSuccess) return ccuda(h, e, "cudaSetDevice");
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, c.device);
  if (e != cudaSuccess) return ccuda(h, e, "cudaGetDeviceProperties");
  if (prop.major != 10) return cfail(h, TTS_AR_ERR_DEVICE, "tts_ar_clvp needs an sm_100 (Blackwell B200) device; kernels are built for sm_100a only");
  if (!encode_fn()) return cfail(h, TTS_AR_ERR_DEVICE, "cuTensorMapEncodeTiled is not available from this driver");
  h->n_sm = prop.multiProcessorCount;
  h->rows_cap = c.max_seqs * std::max(c.max_text_len, c.max_speech_len);
  const size_t R = (size_t)h->rows_cap;
  int rc;
  if ((rc = calloc_dev(h, &h->d_x, R * kCD))) return rc;
  if ((rc = calloc_dev(h, &h->d_xn, R * kCD))) return rc;
  if ((rc = calloc_dev(h, &h->d_ctx, R * kCD))) return rc;
  if ((rc = calloc_dev(h, &h->d_qkv, R * kCQKV))) return rc;
  if ((rc = calloc_dev(h, &h->d_h, R * kCFF))) return rc;
  if ((rc = calloc_dev(h, &h->d_lat_text, (size_t)c.max_seqs * kCD))) return rc;
  if ((rc = calloc_dev(h, &h->d_lat_speech, (size_t)c.max_seqs * kCD))) return rc;
  if ((rc = calloc_dev(h, &h->d_pool, (size_t)c.max_seqs * kPoolSlices * kCD))) return rc;
  if ((rc = calloc_dev(h, &h->d_temp, 4))) return rc;
  if ((rc = calloc_dev(h, &h->d_flags, 4))) return rc;
  e = cudaMemset(h->d_flags, 0, 4 * sizeof(int));
  if (e != cudaSuccess) return ccuda(h, e, "cudaMemset");
  e = cudaMallocHost(reinterpret_cast<void**>(&h->h_flags), 4 * sizeof(int));
  if (e != cudaSuccess) return ccuda(h, e, "cudaMallocHost");
  // rotary table (xtransformers.py:266-276): inv_freq[i] = 10000^(-2 i / 32); row p = cos(p * inv_freq[0..15]), sin(...)
  const int max_len = std::max(c.max_text_len, c.max_speech_len);
  std::vector<float> rot((size_t)max_len * kCRot);
  for (int p = 0; p < max_len; ++p)
    for (int i = 0; i < kCRot / 2; ++i) {
      const float inv_freq = 1.0f / powf(10000.0f, (float)(2 * i) / (float)kCRot);
      const float f = (float)p * inv_freq;
      rot[(size_t)p * kCRot + i] = cosf(f);
      rot[(size_t)p * kCRot + kCRot / 2 + i] = sinf(f);
    }
  if ((rc = calloc_dev(h, &h->d_rot, rot.size()))) return rc;
  e = cudaMemcpy(h->d_rot, rot.data(), rot.size() * sizeof(float), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) return ccuda(h, e, "cudaMemcpy(rotary table)");
  // weights: one bf16 arena per encoder layer + fp32 vectors
  const size_t per_layer = (size_t)kCQKV * kCD + (size_t)kCD * kCD + (size_t)2 * kCFF * kCD + (size_t)kCD * kCFF;
  const size_t per_layer_f = 3 * (size_t)kCD + 2 * kCFF + 2 * (size_t)kCD;     // g_attn, b_out, g_ff | b_glu | b_ff2 (+ pad)
  for (int s = 0; s < 2; ++s) {
    ClvpEncDev& E = h->enc[s];
    E.n_tokens = s == 0 ? c.n_text_tokens : c.n_speech_tokens;
    if ((rc = calloc_dev(h, &E.emb, (size_t)E.n_tokens * kCD))) return rc;
    if ((rc = calloc_dev(h, &E.norm_w, kCD)) || (rc = calloc_dev(h, &E.norm_b, kCD))) return rc;
    if ((rc = calloc_dev(h, &E.w_lat, (size_t)kCD * kCD))) return rc;
    __nv_bfloat16* wt = nullptr;
    float* fv = nullptr;
    if ((rc = calloc_dev(h, &wt, per_layer * c.depth))) return rc;
    if ((rc = calloc_dev(h, &fv, per_layer_f * c.depth))) return rc;
    E.layers.resize(c.depth);
    bool ok = true;
    for (int l = 0; l < c.depth; ++l) {
      ClvpLayerDev& L = E.layers[l];
      __nv_bfloat16* b = wt + per_layer * l;
      L.qkv = b; L.out = L.qkv + (size_t)kCQKV * kCD; L.glu = L.out + (size_t)kCD * kCD; L.ff2 = L.glu + (size_t)2 * kCFF * kCD;
      float* f = fv + per_layer_f * l;
      L.g_attn = f; L.b_out = f + kCD; L.g_ff = f + 2 * kCD; L.b_glu = f + 3 * kCD; L.b_ff2 = L.b_glu + 2 * kCFF;
      for (int k = 0; k < 2; ++k) {
        const uint32_t box = k == 0 ? 128 : 256;
        ok = ok && make_map(&L.m_qkv[k], L.qkv, kCQKV, kCD, box) && make_map(&L.m_out[k], L.out, kCD, kCD, box) &&
             make_map(&L.m_glu[k], L.glu, 2 * kCFF, kCD, box) && make_map(&L.m_ff2[k], L.ff2, kCD, kCFF, box);
      }
    }
    if (!ok) return cfail(h, TTS_AR_ERR_CUDA, "cuTensorMapEncodeTiled failed (weights)");
  }
  if (!make_map(&h->a_xn, h->d_xn, R, kCD, kTcBM) || !make_map(&h->a_ctx, h->d_ctx, R, kCD, kTcBM) ||
      !make_map(&h->a_h, h->d_h, R, kCFF, kTcBM))
    return cfail(h, TTS_AR_ERR_CUDA, "cuTensorMapEncodeTiled failed (activations)");
  e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return ccuda(h, e, "cudaDeviceSynchronize");
  return 0;
}

extern "C" int tts_ar_clvp_create(const TtsClvpConfig* cfg, TtsClvpHandle** out) {
  if (!cfg || !out) return cfail(nullptr, TTS_AR_ERR_BAD_ARG, "null argument");
  if (cfg->struct_size != (int)sizeof(TtsClvpConfig)) return cfail(nullptr, TTS_AR_ERR_BAD_ARG, "TtsClvpConfig.struct_size mismatch");
  if (cfg->dim != kCD || cfg->heads != kCH) return cfail(nullptr, TTS_AR_ERR_BAD_ARG, "clvp: dim must be 768 and heads 12");
  if (cfg->depth < 1 || cfg->depth > 64 || cfg->n_text_tokens < 1 || cfg->n_speech_tokens < 1 || cfg->max_seqs < 1 ||
      cfg->max_text_len < 1 || cfg->max_speech_len < 1 || (long long)cfg->max_seqs * std::max(cfg->max_text_len, cfg->max_speech_len) > (1 << 22))
    return cfail(nullptr, TTS_AR_ERR_BAD_ARG, "TtsClvpConfig out of range");
  TtsClvpHandle* h = new TtsClvpHandle();
  h->cfg = *cfg;
  const int rc = clvp_create_impl(h);
  if (rc != 0) {
    g_clvp_create_error = h->err;
    tts_ar_clvp_destroy(h);
    return rc;
  }
  *out = h;
  return 0;
}

extern "C" int tts_ar_clvp_load_weights(TtsClvpHandle* h, const TtsClvpWeights* w, void* stream) {
  if (!h || !w) return cfail(h, TTS_AR_ERR_BAD_ARG, "null argument");
  if (w->struct_size != (int)sizeof(TtsClvpWeights)) return cfail(h, TTS_AR_ERR_BAD_ARG, "TtsClvpWeights.struct_size mismatch");
  if (w->depth != h->cfg.depth) return cfail(h, TTS_AR_ERR_BAD_ARG, "clvp: layer count mismatch");
  if (w->dtype != 0 && w->dtype != 1) return cfail(h, TTS_AR_ERR_BAD_ARG, "dtype must be 0 (float32) or 1 (bfloat16)");
  if (!w->temperature) return cfail(h, TTS_AR_ERR_BAD_ARG, "null weight pointer");
  cudaError_t e = cudaSetDevice(h->cfg.device);
  if (e != cudaSuccess) return ccuda(h, e, "cudaSetDevice");
  cudaStream_t st = (cudaStream_t)stream;
  const int dt = w->dtype;
  const TtsClvpEncoderWeights* ew[2] = {&w->text, &w->speech};
  for (int s = 0; s < 2; ++s) {
    const TtsClvpEncoderWeights& W = *ew[s];
    ClvpEncDev& E = h->enc[s];
    if (!W.emb || !W.layers || !W.norm_w || !W.norm_b || !W.to_latent_w) return cfail(h, TTS_AR_ERR_BAD_ARG, "null weight pointer");
    launch_to_f32(W.emb, dt, E.emb, (size_t)E.n_tokens * kCD, (size_t)E.n_tokens * kCD, st);
    launch_to_f32(W.norm_w, dt, E.norm_w, kCD, kCD, st);
    launch_to_f32(W.norm_b, dt, E.norm_b, kCD, kCD, st);
    launch_to_f32(W.to_latent_w, dt, E.w_lat, (size_t)kCD * kCD, (size_t)kCD * kCD, st);
    h->launches += 4;
    for (int l = 0; l < h->cfg.depth; ++l) {
      const TtsClvpLayerWeights& lw = W.layers[l];
      const void* ps[] = {lw.attn_norm_g, lw.to_q_w, lw.to_k_w, lw.to_v_w, lw.to_out_w, lw.to_out_b, lw.ff_norm_g, lw.ff_proj_w,
                          lw.ff_proj_b, lw.ff_out_w, lw.ff_out_b};
      for (const void* p : ps)
        if (!p) return cfail(h, TTS_AR_ERR_BAD_ARG, "null weight pointer in layer");
      ClvpLayerDev& L = E.layers[l];
      launch_kmajor_pack(lw.to_q_w, dt, L.qkv, kCD, kCD, 1, st);
      launch_kmajor_pack(lw.to_k_w, dt, L.qkv + (size_t)kCD * kCD, kCD, kCD, 1, st);
      launch_kmajor_pack(lw.to_v_w, dt, L.qkv + (size_t)2 * kCD * kCD, kCD, kCD, 1, st);
      launch_kmajor_pack(lw.to_out_w, dt, L.out, kCD, kCD, 1, st);
      launch_kmajor_pack(lw.ff_out_w, dt, L.ff2, kCFF, kCD, 1, st);
      if (dt == 0)
        clvp_pack_glu_kernel<float><<<2 * kCFF, 192, 0, st>>>((const float*)lw.ff_proj_w, (const float*)lw.ff_proj_b, L.glu, L.b_glu);
      else
        clvp_pack_glu_kernel<__nv_bfloat16><<<2 * kCFF, 192, 0, st>>>((const __nv_bfloat16*)lw.ff_proj_w, (const __nv_bfloat16*)lw.ff_proj_b, L.glu, L.b_glu);
      launch_to_f32(lw.attn_norm_g, dt, L.g_attn, kCD, kCD, st);
      launch_to_f32(lw.to_out_b, dt, L.b_out, kCD, kCD, st);
      launch_to_f32(lw.ff_norm_g, dt, L.g_ff, kCD, kCD, st);
      launch_to_f32(lw.ff_out_b, dt, L.b_ff2, kCD, kCD, st);
      h->launches += 10;
    }
  }
  launch_to_f32(w->temperature, dt, h->d_temp, 1, 1, st);
  h->launches += 1;
  e = cudaGetLastError();
  if (e != cudaSuccess) return ccuda(h, e, "clvp weight packing");
  h->loaded = true;
  return 0;
}

// one encoder over n_seq sequences of `len` tokens -> L2-normalised latents [n_seq][768]
template <typename TokT>
static int clvp_encode(TtsClvpHandle* h, int side, const TokT* tok, int n_seq, int len, int stride, float* lat_out, cudaStream_t st) {
  ClvpEncDev& E = h->enc[side];
  const int M = n_seq * len, n_sm = h->n_sm;
  clvp_embed_kernel<TokT><<<M, 192, 0, st>>>(h->d_x, E.emb, tok, len, stride, E.n_tokens, h->d_flags);
  const int rms_grid = std::max(1, std::min((M + 7) / 8, 8 * n_sm));
  const dim3 attn_grid((len + 63) / 64, kCH, n_seq);
  TcGemmArgs g{};
  g.M = M;
  g.watchdog = h->d_flags + 1;
  g.rot_tab = h->d_rot;
  g.seq_len = len;
  cudaError_t e = cudaSuccess;
  for (int l = 0; l < h->cfg.depth && e == cudaSuccess; ++l) {
    const ClvpLayerDev& L = E.layers[l];
    clvp_rms_kernel<<<rms_grid, 256, 0, st>>>(h->d_x, L.g_attn, h->d_xn, M);
    g.N = kCQKV; g.K = kCD; g.bias = nullptr; g.out_bf16 = h->d_qkv;
    e = launch_tc_auto<TC_EPI_ROT>(h->a_xn, L.m_qkv[0], L.m_qkv[1], g, n_sm, st);
    if (e != cudaSuccess) break;
    full_attn_kernel<kCD><<<attn_grid, 128, 0, st>>>(h->d_qkv, h->d_ctx, len);
    g.N = kCD; g.K = kCD; g.bias = L.b_out; g.out_f32 = h->d_x; g.ldo = kCD;
    e = launch_tc_auto<TC_EPI_RESID>(h->a_ctx, L.m_out[0], L.m_out[1], g, n_sm, st);
    if (e != cudaSuccess) break;
    clvp_rms_kernel<<<rms_grid, 256, 0, st>>>(h->d_x, L.g_ff, h->d_xn, M);
    g.N = 2 * kCFF; g.K = kCD; g.bias = L.b_glu; g.out_bf16 = h->d_h;
    e = launch_tc_auto<TC_EPI_GLU>(h->a_xn, L.m_glu[0], L.m_glu[1], g, n_sm, st);
    if (e != cudaSuccess) break;
    g.N = kCD; g.K = kCFF; g.bias = L.b_ff2; g.out_f32 = h->d_x; g.ldo = kCD;
    e = launch_tc_auto<TC_EPI_RESID>(h->a_h, L.m_ff2[0], L.m_ff2[1], g, n_sm, st);
    h->launches += 7;
  }
  if (e != cudaSuccess) return ccuda(h, e, "clvp encoder kernel launch");
  clvp_pool_partial_kernel<<<dim3(n_seq, kPoolSlices), 256, 0, st>>>(h->d_x, h->d_pool, len);
  clvp_pool_final_kernel<<<n_seq, 256, 0, st>>>(h->d_pool, E.norm_w, E.norm_b, E.w_lat, lat_out, len);
  h->launches += 3;
  e = cudaGetLastError();
  if (e != cudaSuccess) return ccuda(h, e, "clvp encoder");
  return 0;
}

extern "C" int tts_ar_clvp_score(TtsClvpHandle* h, const int32_t* text, int32_t n_text, int32_t text_len, int32_t text_stride,
                                 const int64_t* codes, int32_t n, int32_t codes_len, int32_t codes_stride, float* scores,
                                 float* speech_latents, void* stream) {
  if (!h || !text || !codes || !scores) return cfail(h, TTS_AR_ERR_BAD_ARG, "null argument");
  if (!h->loaded) return cfail(h, TTS_AR_ERR_STATE, "clvp: weights not loaded");
  if (n < 1 || (n_text != 1 && n_text != n) || text_len < 1 || codes_len < 1 || text_stride < text_len || codes_stride < codes_len)
    return cfail(h, TTS_AR_ERR_BAD_ARG, "clvp_score: need n >= 1, n_text in {1, n}, lengths >= 1, strides >= lengths");
  const TtsClvpConfig& c = h->cfg;
  if (n > c.max_seqs || text_len > c.max_text_len || codes_len > c.max_speech_len)
    return cfail(h, TTS_AR_ERR_CAPACITY, "clvp_score: batch / length exceeds what tts_ar_clvp_create was sized for");
  cudaError_t e = cudaSetDevice(c.device);
  if (e != cudaSuccess) return ccuda(h, e, "cudaSetDevice");
  cudaStream_t st = (cudaStream_t)stream;
  int rc = clvp_encode<int32_t>(h, 0, text, n_text, text_len, text_stride, h->d_lat_text, st);
  if (rc) return rc;
  rc = clvp_encode<long long>(h, 1, reinterpret_cast<const long long*>(codes), n, codes_len, codes_stride, h->d_lat_speech, st);
  if (rc) return rc;
  clvp_score_kernel<<<(n + 7) / 8, 256, 0, st>>>(h->d_lat_text, h->d_lat_speech, h->d_temp, n_text, n, scores, speech_latents);
  h->launches += 1;
  e = cudaGetLastError();
  if (e != cudaSuccess) return ccuda(h, e, "clvp_score_kernel launch");
  return 0;
}

extern "C" int tts_ar_clvp_topk(TtsClvpHandle* h, const float* scores, int32_t n, int32_t k, int64_t* indices, void* stream) {
  if (!h || !scores || !indices) return cfail(h, TTS_AR_ERR_BAD_ARG, "null argument");
  if (n < 1 || n > 4096 || k < 1 || k > n) return cfail(h, TTS_AR_ERR_BAD_ARG, "clvp_topk: need 1 <= k <= n <= 4096");
  cudaError_t e = cudaSetDevice(h->cfg.device);
  if (e != cudaSuccess) return ccuda(h, e, "cudaSetDevice");
  clvp_topk_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(scores, n, k, reinterpret_cast<long long*>(indices));
  h->launches += 1;
  e = cudaGetLastError();
  if (e != cudaSuccess) return ccuda(h, e, "clvp_topk_kernel launch");
  return 0;
}

extern "C" int tts_ar_clvp_sync(TtsClvpHandle* h, int64_t* launches, void* stream) {
  if (!h) return TTS_AR_ERR_BAD_ARG;
  cudaError_t e = cudaSetDevice(h->cfg.device);
  if (e != cudaSuccess) return ccuda(h, e, "cudaSetDevice");
  cudaStream_t st = (cudaStream_t)stream;
  e = cudaMemcpyAsync(h->h_flags, h->d_flags, 2 * sizeof(int), cudaMemcpyDeviceToHost, st);
  if (e != cudaSuccess) return ccuda(h, e, "cudaMemcpyAsync(flags)");
  e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return ccuda(h, e, "cudaStreamSynchronize");
  if (launches) *launches = h->launches;
  const int bad = h->h_flags[0], wd = h->h_flags[1];
  if (bad || wd) {
    e = cudaMemsetAsync(h->d_flags, 0, 2 * sizeof(int), st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return ccuda(h, e, "clearing the flags");
  }
  if (wd) return cfail(h, TTS_AR_ERR_WATCHDOG, "clvp: a device-side wait timed out; results are invalid");
  if (bad) return cfail(h, TTS_AR_ERR_BAD_ARG, "clvp: a token id was outside its embedding table");
  return 0;
}
