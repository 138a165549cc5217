// dense_pass.cu -- the prompt-shaped passes (rows = positions of whole sequences) as a pipeline of dense kernels:
//     ln_1 -> c_attn (tcgen05, K/V page scatter) -> causal attention -> attn.c_proj (+residual)
//          -> ln_2 -> c_fc (tcgen05, gelu_new) -> mlp.c_proj (+residual)          x n_layers
//     tail: ln_f + final_norm of the last row per utterance -> mel head (tcgen05) -> logits      (prompt pass)
//        or ln_f + final_norm of every mel row -> latents                                          (latent re-forward)
// Replaces: GPT2InferenceModel.forward prefill branch (reference unified_voice.py:203-211, 216-239) and
// UnifiedVoice.forward(return_latent=True) / get_logits (unified_voice.py:508-533, 544-600).
// The GEMMs are real dense contractions here (68 .. 48,896 rows), so they run on tcgen05 / TMEM with TMA-staged
// operands (tc_gemm.cuh); the weights are read from a K-major bf16 image built once at load time.
#include <cstdio>
#include <cstring>
#include <algorithm>
#include "tc_launch.cuh"
#include "ar_attn.cuh"
#include "tts_ar_host.h"

namespace tts {

// ---- LayerNorm of rows -> bf16 row-major A operand -------------------------------------------------------------------
// One warp per row, fp32 statistics (two-pass), eps 1e-5.
//   mode 0: rows [0, n) of x with (g1, b1)                                   -> xn[row]
//   mode 1: rows head_rows[i], ln_f then final_norm (quirk Q1)               -> xn[i] and cur_latent[i] (fp32)
//   mode 2: rows with latent_dst[row] >= 0, ln_f then final_norm             -> latent_out[latent_dst[row]] (fp32)
struct LnArgs {
  const float* x;
  const float *g1, *b1, *g2, *b2;
  __nv_bfloat16* xn;
  const int* head_rows;
  float* cur_latent;
  const int* latent_dst;
  float* latent_out;
  int n, mode;
};

__global__ void __launch_bounds__(256) ln_rows_kernel(const LnArgs a) {
  const int lane = threadIdx.x & 31;
  const int gw = blockIdx.x * 8 + (threadIdx.x >> 5), nw = gridDim.x * 8;
  for (int i = gw; i < a.n; i += nw) {
    int src = i, dst = -1;
    if (a.mode == 1) src = a.head_rows[i];
    if (a.mode == 2) {
      dst = a.latent_dst[i];
      if (dst < 0) continue;
    }
    const float* xp = a.x + (size_t)src * kD;
    float4 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = *reinterpret_cast<const float4*>(xp + 128 * j + 4 * lane);
    ln_normalise(v, a.g1, a.b1, lane);
    if (a.mode != 0) ln_normalise(v, a.g2, a.b2, lane);
    if (a.mode == 2) {
      float* lp = a.latent_out + (size_t)dst * kD;
#pragma unroll
      for (int j = 0; j < 8; ++j) *reinterpret_cast<float4*>(lp + 128 * j + 4 * lane) = v[j];
      continue;
    }
    if (a.mode == 1) {
      float* lp = a.cur_latent + (size_t)i * kD;
#pragma unroll
      for (int j = 0; j < 8; ++j) *reinterpret_cast<float4*>(lp + 128 * j + 4 * lane) = v[j];
    }
    __nv_bfloat16* op = a.xn + (size_t)i * kD;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      *reinterpret_cast<uint2*>(op + 128 * j + 4 * lane) = make_uint2(pack_bf16x2(v[j].x, v[j].y), pack_bf16x2(v[j].z, v[j].w));
  }
}

// ---- causal attention of prompt-shaped rows -> ctx (bf16 row-major) ----------------------------------------------------
// One warp per unit.  use_tiles: unit = (16-row query tile, head) on mma.m16n8k16 (ar_attn.cuh attn_tile16; tiles that
// straddle two sequences fall back to one row at a time); else unit = (row, head) as in decode (short prompts: more
// warps in flight).
__global__ void __launch_bounds__(256) dense_attn_kernel(const __grid_constant__ Params P, int layer, int use_tiles) {
  const int lane = threadIdx.x & 31;
  const int rg = lane >> 3, dc = lane & 7;
  const int gw = blockIdx.x * 8 + (threadIdx.x >> 5), nw = gridDim.x * 8;
  const int nrows = P.n_rows;
  if (use_tiles) {
    const int n_tiles = (nrows + 15) >> 4;
    for (int ut = gw; ut < n_tiles * kH; ut += nw) {
      const int tile = ut / kH, head = ut - tile * kH;
      const int r0 = 16 * tile, nq = min(16, nrows - r0);
      const int pa = P.row_pos[r0], pb = P.row_pos[r0 + nq - 1];
      if (pa >= 0 && P.row_utt[r0] == P.row_utt[r0 + nq - 1] && pb == pa + nq - 1) {
        attn_tile16(P, layer, r0, nq, head);
      } else {
        for (int row = r0; row < r0 + nq; ++row) {
          const int p1 = P.row_pos[row] + 1;
          if (p1 <= 0) continue;
          float m, l, acc[8];
          attn_unit(P, layer, row, head, 0, p1, m, l, acc);
          if (rg == 0) {
            const float inv = 1.0f / l;
#pragma unroll
            for (int i = 0; i < 8; i += 2) write_ctx_pair(P, row, head * kHD + dc * 8 + i, acc[i] * inv, acc[i + 1] * inv);
          }
        }
      }
    }
    return;
  }
  for (int u = gw; u < nrows * kH; u += nw) {
    const int row = u / kH, head = u - row * kH;
    const int p1 = P.row_pos[row] + 1;
    if (p1 <= 0) continue;
    float m, l, acc[8];
    attn_unit(P, layer, row, head, 0, p1, m, l, acc);
    if (rg == 0) {
      const float inv = 1.0f / l;
#pragma unroll
      for (int i = 0; i < 8; i += 2) write_ctx_pair(P, row, head * kHD + dc * 8 + i, acc[i] * inv, acc[i + 1] * inv);
    }
  }
}

// ---- GEMM launch -----------------------------------------------------------------------------------------------------
}  // namespace tts

using namespace tts;

// ---- handle-side state ---------------------------------------------------------------------------------------------------
int tts::dense_create(TtsArHandle* h) {
  const TtsArConfig& c = h->cfg;
  DenseW& d = h->dense;
  d.ok = false;
  if (!encode_fn()) return fail(h, TTS_AR_ERR_DEVICE, "cuTensorMapEncodeTiled is not available from this driver");
  const size_t per_layer = (size_t)kD3 * kD + (size_t)kD * kD + (size_t)kD4 * kD + (size_t)kD * kD4;
  d.per_layer = per_layer;
  const size_t head_elems = (size_t)c.n_mel_codes * kD;
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, (per_layer * c.n_layers + head_elems) * sizeof(__nv_bfloat16));
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaMalloc(K-major weight image)");
  h->allocs.push_back(p);
  d.wt = reinterpret_cast<__nv_bfloat16*>(p);
  d.wt_head = d.wt + per_layer * c.n_layers;
  d.layers.resize(c.n_layers);
  bool ok = true;
  for (int l = 0; l < c.n_layers; ++l) {
    __nv_bfloat16* b = d.wt + per_layer * l;
    DenseW::Layer& L = d.layers[l];
    L.qkv = b; L.proj = L.qkv + (size_t)kD3 * kD; L.fc = L.proj + (size_t)kD * kD; L.proj2 = L.fc + (size_t)kD4 * kD;
    ok = ok && make_map(&L.m_qkv[0], L.qkv, kD3, kD, 128) && make_map(&L.m_qkv[1], L.qkv, kD3, kD, 256);
    ok = ok && make_map(&L.m_proj[0], L.proj, kD, kD, 128) && make_map(&L.m_proj[1], L.proj, kD, kD, 256);
    ok = ok && make_map(&L.m_fc[0], L.fc, kD4, kD, 128) && make_map(&L.m_fc[1], L.fc, kD4, kD, 256);
    ok = ok && make_map(&L.m_proj2[0], L.proj2, kD, kD4, 128) && make_map(&L.m_proj2[1], L.proj2, kD, kD4, 256);
  }
  ok = ok && make_map(&d.m_head[0], d.wt_head, c.n_mel_codes, kD, 128) && make_map(&d.m_head[1], d.wt_head, c.n_mel_codes, kD, 256);
  // A operands: the bf16 activation buffers, row-major in this path
  ok = ok && make_map(&d.a_xn, h->d_xn, h->rows_cap, kD, kTcBM);
  ok = ok && make_map(&d.a_ctx, h->d_ctx, h->rows_cap, kD, kTcBM);
  ok = ok && make_map(&d.a_h, h->d_h, h->rows_cap, kD4, kTcBM);
  if (!ok) return fail(h, TTS_AR_ERR_CUDA, "cuTensorMapEncodeTiled failed");
  d.ok = true;
  return 0;
}

void tts::dense_pack_layer(TtsArHandle* h, int l, const TtsArLayerWeights& lw, int dtype, cudaStream_t st) {
  DenseW::Layer& L = h->dense.layers[l];
  launch_kmajor_pack(lw.c_attn_w, dtype, L.qkv, kD, kD3, 0, st);
  launch_kmajor_pack(lw.attn_proj_w, dtype, L.proj, kD, kD, 0, st);
  launch_kmajor_pack(lw.c_fc_w, dtype, L.fc, kD, kD4, 0, st);
  launch_kmajor_pack(lw.mlp_proj_w, dtype, L.proj2, kD4, kD, 0, st);
  h->launches += 4;
}

void tts::dense_pack_head(TtsArHandle* h, const void* mel_head_w, int dtype, cudaStream_t st) {
  launch_kmajor_pack(mel_head_w, dtype, h->dense.wt_head, kD, h->cfg.n_mel_codes, 1, st);
  h->launches += 1;
}

// Runs the layers + tail over the rows described by P (decode_mode 0: row_utt / row_pos / prefix pages are set up by
// run_prompt_pass, x holds the embedded rows).
int tts::run_dense_pass(TtsArHandle* h, const Params& Pin, cudaStream_t st) {
  DenseW& d = h->dense;
  if (!d.ok) return fail(h, TTS_AR_ERR_STATE, "dense path not initialised");
  const TtsArConfig& c = h->cfg;
  Params P = Pin;
  P.ctx_row_major = 1;
  const int n = P.n_rows, n_sm = h->G;
  const int ln_grid = std::max(1, std::min((n + 7) / 8, 4 * n_sm));
  const int n_tiles16 = (n + 15) / 16;
  const int use_tiles = h->opt_attn_tiles == 2 ? 1 : h->opt_attn_tiles == 1 ? 0 : ((n_tiles16 * kH >= n_sm * kNW) ? 1 : 0);
  const int attn_units = use_tiles ? n_tiles16 * kH : n * kH;
  const int attn_grid = std::max(1, std::min((attn_units + 7) / 8, 8 * n_sm));
  TcGemmArgs g{};
  g.M = n;
  g.watchdog = h->d_watchdog;
  g.kv = h->d_kv; g.row_utt = P.row_utt; g.row_pos = P.row_pos; g.prefix_pt = P.prefix_pt; g.n_pages = h->n_pages; g.q = h->d_q;
  LnArgs la{};
  la.x = h->d_x; la.xn = h->d_xn; la.n = n; la.mode = 0;
  cudaError_t e = cudaSuccess;
  for (int l = 0; l < c.n_layers && e == cudaSuccess; ++l) {
    const LayerDev& ld = h->layers_host[l];
    const DenseW::Layer& L = d.layers[l];
    la.g1 = ld.ln1_g; la.b1 = ld.ln1_b;
    ln_rows_kernel<<<ln_grid, 256, 0, st>>>(la);
    g.N = kD3; g.K = kD; g.bias = ld.b_qkv; g.layer = l;
    e = launch_tc_auto<TC_EPI_QKV>(d.a_xn, L.m_qkv[0], L.m_qkv[1], g, n_sm, st);
    if (e != cudaSuccess) break;
    dense_attn_kernel<<<attn_grid, 256, 0, st>>>(P, l, use_tiles);
    g.N = kD; g.K = kD; g.bias = ld.b_proj; g.out_f32 = h->d_x; g.ldo = kD;
    e = launch_tc_auto<TC_EPI_RESID>(d.a_ctx, L.m_proj[0], L.m_proj[1], g, n_sm, st);
    if (e != cudaSuccess) break;
    la.g1 = ld.ln2_g; la.b1 = ld.ln2_b;
    ln_rows_kernel<<<ln_grid, 256, 0, st>>>(la);
    g.N = kD4; g.K = kD; g.bias = ld.b_fc; g.out_bf16 = h->d_h;
    e = launch_tc_auto<TC_EPI_GELU>(d.a_xn, L.m_fc[0], L.m_fc[1], g, n_sm, st);
    if (e != cudaSuccess) break;
    g.N = kD; g.K = kD4; g.bias = ld.b_proj2; g.out_f32 = h->d_x; g.ldo = kD;
    e = launch_tc_auto<TC_EPI_RESID>(d.a_h, L.m_proj2[0], L.m_proj2[1], g, n_sm, st);
    h->launches += 7;
  }
  if (e != cudaSuccess) return cuda_fail(h, e, "dense pass kernel launch");
  LnArgs lt{};
  lt.x = h->d_x; lt.xn = h->d_xn; lt.g1 = h->d_lnf_g; lt.b1 = h->d_lnf_b; lt.g2 = h->d_fn_g; lt.b2 = h->d_fn_b;
  if (P.tail == 1) {
    lt.mode = 1; lt.n = P.n_head_rows; lt.head_rows = P.head_rows; lt.cur_latent = h->d_cur_latent;
    ln_rows_kernel<<<std::max(1, (lt.n + 7) / 8), 256, 0, st>>>(lt);
    TcGemmArgs gh{};
    gh.M = P.n_head_rows; gh.N = c.n_mel_codes; gh.K = kD; gh.bias = h->d_b_head; gh.out_f32 = h->d_logits; gh.ldo = h->v_stride;
    gh.watchdog = h->d_watchdog;
    e = launch_tc<128, TC_EPI_STORE>(d.a_xn, d.m_head[0], gh, n_sm, st);
    if (e != cudaSuccess) return cuda_fail(h, e, "mel head GEMM launch");
    h->launches += 2;
  } else if (P.tail == 2) {
    lt.mode = 2; lt.n = n; lt.latent_dst = P.latent_dst; lt.latent_out = P.latent_out;
    ln_rows_kernel<<<ln_grid, 256, 0, st>>>(lt);
    h->launches += 1;
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(h, e, "dense pass");
  return 0;
}

// Parity hook for the tensor-core GEMM alone: out[M][N] (fp32) = A[M][K] . Wt[N][K]^T + bias, bf16 operands (device
// pointers, row-major, K multiple of 64, 16-byte aligned).  bn = 128 or 256 picks the tile width.
extern "C" int tts_ar_debug_gemm(TtsArHandle* h, const void* a_bf16, const void* wt_bf16, const float* bias, float* out,
                                 int32_t M, int32_t N, int32_t K, int32_t bn, void* stream) {
  if (!h || !a_bf16 || !wt_bf16 || !out) return TTS_AR_ERR_BAD_ARG;
  if (M < 1 || N < 1 || K < kTcBK || K % kTcBK || (bn != 128 && bn != 256)) return fail(h, TTS_AR_ERR_BAD_ARG, "debug_gemm: bad shape");
  cudaError_t ce = cudaSetDevice(h->cfg.device);
  if (ce != cudaSuccess) return cuda_fail(h, ce, "cudaSetDevice");
  CUtensorMap ma, mb;
  if (!make_map(&ma, a_bf16, M, K, kTcBM) || !make_map(&mb, wt_bf16, N, K, bn))
    return fail(h, TTS_AR_ERR_CUDA, "cuTensorMapEncodeTiled failed");
  TcGemmArgs g{};
  g.M = M; g.N = N; g.K = K; g.bias = bias; g.out_f32 = out; g.ldo = N; g.watchdog = h->d_watchdog;
  cudaError_t e = (bn == 256) ? launch_tc<256, TC_EPI_STORE>(ma, mb, g, h->G, (cudaStream_t)stream)
                              : launch_tc<128, TC_EPI_STORE>(ma, mb, g, h->G, (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(h, e, "tc_gemm_kernel launch");
  h->launches += 1;
  return 0;
}

// ---- allocation-free GEMM entry point -------------------------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same tcgen05 kernel as the dense passes and
// tts_ar_debug_gemm, but the scratch a pass needs -- the bf16 row-major A operand that ln_rows_kernel writes into the
// cudaMalloc'ed h->d_xn there, and the watchdog word -- lives in a file-scope __device__ array: no handle, no allocation.
// Every GPU that loads the library holds kNoAllocScratchBytes (4 MiB + 256 B) until exit.  All calls share the array:
// calls that overlap on one GPU (two streams) race on it; serialise them on one stream.
namespace tts {

constexpr int kNoAllocHeaderBytes = 256;                        // [0] watchdog word (int), rest reserved
constexpr long long kNoAllocMaxElems = 2ll * 1024 * 1024;       // M * K <= this many bf16 elements (4 MiB)
constexpr long long kNoAllocScratchBytes = kNoAllocHeaderBytes + 2 * kNoAllocMaxElems;
constexpr int kNoAllocMaxDevices = 32;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

// In host code `g_noalloc_scratch` is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress,
// once per device with that device current.
static void* g_noalloc_ptr[kNoAllocMaxDevices] = {};
static int g_noalloc_sms[kNoAllocMaxDevices] = {};

static cudaError_t noalloc_setup(int dev, void** ptr, int* n_sm, bool* is_sm100) {
  *is_sm100 = true;
  if (!g_noalloc_ptr[dev]) {
    int major = 0, sms = 0;
    cudaError_t e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (e != cudaSuccess) return e;
    if (major != 10) {
      *is_sm100 = false;
      return cudaSuccess;
    }
    e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_noalloc_sms[dev] = sms;
    g_noalloc_ptr[dev] = p;
  }
  *ptr = g_noalloc_ptr[dev];
  *n_sm = g_noalloc_sms[dev];
  return cudaSuccess;
}

// fp32 activations -> the bf16 A operand in scratch (round to nearest even, as ln_rows_kernel and torch do); clears the
// watchdog word, which keeps the previous call's value otherwise.
__global__ void __launch_bounds__(256) noalloc_stage_kernel(const float4* __restrict__ a, uint2* __restrict__ dst, int* hdr, long long n4) {
  if (blockIdx.x == 0 && threadIdx.x == 0) hdr[0] = 0;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 v = a[i];
    dst[i] = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  }
}

}  // namespace tts

extern "C" int tts_ar_gemm_f32_noalloc(const float* a_f32, const void* wt_bf16, const float* bias, float* out, int32_t M,
                                       int32_t N, int32_t K, int32_t bn, int32_t* status_dev, void* stream) {
  if (!a_f32 || !wt_bf16 || !out) return fail(nullptr, TTS_AR_ERR_BAD_ARG, "gemm_f32_noalloc: null argument");
  if (M < 1 || N < 1 || K < kTcBK || K % kTcBK || (bn != 128 && bn != 256))
    return fail(nullptr, TTS_AR_ERR_BAD_ARG, "gemm_f32_noalloc: need M, N >= 1, K a multiple of 64, bn 128 or 256");
  if ((long long)M * K > kNoAllocMaxElems) {
    char msg[192];
    snprintf(msg, sizeof msg, "gemm_f32_noalloc: M * K = %lld exceeds the static scratch of %lld bf16 elements (4 MiB)",
             (long long)M * K, kNoAllocMaxElems);
    return fail(nullptr, TTS_AR_ERR_CAPACITY, msg);
  }
  if ((reinterpret_cast<uintptr_t>(a_f32) | reinterpret_cast<uintptr_t>(wt_bf16)) & 15)
    return fail(nullptr, TTS_AR_ERR_BAD_ARG, "gemm_f32_noalloc: a_f32 and wt_bf16 must be 16-byte aligned");
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaGetDevice");
  if (dev < 0 || dev >= kNoAllocMaxDevices) return fail(nullptr, TTS_AR_ERR_DEVICE, "gemm_f32_noalloc: device ordinal out of range");
  void* scratch = nullptr;
  int n_sm = 0;
  bool is_sm100 = true;
  e = noalloc_setup(dev, &scratch, &n_sm, &is_sm100);
  if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaGetSymbolAddress(g_noalloc_scratch)");
  if (!is_sm100) return fail(nullptr, TTS_AR_ERR_DEVICE, "gemm_f32_noalloc: needs an sm_100 (Blackwell B200) device");
  int* hdr = reinterpret_cast<int*>(scratch);
  __nv_bfloat16* a_bf16 = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<unsigned char*>(scratch) + kNoAllocHeaderBytes);
  CUtensorMap ma, mb;
  if (!make_map(&ma, a_bf16, M, K, kTcBM) || !make_map(&mb, wt_bf16, N, K, bn))
    return fail(nullptr, TTS_AR_ERR_CUDA, "gemm_f32_noalloc: cuTensorMapEncodeTiled failed");
  cudaStream_t st = (cudaStream_t)stream;
  const long long n4 = (long long)M * K / 4;
  const int grid = (int)std::max(1ll, std::min((n4 + 255) / 256, (long long)n_sm * 8));
  noalloc_stage_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const float4*>(a_f32), reinterpret_cast<uint2*>(a_bf16), hdr, n4);
  e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(nullptr, e, "noalloc_stage_kernel launch");
  TcGemmArgs g{};
  g.M = M; g.N = N; g.K = K; g.bias = bias; g.out_f32 = out; g.ldo = N; g.watchdog = hdr;
  e = (bn == 256) ? launch_tc<256, TC_EPI_STORE>(ma, mb, g, n_sm, st) : launch_tc<128, TC_EPI_STORE>(ma, mb, g, n_sm, st);
  if (e != cudaSuccess) return cuda_fail(nullptr, e, "tc_gemm_kernel launch");
  if (status_dev) {
    e = cudaMemcpyAsync(status_dev, hdr, sizeof(int), cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaMemcpyAsync(status)");
  }
  return 0;
}
