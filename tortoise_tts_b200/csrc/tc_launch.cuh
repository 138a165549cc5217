// tc_launch.cuh -- host side of the tcgen05 GEMM (tc_gemm.cuh): TMA tensor maps, the K-major bf16 weight image, launches.
// Shared by the dense passes of the AR model (dense_pass.cu) and the CLVP scorer (clvp.cu).
#pragma once
#include <algorithm>
#include "tc_gemm.cuh"

namespace tts {

// ---- tensor maps ---------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
    cudaGetLastError();
  }
  return fn;
}

// bf16 matrix [rows][K] (K contiguous); box = 64 (k) x box_rows, 128-byte swizzle; out-of-range rows read as zero
static bool make_map(CUtensorMap* m, const void* base, uint64_t rows, uint64_t K, uint32_t box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  const cuuint64_t dims[2] = {K, rows};
  const cuuint64_t strides[1] = {K * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kTcBK, box_rows};
  const cuuint32_t es[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// ---- K-major weight image: HF Conv1D [K, N] (or nn.Linear [N, K]) -> bf16 [N][K] -------------------------------------
template <typename T>
__global__ void kmajor_pack_kernel(const T* __restrict__ src, __nv_bfloat16* __restrict__ dst, int K, int N, int transposed) {
  __shared__ float tile[32][33];
  const int k0 = blockIdx.x * 32, n0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  if (transposed) {   // src already [N][K]: convert only
    for (int r = ty; r < 32; r += 8) {
      const int n = n0 + r, k = k0 + tx;
      if (n < N && k < K) dst[(size_t)n * K + k] = __float2bfloat16((float)src[(size_t)n * K + k]);
    }
    return;
  }
  for (int r = ty; r < 32; r += 8) {
    const int k = k0 + r, n = n0 + tx;
    tile[r][tx] = (k < K && n < N) ? (float)src[(size_t)k * N + n] : 0.f;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int n = n0 + r, k = k0 + tx;
    if (n < N && k < K) dst[(size_t)n * K + k] = __float2bfloat16(tile[tx][r]);
  }
}

static void launch_kmajor_pack(const void* src, int dtype, __nv_bfloat16* dst, int K, int N, int transposed, cudaStream_t st) {
  dim3 grid((K + 31) / 32, (N + 31) / 32);
  if (dtype == 0)
    kmajor_pack_kernel<float><<<grid, 256, 0, st>>>((const float*)src, dst, K, N, transposed);
  else
    kmajor_pack_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)src, dst, K, N, transposed);
}

template <int BN, int EPI>
static cudaError_t launch_tc(const CUtensorMap& ma, const CUtensorMap& mb, const TcGemmArgs& g, int n_sm, cudaStream_t st) {
  static int attr_dev_mask = 0;   // per-device opt-in (bit per device ordinal)
  int dev = 0;
  cudaGetDevice(&dev);
  if (!(attr_dev_mask & (1 << dev))) {
    cudaError_t e = cudaFuncSetAttribute(tc_gemm_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TcSmem<BN>::kBytes);
    if (e != cudaSuccess) return e;
    attr_dev_mask |= 1 << dev;
  }
  const int tiles = ((g.M + kTcBM - 1) / kTcBM) * ((g.N + BN - 1) / BN);
  const int grid = std::max(1, std::min(tiles, n_sm));
  tc_gemm_kernel<BN, EPI><<<grid, 192, TcSmem<BN>::kBytes, st>>>(ma, mb, g);
  return cudaGetLastError();
}

template <int EPI>
static cudaError_t launch_tc_auto(const CUtensorMap& a_map, const CUtensorMap& b128, const CUtensorMap& b256, const TcGemmArgs& g,
                                  int n_sm, cudaStream_t st) {
  // wide tiles once there are enough of them to fill the machine; narrow ones for short prompts, and when the wide tiling
  // quantises badly over the SMs (CLVP: 63 x 3 tiles of N = 768 are 1.28 waves; 63 x 6 narrow ones 2.55 at ~0.55 of the time each)
  const int m_tiles = (g.M + kTcBM - 1) / kTcBM;
  const int t256 = m_tiles * ((g.N + 255) / 256), t128 = m_tiles * ((g.N + 127) / 128);
  const int w256 = (t256 + n_sm - 1) / n_sm, w128 = (t128 + n_sm - 1) / n_sm;
  if (t256 >= n_sm && 100 * w256 <= 55 * w128) return launch_tc<256, EPI>(a_map, b256, g, n_sm, st);
  return launch_tc<128, EPI>(a_map, b128, g, n_sm, st);
}

}  // namespace tts
