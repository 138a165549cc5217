// layout.h -- index arithmetic shared by the CUDA kernels, the host runtime and the CPU unit tests.
// Everything here is __host__ __device__ and free of CUDA types so tests/test_layout.py can compile it
// with gcc and check the packing / fragment maps without a GPU.
//
// Geometry is the reference's UnifiedVoice GPT-2 (tortoise_tts/models/unified_voice.py:335-354,
// 405-418): d_model 1024, 16 heads x 64, MLP 4096, Conv1D weights stored [in, out].
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define TTS_HD __host__ __device__ __forceinline__
#else
#define TTS_HD static inline
#endif

namespace tts {

constexpr int kD = 1024;          // model dim
constexpr int kH = 16;            // heads
constexpr int kHD = 64;           // head dim
constexpr int kD3 = 3 * kD;       // c_attn out
constexpr int kD4 = 4 * kD;       // c_fc out
constexpr int kPage = 32;         // KV positions per page
constexpr int kMaxPagesPerSeg = 32;   // page-table width per segment (32*32 = 1024 >= ctx 1012)
constexpr int kChunkK = 1024;     // k extent of one streamed weight chunk
constexpr int kChunkN = 8;        // n extent (one mma n8 tile)
constexpr int kChunkBytes = kChunkK * kChunkN * 2;   // 16 KiB of bf16
constexpr int kKbPerChunk = kChunkK / 16;            // 64 k16-blocks
constexpr int kNW = 8;            // consumer warps per CTA (8 + 1 producer = 9 warps -> 168 registers/thread, no spills)
constexpr int kConsumerThreads = kNW * 32;
constexpr int kThreads = kConsumerThreads + 32;      // + 1 producer warp
constexpr int kRowGroup = 64;     // rows handled per GEMM instruction (<= 4 M-tiles of 16)
constexpr int kARowBytes = kD * 2 + 16;              // padded bf16 row of the LN'd A operand in smem
constexpr int kPartStride = 72;   // floats per split-KV partial: m, l, pad[6], acc[64]

// ---- streamed weight chunk -------------------------------------------------------------------------
// A chunk holds W[k0 .. k0+1024) x [n0 .. n0+8) (bf16) in mma.m16n8k16 B-fragment order so a warp reads
// its fragments with conflict-free 128-bit shared loads:
//   chunk[slot s (0..31)][lane (0..31)] = uint4 { b0(kb=2s), b1(kb=2s), b0(kb=2s+1), b1(kb=2s+1) }
//   b0(kb) = { W[k0 + 16kb + 2t    ][n0 + g], W[.. + 2t + 1][n0 + g] }   g = lane >> 2, t = lane & 3
//   b1(kb) = { W[k0 + 16kb + 2t + 8][n0 + g], W[.. + 2t + 9][n0 + g] }
// Returns the bf16 element offset inside the chunk of W[k0 + k][n0 + n], k < 1024, n < 8.
TTS_HD int chunk_elem_offset(int k, int n) {
  int kb = k >> 4, kk = k & 15;
  int s = kb >> 1, odd = kb & 1;
  int t = (kk & 7) >> 1, half = kk & 1, hi = kk >> 3;   // hi: b1 vs b0
  int lane = n * 4 + t;
  int word = odd * 2 + hi;                              // 32-bit word inside the uint4
  return ((s * 32 + lane) * 4 + word) * 2 + half;
}

// Packed weight matrix [K, N] (K multiple of 1024): tile-major, chunk index = n_tile * (K/1024) + kq.
TTS_HD int64_t packed_chunk_index(int n_tile, int kq, int K) { return (int64_t)n_tile * (K / kChunkK) + kq; }

// ---- A-operand fragment layout in global memory (attention context, MLP hidden) --------------------
// mma.m16n8k16 A fragment, one uint4 per lane per (m-tile, k16-block):
//   a0 = A[g][2t..2t+1], a1 = A[g+8][2t..], a2 = A[g][2t+8..], a3 = A[g+8][2t+8..]
// Storage: [m_tile][kb (K/16)][lane][8 bf16].  Returns the bf16 element offset of A[row][col].
TTS_HD int64_t afrag_elem_offset(int row, int col, int K) {
  int mt = row >> 4, rr = row & 15;
  int kb = col >> 4, kk = col & 15;
  int lane = (rr & 7) * 4 + ((kk & 7) >> 1);
  int reg = (rr >> 3) + 2 * (kk >> 3);
  return (((int64_t)mt * (K >> 4) + kb) * 32 + lane) * 8 + reg * 2 + (kk & 1);
}

// ---- KV page pool -----------------------------------------------------------------------------------------------------
// pool[layer][head][page][kv (0=K,1=V)][slot (32)][64] bf16.  Returns the bf16 element offset of [..][0].
// Head-major: the K and V blocks of one (head, page) are adjacent (8 KiB), and consecutive pages of a sequence (the host
// allocates them consecutively) are one contiguous byte range per head -- one bulk L2 prefetch covers a whole segment.
constexpr int kKvVOffset = kPage * kHD;            // elements from a page's K block to its V block
constexpr int kKvPageStride = 2 * kPage * kHD;     // elements between consecutive pages of one (layer, head)
TTS_HD int64_t kv_row_offset(int layer, int n_pages, int page, int kv, int head, int slot) {
  return (((((int64_t)layer * kH + head) * n_pages + page) * 2 + kv) * kPage + slot) * kHD;
}

// ---- static tile schedule ---------------------------------------------------------------------------
// n_tiles n8-tiles of a GEMM are dealt to G CTAs in contiguous ranges; `rot` rotates which CTAs get the
// extra tile so that bytes per CTA even out across consecutive GEMMs of a layer.
TTS_HD void tile_range(int n_tiles, int G, int rot, int cta, int* t0, int* cnt) {
  int r = cta - rot;
  r %= G;
  if (r < 0) r += G;
  int base = n_tiles / G, extra = n_tiles % G;
  *cnt = base + (r < extra ? 1 : 0);
  *t0 = r * base + (r < extra ? r : extra);
}
TTS_HD int next_rot(int n_tiles, int G, int rot) { return (rot + n_tiles % G) % G; }

// ---- head-cluster decode kernel (ar_decode.cu) ---------------------------------------------------------
// 32 clusters x 4 CTAs (B200 co-schedules 33 clusters of 4 but only 15 of 8): cluster c owns attention head
// h = c >> 1 for the 8 rows [8*(c&1), 8*(c&1)+8).  CTA rank r computes 16 q/k/v columns of the head
// (n8-tiles h*8 + 2r, 2r+1 of the q, k and v thirds of c_attn -- the sibling cluster streams the same weights,
// normally an L2 hit), the attention of rows 8*(c&1) + 2r, +1, and output columns [256r, 256r+256) of the
// split-K attn.c_proj product  ctx[rows, 64h..64h+64) . W[64h..64h+64, :].
constexpr int kClusterSize = 4;
constexpr int kDecClusters = 2 * kH;
constexpr int kDecCtas = kClusterSize * kDecClusters;   // 128
constexpr int kDecRows = 16;                            // rows (sequences) of one m-tile of the cluster kernels
constexpr int kDecMaxRows = 64;                         // most rows one launch of them handles (MODE 4); larger batches run as groups
constexpr int kProjkChunks = kH * 8;                    // split-K chunks per layer: (head, 128-column block)
// attn.c_proj split-K chunk (16 KiB) of (head h, column block cb): W[64h + k][128cb + n], k < 64, n < 128, stored
//   chunk[tile (16)][s (2)][lane (32)] = uint4 { b0(kb=2s), b1(kb=2s), b0(kb=2s+1), b1(kb=2s+1) }
// with b0/b1 as in chunk_elem_offset.  Returns the bf16 element offset inside the chunk.
TTS_HD int projk_elem_offset(int k, int n) {
  int tile = n >> 3, g = n & 7;
  int kb = k >> 4, kk = k & 15;
  int s = kb >> 1, odd = kb & 1;
  int t = (kk & 7) >> 1, half = kk & 1, hi = kk >> 3;
  int lane = g * 4 + t;
  int word = odd * 2 + hi;
  return (((tile * 2 + s) * 32 + lane) * 4 + word) * 2 + half;
}

// monotone float -> uint key (larger float <=> larger key); -inf maps below every finite value.
TTS_HD uint32_t float_key(uint32_t bits) { return (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u); }

}  // namespace tts
