/* layout_check.c -- TEST INFRASTRUCTURE (see oracle/__init__.py): exposes the product's index arithmetic
 * (tortoise_tts_b200/csrc/layout.h, host+device inline functions) to the CPU tests, so the packing,
 * fragment and page maps can be checked for bijectivity / coverage without a GPU.  Built by
 * __graft_entry__.build() into oracle/_ref/liblayout_check.so (compiled as C++: layout.h uses constexpr). */
#include "layout.h"

extern "C" {
int lc_chunk_elem_offset(int k, int n) { return tts::chunk_elem_offset(k, n); }
long long lc_afrag_elem_offset(int row, int col, int K) { return tts::afrag_elem_offset(row, col, K); }
long long lc_kv_row_offset(int layer, int n_pages, int page, int kv, int head, int slot) {
  return tts::kv_row_offset(layer, n_pages, page, kv, head, slot);
}
void lc_tile_range(int n_tiles, int G, int rot, int cta, int* t0, int* cnt) { tts::tile_range(n_tiles, G, rot, cta, t0, cnt); }
int lc_next_rot(int n_tiles, int G, int rot) { return tts::next_rot(n_tiles, G, rot); }
unsigned lc_float_key(unsigned bits) { return tts::float_key(bits); }
int lc_projk_elem_offset(int k, int n) { return tts::projk_elem_offset(k, n); }
int lc_chunk_bytes(void) { return tts::kChunkBytes; }
}
