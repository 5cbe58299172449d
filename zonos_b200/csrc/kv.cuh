// Paged KV-cache addressing shared by the in_proj epilogue (writer) and the attention kernels (readers).
#pragma once
#include "common.cuh"

namespace zb {

// Pool: [layer][page][kv(2)][n_kv_heads][ZB200_PAGE positions][128 dims] bf16.  Inside a position row the sixteen
// 16-byte chunks are XOR-swizzled with (pos & 7), so a whole (K or V, head) page -- 16 KB, contiguous -- lands in
// shared memory through one 1-D TMA bulk copy already bank-conflict-free for ldmatrix.
// This replaces the reference's dense per-layer [B2, max_seq, 2, Hkv, D] tensor (zonos/backbone/_torch.py:96-97).
struct KvLayout {
  bf16* pool;
  const int* block_table;  // [n_seq][pages_per_seq] -> page id
  int pages_per_seq, num_pages, n_kv_heads, layer;
  __device__ __forceinline__ bf16* row(int seq, int pos, int kv, int head) const {
    const int page = block_table[seq * pages_per_seq + (pos >> 6)];
    const size_t idx = ((((size_t)layer * num_pages + page) * 2 + kv) * n_kv_heads + head) * 64 + (pos & 63);
    return pool + idx * 128;
  }
};

}  // namespace zb
