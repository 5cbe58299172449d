// One-time weight repack into the shared-memory images the tcgen05 GEMM streams with 1-D TMA bulk copies.
// Layout: [n_tile][k_block][tile_n rows x 128 B], rows in the canonical SWIZZLE_128B order (common.cuh:sw128_offset).
#include "../../include/zonos_b200.h"
#include "common.cuh"

namespace zb {

struct RowMap {
  int mode, n_rows, tile_n, rows_per_head, rows_per_head_padded;
  // source row for packed row (tile j, r), or -1 for zero padding
  __host__ __device__ long long src_row(int j, int r) const {
    if (mode == 0) {
      const long long g = (long long)j * tile_n + r;
      return g < n_rows ? g : -1;
    } else if (mode == 1) {
      const int half = tile_n / 2;
      const long long g = (r < half) ? (long long)j * half + r : (long long)n_rows / 2 + (long long)j * half + (r - half);
      return g < n_rows ? g : -1;
    } else {
      const long long p = (long long)j * tile_n + r;  // padded row index over all heads
      const int head = (int)(p / rows_per_head_padded), rr = (int)(p % rows_per_head_padded);
      if (rr >= rows_per_head) return -1;
      const long long g = (long long)head * rows_per_head + rr;
      return g < n_rows ? g : -1;
    }
  }
  __host__ __device__ int n_tiles() const {
    if (mode == 0) return (n_rows + tile_n - 1) / tile_n;
    if (mode == 1) return (n_rows / 2 + tile_n / 2 - 1) / (tile_n / 2);
    return (n_rows / rows_per_head) * rows_per_head_padded / tile_n;
  }
};

// one thread per 16-byte chunk of the destination
__global__ void pack_weight_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, RowMap map, int k_blocks,
                                   long long total_chunks) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total_chunks;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i & 7);
    long long t = i >> 3;
    const int r = (int)(t % map.tile_n);
    t /= map.tile_n;
    const int kb = (int)(t % k_blocks);
    const int j = (int)(t / k_blocks);
    const long long srow = map.src_row(j, r);
    uint4 v = make_uint4(0, 0, 0, 0);
    if (srow >= 0) v = src[(srow * k_blocks + kb) * 8 + c];
    const size_t tile_base = ((size_t)j * k_blocks + kb) * (size_t)map.tile_n * 128;
    *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(dst) + tile_base + sw128_offset(r, c)) = v;
  }
}

}  // namespace zb

extern "C" int64_t zb200_packed_weight_bytes(int n_rows, int k, int tile_n, int row_mode, int rows_per_head,
                                             int rows_per_head_padded) {
  if (k <= 0 || k % 64 || tile_n <= 0 || tile_n % 16) return -1;
  zb::RowMap m{row_mode, n_rows, tile_n, rows_per_head, rows_per_head_padded};
  if (row_mode == 2 && (rows_per_head <= 0 || rows_per_head_padded % tile_n || n_rows % rows_per_head)) return -1;
  return (int64_t)m.n_tiles() * (k / 64) * tile_n * 128;
}

extern "C" int zb200_pack_weight(const void* src, void* dst, int n_rows, int k, int tile_n, int row_mode,
                                 int rows_per_head, int rows_per_head_padded, void* stream) {
  const int64_t bytes = zb200_packed_weight_bytes(n_rows, k, tile_n, row_mode, rows_per_head, rows_per_head_padded);
  ZB_REQUIRE(bytes > 0, "pack_weight: bad shape n=%d k=%d tile_n=%d mode=%d", n_rows, k, tile_n, row_mode);
  zb::RowMap m{row_mode, n_rows, tile_n, rows_per_head, rows_per_head_padded};
  const long long chunks = bytes / 16;
  const int threads = 256;
  const int blocks = (int)((chunks + threads - 1) / threads < 148 * 16 ? (chunks + threads - 1) / threads : 148 * 16);
  zb::pack_weight_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>((const uint4*)src, (uint4*)dst, m, k / 64, chunks);
  ZB_CHECK_LAUNCH("pack_weight_kernel");
  return ZB200_OK;
}
