// Persistent tcgen05 GEMM for many-row problems (the prefill: M = 2B*S rows):  C[M, N] = A[M, K] * W[N, K]^T, bf16/fp16
// inputs, fp32 accumulate, fused epilogue functor (the same functors as gemm_tc.cuh: row == TMEM lane, 64 columns).
//
// What it changes against gemm_tc.cuh (one tile per CTA, A gathered by 128 cp.async threads that then run the epilogue):
//   * one CTA per SM walks a static tile schedule (tile = m_tile * n_groups + column group, 128 x up to 256 columns);
//   * the A k-block (128 rows x 64 columns) is ONE 2-D TMA tensor copy into the SWIZZLE_128B layout (rows past M are
//     zero-filled by the TMA unit), issued by the same warp that issues the weight bulk copies;
//   * two TMEM accumulators (2 x 256 columns): the epilogue of tile i (8 warps) overlaps the MMAs of tile i+1, so a
//     K = 2048 GEMM with a heavy epilogue (fc1: SwiGLU + stores, in_proj: RoPE + KV page writes) no longer serialises
//     ~16k cycles of MMA with ~8k cycles of epilogue per tile.
//
// Warp roles (320 threads): warps 0-7 epilogue (warp w: TMEM lanes 32*(w%4).., 64-column tiles of parity w/4),
// warp 8 producer (lane 0: A tensor copy, lanes 1-4: one weight tile each), warp 9 MMA issuer (+ TMEM alloc).
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "gemm_tc.cuh"

namespace zb {

constexpr int GB_THREADS = 320;
constexpr int GB_EPI_WARPS = 8;
constexpr int GB_NT = 64;
constexpr int GB_MAX_STAGES = 6;
constexpr int GB_SMEM_BUDGET = 200 * 1024;
constexpr int GB_BAR_BYTES = (2 * GB_MAX_STAGES + 4) * 8 + 16;

struct GemmBigArgs {
  const uint8_t* Wp;  // packed weights: [n_tiles][k_blocks][64*128 bytes] (pack.cu)
  int n_tiles;        // N tiles of 64 columns
  int tpc, n_groups;  // 64-column tiles per CTA tile (<= 4), ceil(n_tiles / tpc)
  int k_blocks;       // K / 64
  int stages, stage_bytes;
  int m_valid, total_tiles;
};

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}

template <int AB_FMT, class Epilogue>
__global__ void __launch_bounds__(GB_THREADS, 1)
gemm_big_kernel(const __grid_constant__ CUtensorMap amap, const GemmBigArgs g, const Epilogue epi) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + g.stages * g.stage_bytes);
  uint64_t* empty_bar = full_bar + GB_MAX_STAGES;
  uint64_t* t_full = empty_bar + GB_MAX_STAGES;
  uint64_t* t_empty = t_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < GB_MAX_STAGES; ++s) mbar_init(&full_bar[s], 1), mbar_init(&empty_bar[s], 1);
    for (int s = 0; s < 2; ++s) mbar_init(&t_full[s], 1), mbar_init(&t_empty[s], GB_EPI_WARPS);
    fence_barrier_init();
  }
  if (warp == 9) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 8) {
    // ---------------------------------------------------------------- producer: A box + weight tiles of every k-block
    if (lane < 5) {
      if (lane == 0) asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&amap)) : "memory");
      int s = 0;
      uint32_t empty_parity = 1;
      const size_t tile_stride = (size_t)g.k_blocks * (GB_NT * 128);  // bytes between consecutive 64-column tiles
      for (int tile = blockIdx.x; tile < g.total_tiles; tile += gridDim.x) {
        const int m_tile = tile / g.n_groups, group = tile - m_tile * g.n_groups;
        const int n_tile0 = group * g.tpc;
        const int tiles_here = min(g.tpc, g.n_tiles - n_tile0);
        const uint32_t stage_tx = GEMM_A_STAGE_BYTES + tiles_here * GB_NT * 128;
        const uint8_t* wsrc = g.Wp + (size_t)(n_tile0 + lane - 1) * tile_stride;
        for (int kb = 0; kb < g.k_blocks; ++kb) {
          mbar_wait(&empty_bar[s], empty_parity);
          uint8_t* stage = smem + s * g.stage_bytes;
          if (lane == 0) {
            mbar_arrive_expect_tx(&full_bar[s], stage_tx);
            tma_load_2d(stage, &amap, kb * 64, m_tile * GEMM_BM, &full_bar[s]);
          } else if (lane <= tiles_here) {
            bulk_g2s(stage + GEMM_A_STAGE_BYTES + (lane - 1) * GB_NT * 128, wsrc + (size_t)kb * (GB_NT * 128), GB_NT * 128,
                     &full_bar[s]);
          }
          if (++s == g.stages) s = 0, empty_parity ^= 1;
        }
      }
    }
  } else if (warp == 9) {
    // ---------------------------------------------------------------- MMA issuer (converged warp, one elected lane)
    int s = 0, it = 0;
    uint32_t full_parity = 0;
    const uint32_t ring_addr = smem_u32(smem);
    const bool leader = elect_one();
    for (int tile = blockIdx.x; tile < g.total_tiles; tile += gridDim.x, ++it) {
      const int group = tile % g.n_groups;
      const int tiles_here = min(g.tpc, g.n_tiles - group * g.tpc);
      const uint32_t idesc = umma_idesc(GEMM_BM, tiles_here * GB_NT, AB_FMT);
      const int buf = it & 1;
      mbar_wait(&t_empty[buf], ((it >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + buf * 256;
      for (int kb = 0; kb < g.k_blocks; ++kb) {
        mbar_wait(&full_bar[s], full_parity);
        tc_fence_after();
        if (leader) {
          const uint32_t a_addr = ring_addr + s * g.stage_bytes;
          const uint64_t adesc = umma_desc_sw128(a_addr);
          const uint64_t bdesc = umma_desc_sw128(a_addr + GEMM_A_STAGE_BYTES);
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k)
            umma_f16(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0);
          umma_commit(&empty_bar[s]);
        }
        __syncwarp();
        if (++s == g.stages) s = 0, full_parity ^= 1;
      }
      if (leader) umma_commit(&t_full[buf]);
      __syncwarp();
    }
  } else {
    // ---------------------------------------------------------------- epilogue: TMEM -> functor (row == TMEM lane)
    const int q = warp & 3, par = warp >> 2;
    const int row = q * 32 + lane;
    int it = 0;
    for (int tile = blockIdx.x; tile < g.total_tiles; tile += gridDim.x, ++it) {
      const int m_tile = tile / g.n_groups, group = tile - m_tile * g.n_groups;
      const int n_tile0 = group * g.tpc;
      const int tiles_here = min(g.tpc, g.n_tiles - n_tile0);
      const bool row_valid = m_tile * GEMM_BM + row < g.m_valid;
      const int buf = it & 1;
      mbar_wait(&t_full[buf], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * 256 + ((uint32_t)(q * 32) << 16);
      bool released = false;
      for (int jt = par; jt < tiles_here; jt += 2) {
        float acc[GB_NT];
        tmem_ld32(taddr + jt * GB_NT, acc);
        tmem_ld32(taddr + jt * GB_NT + 32, acc + 32);
        tmem_ld_wait();
        if (jt + 2 >= tiles_here) {  // last TMEM read of this warp: hand the accumulator back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&t_empty[buf]);
          released = true;
        }
        epi(row, row_valid, m_tile, n_tile0 + jt, acc);
      }
      if (!released) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&t_empty[buf]);
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------ launch helper
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// A: row-major [m_rows, K] 16-bit elements with row pitch lda_bytes (a multiple of 16).  Returns cudaErrorNotSupported
// when the tensor map cannot be built (the caller falls back to the per-tile kernel).
template <int AB_FMT, class Epilogue>
cudaError_t launch_gemm_big(const void* A, size_t lda_bytes, int m_rows, int k, const void* Wp, int n_tiles, const Epilogue& epi,
                            int n_sms, cudaStream_t stream) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return cudaErrorNotSupported;
  CUtensorMap amap;
  const cuuint64_t gdim[2] = {(cuuint64_t)k, (cuuint64_t)m_rows};
  const cuuint64_t gstr[1] = {(cuuint64_t)lda_bytes};
  const cuuint32_t box[2] = {64, GEMM_BM};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult cr = enc(&amap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(A), gdim, gstr, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (cr != CUDA_SUCCESS) return cudaErrorNotSupported;

  GemmBigArgs g{};
  g.Wp = (const uint8_t*)Wp, g.n_tiles = n_tiles, g.k_blocks = k / 64, g.m_valid = m_rows;
  const int max_tpc = 4;
  const int groups = (n_tiles + max_tpc - 1) / max_tpc;
  g.tpc = (n_tiles + groups - 1) / groups;
  g.n_groups = (n_tiles + g.tpc - 1) / g.tpc;
  g.stage_bytes = GEMM_A_STAGE_BYTES + g.tpc * GB_NT * 128;
  g.stages = GB_SMEM_BUDGET / g.stage_bytes < GB_MAX_STAGES ? GB_SMEM_BUDGET / g.stage_bytes : GB_MAX_STAGES;
  g.total_tiles = ((m_rows + GEMM_BM - 1) / GEMM_BM) * g.n_groups;
  auto kern = gemm_big_kernel<AB_FMT, Epilogue>;
  static bool configured = false;  // per template instantiation
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GB_SMEM_BUDGET + GB_BAR_BYTES + 1024);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  const int grid = g.total_tiles < n_sms ? g.total_tiles : n_sms;
  kern<<<grid, GB_THREADS, g.stages * g.stage_bytes + GB_BAR_BYTES + 1024, stream>>>(amap, g, epi);
  return cudaGetLastError();
}

}  // namespace zb
