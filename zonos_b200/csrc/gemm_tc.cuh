// tcgen05 / TMEM skinny-and-tall GEMM core:  C[M, N] = A[M, K] * W[N, K]^T,  16-bit inputs, fp32 accumulate.
//
//   * A (activations) is row-major in global memory; 4 producer warps gather 128-row x 64-col k-blocks with 16-byte
//     cp.async into the canonical SWIZZLE_128B K-major layout (the gather is a functor, so the same core serves plain
//     rows, "last position" rows, cond/uncond interleave and the shifted taps of the DAC convolutions).
//   * W (weights) is pre-packed ONCE at load time (pack.cu) into the exact shared-memory image of each
//     (n-tile, k-block): one contiguous NT*128-byte chunk per pipeline stage, fetched by a single 1-D TMA bulk copy.
//   * One elected thread issues tcgen05.mma (M=128, N=NT, K=16) into a TMEM accumulator; tcgen05.commit releases
//     smem stages and signals the epilogue.
//   * Optional split-K: each split stores its fp32 partial tile to an L2-resident workspace; the last CTA to arrive
//     (ticket counter) sums the partials in fixed split order (deterministic) and runs the fused epilogue.
//
// Warp roles (192 threads): warps 0-3 = A producers, then epilogue (warp w owns TMEM lanes 32w..32w+31);
// warp 4 = weight TMA producer (+ TMEM alloc/free); warp 5 = MMA issuer.
#pragma once
#include "common.cuh"

namespace zb {

constexpr int GEMM_BM = 128;      // rows per M tile == TMEM lanes
constexpr int GEMM_BK = 64;       // 16-bit elements per k-block == one 128-byte swizzle row
constexpr int GEMM_THREADS = 192;
constexpr int GEMM_A_STAGE_BYTES = GEMM_BM * 128;

struct GemmCore {
  const void* Wp;        // packed weights: [n_tiles][k_blocks][NT*128 bytes]
  int k_blocks;          // K / 64
  int kb_per_split;      // k-blocks per split (last split may be shorter)
  int splits;            // gridDim.y
  int m_valid;           // rows of A that exist (over all m tiles)
  float* partial_ws;     // [m_tiles*n_tiles][splits][NT/4][128] float4, only when splits > 1
  int* tickets;          // [m_tiles*n_tiles], zero-initialised, self-resetting
  int stream_weights;    // 1: weights are read once per launch -> L2 evict_first
  int n_tiles;           // N tiles; blockIdx.x = m_tile * n_tiles + n_tile (set by launch_gemm_tc)
};

template <int NT, int STAGES>
struct GemmSmem {
  static constexpr int B_STAGE_BYTES = NT * 128;
  static constexpr int STAGE_BYTES = GEMM_A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + (2 * STAGES + 1) * 8 + 16;
  static constexpr int DYN_BYTES = TOTAL + 1024;  // slack for manual 1024-byte alignment
};

// ALoader:  __device__ const void* src(int tile_row /*0..127*/, int m_tile, int kb, int chunk /*0..7*/, bool& valid)
//           returns the global address of 16 bytes; valid=false -> zero fill.  rows_to_load(m_tile) bounds the gather.
// Epilogue: __device__ void operator()(int tile_row, bool row_valid, int m_tile, int n_tile, float* acc /*NT*/) const
//           (called by all 128 epilogue threads so that it may use warp shuffles)
template <int NT, int STAGES, int AB_FMT, class ALoader, class Epilogue>
__global__ void __launch_bounds__(GEMM_THREADS, 1) gemm_tc_kernel(GemmCore g, ALoader aload, Epilogue epi) {
  using SM = GemmSmem<NT, STAGES>;
  static_assert(NT % 16 == 0 && NT >= 16 && NT <= 256, "UMMA N for M=128");
  constexpr int LAG = STAGES >= 4 ? 2 : 1;
  constexpr uint32_t TMEM_COLS = NT <= 32 ? 32 : NT <= 64 ? 64 : NT <= 128 ? 128 : 256;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + SM::BAR_OFFSET);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* accum_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  __shared__ int s_is_last;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m_tile = blockIdx.x / g.n_tiles, n_tile = blockIdx.x - m_tile * g.n_tiles, split = blockIdx.y;
  const int kb0 = split * g.kb_per_split;
  const int kb1 = min(g.k_blocks, kb0 + g.kb_per_split);
  const int nkb = kb1 - kb0;

  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 128 + 1);  // 128 A-producer threads + 1 expect_tx arrive for the weight TMA
      mbar_init(&empty_bar[s], 1);       // one tcgen05.commit
    }
    mbar_init(accum_bar, 1);
    fence_barrier_init();
  }
  if (warp == 4) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 4) {
    // ------------------------------------------------------------------ A producers (cp.async gather -> SW128 tile)
    const int rows = aload.rows(m_tile, g.m_valid);
    int arrived = 0;
    for (int i = 0; i < nkb; ++i) {
      const int s = i % STAGES;
      mbar_wait(&empty_bar[s], ((i / STAGES) & 1) ^ 1);
      uint8_t* a_tile = smem + s * SM::STAGE_BYTES;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int idx = j * 128 + tid;
        const int r = idx >> 3, c = idx & 7;
        if (r < rows) {
          bool valid = true;
          const void* src = aload.src(r, m_tile, kb0 + i, c, valid);
          cp_async16(a_tile + sw128_offset(r, c), src, valid ? 16u : 0u);
        }
      }
      cp_async_commit();
      if (i >= LAG) {
        cp_async_wait<LAG>();
        fence_proxy_async();
        mbar_arrive(&full_bar[arrived % STAGES]);
        ++arrived;
      }
    }
    cp_async_wait<0>();
    fence_proxy_async();
    for (; arrived < nkb; ++arrived) mbar_arrive(&full_bar[arrived % STAGES]);

    // ------------------------------------------------------------------ epilogue (same 128 threads: row == TMEM lane)
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    float acc[NT];
    const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16);
#pragma unroll
    for (int c = 0; c < NT; c += 32) tmem_ld32(taddr + c, acc + c);
    tmem_ld_wait();
    const int row = tid;  // tile row
    const bool row_valid = row < rows;

    if (g.splits > 1) {
      const int tile_id = m_tile * g.n_tiles + n_tile;
      float4* ws = reinterpret_cast<float4*>(g.partial_ws) + (size_t)tile_id * g.splits * (NT / 4) * 128;
      if (row_valid) {
        float4* mine = ws + (size_t)split * (NT / 4) * 128;
#pragma unroll
        for (int c = 0; c < NT / 4; ++c)
          mine[c * 128 + row] = make_float4(acc[4 * c], acc[4 * c + 1], acc[4 * c + 2], acc[4 * c + 3]);
      }
      __threadfence();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (tid == 0) {
        const int old = atomicAdd(&g.tickets[tile_id], 1);
        s_is_last = (old == g.splits - 1);
        if (s_is_last) g.tickets[tile_id] = 0;  // self-reset for the next launch / graph replay
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (s_is_last) {
        __threadfence();
        if (row_valid) {
#pragma unroll
          for (int c = 0; c < NT; ++c) acc[c] = 0.f;
          for (int sp = 0; sp < g.splits; ++sp) {  // fixed order -> bitwise reproducible
            const float4* part = ws + (size_t)sp * (NT / 4) * 128;
#pragma unroll
            for (int c = 0; c < NT / 4; ++c) {
              const float4 v = __ldcg(&part[c * 128 + row]);
              acc[4 * c] += v.x, acc[4 * c + 1] += v.y, acc[4 * c + 2] += v.z, acc[4 * c + 3] += v.w;
            }
          }
        }
        epi(row, row_valid, m_tile, n_tile, acc);
      }
    } else {
      epi(row, row_valid, m_tile, n_tile, acc);
    }
    tc_fence_before();
  } else if (warp == 4) {
    // ------------------------------------------------------------------ weight producer: one bulk copy per stage
    if (lane == 0) {
      const uint64_t pol = l2_policy_evict_first();
      const uint8_t* wsrc =
          reinterpret_cast<const uint8_t*>(g.Wp) + ((size_t)n_tile * g.k_blocks + kb0) * SM::B_STAGE_BYTES;
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES;
        mbar_wait(&empty_bar[s], ((i / STAGES) & 1) ^ 1);
        mbar_arrive_expect_tx(&full_bar[s], SM::B_STAGE_BYTES);
        uint8_t* dst = smem + s * SM::STAGE_BYTES + GEMM_A_STAGE_BYTES;
        if (g.stream_weights)
          bulk_g2s_hint(dst, wsrc + (size_t)i * SM::B_STAGE_BYTES, SM::B_STAGE_BYTES, &full_bar[s], pol);
        else
          bulk_g2s(dst, wsrc + (size_t)i * SM::B_STAGE_BYTES, SM::B_STAGE_BYTES, &full_bar[s]);
      }
    }
  } else {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc(GEMM_BM, NT, AB_FMT);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES;
        mbar_wait(&full_bar[s], (i / STAGES) & 1);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + s * SM::STAGE_BYTES);
        const uint64_t adesc = umma_desc_sw128(a_addr);
        const uint64_t bdesc = umma_desc_sw128(a_addr + GEMM_A_STAGE_BYTES);
#pragma unroll
        for (int k = 0; k < GEMM_BK / 16; ++k)  // +32 bytes per K=16 step inside the 128-byte swizzle row
          umma_f16(tmem_base, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (i | k) != 0);
        umma_commit(&empty_bar[s]);
      }
      umma_commit(accum_bar);
    }
  }
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------ A loaders
// Plain row-major A [M, lda]; tile row r of m-tile z is global row z*128 + r.
struct ARowMajor {
  const uint8_t* base;
  size_t row_bytes;
  __device__ int rows(int m_tile, int m_valid) const { return min(GEMM_BM, m_valid - m_tile * GEMM_BM); }
  __device__ const void* src(int r, int m_tile, int kb, int c, bool& valid) const {
    return base + (size_t)(m_tile * GEMM_BM + r) * row_bytes + (size_t)kb * 128 + c * 16;
  }
};
// Row r of the tile is global row  map_mul * (z*128 + r) + map_add  (e.g. the last position of each sequence).
struct ARowStrided {
  const uint8_t* base;
  size_t row_bytes;
  int map_mul, map_add;
  __device__ int rows(int m_tile, int m_valid) const { return min(GEMM_BM, m_valid - m_tile * GEMM_BM); }
  __device__ const void* src(int r, int m_tile, int kb, int c, bool& valid) const {
    return base + (size_t)((m_tile * GEMM_BM + r) * map_mul + map_add) * row_bytes + (size_t)kb * 128 + c * 16;
  }
};
// CFG pairing: even tile rows are the conditional rows, odd rows the unconditional ones (global row B + u),
// so that the epilogue finds (cond, uncond) of utterance u in adjacent lanes.
struct ARowCfgPairs {
  const uint8_t* base;
  size_t row_bytes;
  int batch;  // B utterances; rows [0,B) cond, [B,2B) uncond
  __device__ int rows(int m_tile, int m_valid) const { return min(GEMM_BM, m_valid - m_tile * GEMM_BM); }
  __device__ const void* src(int r, int m_tile, int kb, int c, bool& valid) const {
    const int u = m_tile * (GEMM_BM / 2) + (r >> 1);
    const int grow = (r & 1) ? batch + u : u;
    return base + (size_t)grow * row_bytes + (size_t)kb * 128 + c * 16;
  }
};

// ------------------------------------------------------------------------------------------------ launch helper
template <int NT, int STAGES, int AB_FMT, class ALoader, class Epilogue>
cudaError_t launch_gemm_tc(const GemmCore& g_in, const ALoader& al, const Epilogue& epi, int n_tiles, int m_tiles,
                           cudaStream_t stream) {
  GemmCore g = g_in;
  g.n_tiles = n_tiles;
  using SM = GemmSmem<NT, STAGES>;
  auto kern = gemm_tc_kernel<NT, STAGES, AB_FMT, ALoader, Epilogue>;
  static bool configured = false;  // per template instantiation
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::DYN_BYTES);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  dim3 grid((unsigned)n_tiles * (unsigned)m_tiles, g.splits, 1);
  kern<<<grid, GEMM_THREADS, SM::DYN_BYTES, stream>>>(g, al, epi);
  return cudaGetLastError();
}

}  // namespace zb
