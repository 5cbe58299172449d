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
  int n_tiles;           // N tiles of NT columns (set by launch_gemm_tc)
  int tpc;               // N tiles per CTA (1..4, UMMA N = tpc*NT <= 256): the A k-block is read once for tpc tiles
  int n_groups;          // ceil(n_tiles / tpc); blockIdx.x = m_tile * n_groups + group
  int stages;            // pipeline stages actually used (<= STAGES)
};

template <int NT, int STAGES>
struct GemmSmem {
  static constexpr int B_TILE_BYTES = NT * 128;
  static constexpr int SMEM_BUDGET = 200 * 1024;
  static constexpr int BAR_BYTES = (2 * STAGES + 1) * 8 + 16;
  __host__ __device__ static constexpr int stage_bytes(int tpc) { return GEMM_A_STAGE_BYTES + tpc * B_TILE_BYTES; }
  __host__ __device__ static constexpr int stages(int tpc) {
    return SMEM_BUDGET / stage_bytes(tpc) < STAGES ? SMEM_BUDGET / stage_bytes(tpc) : STAGES;
  }
  __host__ __device__ static constexpr int dyn_bytes(int tpc) {  // + slack for manual 1024-byte alignment
    return stages(tpc) * stage_bytes(tpc) + BAR_BYTES + 1024;
  }
};

// ALoader:  __device__ const void* src(int tile_row /*0..127*/, int m_tile, int kb, int chunk /*0..7*/, bool& valid)
//           returns the global address of 16 bytes; valid=false -> zero fill.  rows_to_load(m_tile) bounds the gather.
// Epilogue: __device__ void operator()(int tile_row, bool row_valid, int m_tile, int n_tile, float* acc /*NT*/) const
//           (called by all 128 epilogue threads so that it may use warp shuffles)
template <int NT, int STAGES, int AB_FMT, int MINB, class ALoader, class Epilogue>
__global__ void __launch_bounds__(GEMM_THREADS, MINB) gemm_tc_kernel(GemmCore g, ALoader aload, Epilogue epi) {
  using SM = GemmSmem<NT, STAGES>;
  static_assert(NT % 16 == 0 && NT >= 16 && NT <= 256, "UMMA N for M=128");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int stages = g.stages, stage_bytes = SM::stage_bytes(g.tpc);
  const int lag = stages >= 4 ? 2 : 1;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + stages * stage_bytes);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* accum_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  __shared__ int s_is_last;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m_tile = blockIdx.x / g.n_groups, group = blockIdx.x - m_tile * g.n_groups, split = blockIdx.y;
  const int n_tile0 = group * g.tpc;
  const int tiles_here = min(g.tpc, g.n_tiles - n_tile0);
  const int n_cols = tiles_here * NT;  // UMMA N of this CTA
  const uint32_t tmem_cols = n_cols <= 32 ? 32 : n_cols <= 64 ? 64 : n_cols <= 128 ? 128 : 256;
  const int kb0 = split * g.kb_per_split;
  const int kb1 = min(g.k_blocks, kb0 + g.kb_per_split);
  const int nkb = kb1 - kb0;

  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 128 + 1);  // 128 A-producer threads + 1 expect_tx arrive for the weight TMA
      mbar_init(&empty_bar[s], 1);       // one tcgen05.commit
    }
    mbar_init(accum_bar, 1);
    fence_barrier_init();
  }
  if (warp == 4) tmem_alloc(tmem_slot, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 4) {
    // ------------------------------------------------------------------ A producers (cp.async gather -> SW128 tile)
    const int rows = aload.rows(m_tile, g.m_valid);
    int arrived = 0;
    for (int i = 0; i < nkb; ++i) {
      const int s = i % stages;
      mbar_wait(&empty_bar[s], ((i / stages) & 1) ^ 1);
      uint8_t* a_tile = smem + s * stage_bytes;
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
      if (i >= lag) {
        if (lag == 2) cp_async_wait<2>(); else cp_async_wait<1>();
        fence_proxy_async();
        mbar_arrive(&full_bar[arrived % stages]);
        ++arrived;
      }
    }
    cp_async_wait<0>();
    fence_proxy_async();
    for (; arrived < nkb; ++arrived) mbar_arrive(&full_bar[arrived % stages]);

    // ------------------------------------------------------------------ epilogue (same 128 threads: row == TMEM lane)
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    const int row = tid;  // tile row
    const bool row_valid = row < rows;
    const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16);
    for (int jt = 0; jt < tiles_here; ++jt) {
      const int n_tile = n_tile0 + jt;
      float acc[NT];
#pragma unroll
      for (int c = 0; c < NT; c += 32) tmem_ld32(taddr + jt * NT + c, acc + c);
      tmem_ld_wait();

      if (g.splits > 1) {  // (tpc == 1 whenever K is split)
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
    }
    tc_fence_before();
  } else if (warp == 4) {
    // ------------------------------------------------------------------ weight producer: one bulk copy per tile and stage
    // (lane jt copies tile jt: a single thread only gets a few bulk copies per microsecond through the TMA unit)
    if (lane < 4) {
      const uint64_t pol = l2_policy_evict_first();
      const uint8_t* wbase = reinterpret_cast<const uint8_t*>(g.Wp);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % stages;
        mbar_wait(&empty_bar[s], ((i / stages) & 1) ^ 1);
        if (lane == 0) mbar_arrive_expect_tx(&full_bar[s], tiles_here * SM::B_TILE_BYTES);
        if (lane < tiles_here) {
          uint8_t* dst = smem + s * stage_bytes + GEMM_A_STAGE_BYTES + lane * SM::B_TILE_BYTES;
          const uint8_t* src = wbase + ((size_t)(n_tile0 + lane) * g.k_blocks + kb0 + i) * SM::B_TILE_BYTES;
          if (g.stream_weights)
            bulk_g2s_hint(dst, src, SM::B_TILE_BYTES, &full_bar[s], pol);
          else
            bulk_g2s(dst, src, SM::B_TILE_BYTES, &full_bar[s]);
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ MMA issuer
    // The warp stays converged (uniform control flow keeps descriptors in uniform registers and the issue sequence
    // short); one elected lane issues the MMAs and the commits.
    {
      const uint32_t idesc = umma_idesc(GEMM_BM, n_cols, AB_FMT);
      const bool leader = elect_one();
      for (int i = 0; i < nkb; ++i) {
        const int s = i % stages;
        mbar_wait(&full_bar[s], (i / stages) & 1);
        tc_fence_after();
        if (leader) {
          const uint32_t a_addr = smem_u32(smem + s * stage_bytes);
          const uint64_t adesc = umma_desc_sw128(a_addr);
          const uint64_t bdesc = umma_desc_sw128(a_addr + GEMM_A_STAGE_BYTES);
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k)  // +32 bytes per K=16 step inside the 128-byte swizzle row
            umma_f16(tmem_base, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (i | k) != 0);
          umma_commit(&empty_bar[s]);
        }
        __syncwarp();
      }
      if (leader) umma_commit(accum_bar);
      __syncwarp();
    }
  }
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
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
// Tiles per CTA: as many N tiles as one UMMA takes (N <= 256) when there are several M tiles to keep the SMs busy --
// the A k-block (16 KB through the L2 fabric) is then amortised over up to 4x the weight bytes and flops.
template <int NT>
inline int gemm_tc_tpc(int n_tiles, int m_tiles, int splits, int k_blocks) {
  if (splits > 1 || m_tiles < 2) return 1;
  if (k_blocks <= 4) return 1;  // short K: epilogue / latency bound -> small stages, several CTAs per SM instead
  const int max_tpc = 256 / NT < 4 ? 256 / NT : 4;
  const int groups = (n_tiles + max_tpc - 1) / max_tpc;
  return (n_tiles + groups - 1) / groups;
}

template <int NT, int STAGES, int AB_FMT, int MINB = 1, class ALoader, class Epilogue>
cudaError_t launch_gemm_tc(const GemmCore& g_in, const ALoader& al, const Epilogue& epi, int n_tiles, int m_tiles,
                           cudaStream_t stream) {
  GemmCore g = g_in;
  using SM = GemmSmem<NT, STAGES>;
  g.n_tiles = n_tiles;
  g.tpc = gemm_tc_tpc<NT>(n_tiles, m_tiles, g.splits, g.k_blocks);
  g.n_groups = (n_tiles + g.tpc - 1) / g.tpc;
  g.stages = SM::stages(g.tpc);
  const int kb_cta = g.kb_per_split < g.k_blocks ? g.kb_per_split : g.k_blocks;
  if (g.stages > kb_cta) g.stages = kb_cta < 2 ? 2 : kb_cta;  // no more stages than k-blocks: leaves room for co-resident CTAs
  auto kern = gemm_tc_kernel<NT, STAGES, AB_FMT, MINB, ALoader, Epilogue>;
  static bool configured = false;  // per template instantiation
  if (!configured) {
    int max_dyn = 0;
    for (int t = 1; t <= 4 && t * NT <= 256; ++t) max_dyn = SM::dyn_bytes(t) > max_dyn ? SM::dyn_bytes(t) : max_dyn;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, max_dyn);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  dim3 grid((unsigned)g.n_groups * (unsigned)m_tiles, g.splits, 1);
  kern<<<grid, GEMM_THREADS, g.stages * SM::stage_bytes(g.tpc) + SM::BAR_BYTES + 1024, stream>>>(g, al, epi);
  return cudaGetLastError();
}

}  // namespace zb
