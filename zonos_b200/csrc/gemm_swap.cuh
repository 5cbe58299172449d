// Weight-streaming GEMM for M <= 8 activation rows (batch 1..4 with CFG):  C[M, N] = A[M, K] * W[N, K]^T, operands SWAPPED.
//
// With two to eight activation rows the tensor core is only there to make the per-byte instruction cost zero, and the
// orientation decides what that costs: with the activations as the 128-row UMMA operand every MMA reads 4 KB of mostly
// stale shared memory per 2 KB of weights.  Here the WEIGHTS are the M=128 operand (two adjacent packed 64-row tiles,
// 16 KB per k-block, every byte useful) and the activation rows are the N=16 operand (1 KB per k-block), so a 32 KB
// weight stage costs eight small MMAs and the accumulator is D[128 output features (TMEM lanes), 16 batch columns].
//
//   * unit = (group of two n_tiles, pair of k-blocks) = 32 KB of weights; units numbered group-major, CTA c of G owns
//     [c*U/G, (c+1)*U/G) (stream-K), fragments / tickets as in gemm_skinny.cuh;
//   * a partial fragment is only 128 features x 8 rows of fp32 (4 KB): the last contributor to arrive loads ALL
//     contributors' partials in one batch of independent loads (one L2 round trip), sums them in CTA order (bitwise
//     reproducible) and runs the fused epilogue;
//   * epilogues are "transposed": thread = output feature, acc[b] = batch row b.  Pairs needed by RoPE sit in adjacent
//     lanes (shuffle), value/gate of the gated MLP 32 lanes apart (exchanged through 4 KB of shared memory), and the
//     cond/uncond rows of classifier-free guidance in the same thread.
//   * 224 threads, 109 KB of shared memory: two CTAs per SM, the next GEMM of the step prefetches under PDL.
#pragma once
#include "common.cuh"
#include "gemm_skinny.cuh"

namespace zb {

constexpr int SWP_THREADS = 224;
constexpr int SWP_W_STAGE = 4 * SK_TILE_BYTES;  // [kk(2)][tile(2)][8 KB]
constexpr int SWP_A_STAGE = 2 * 1024;           // [kk(2)][8 rows x 128 B]
constexpr int SWP_SW = 3, SWP_SA = 4;
constexpr int SWP_A_BYTES = SWP_SA * SWP_A_STAGE, SWP_W_BYTES = SWP_SW * SWP_W_STAGE;
constexpr int SWP_SCRATCH_OFFSET = SWP_W_BYTES + SWP_A_BYTES + 1024;  // +1 KB: the N=16 operand read of the last stage
constexpr int SWP_BAR_OFFSET = SWP_SCRATCH_OFFSET + 128 * 8 * 4;     // 4 KB epilogue scratch
constexpr int SWP_DYN_BYTES = SWP_BAR_OFFSET + (2 * SWP_SW + 2 * SWP_SA + 4) * 8 + 16 + 1024;
constexpr int SWP_MAXC = 8;  // contributors summed per batch of loads
static_assert(2 * (SWP_DYN_BYTES + 1024) <= 227 * 1024, "two CTAs per SM");

struct SwapCore {
  const void* Wp;     // packed weights [n_tiles][k_blocks][8 KB] (tile 64)
  int n_tiles, k_blocks;  // k_blocks even
  int m_valid;        // activation rows, 1..8
  float* partial_ws;  // [gridDim.x][2][2][128] float4
  int* tickets;       // [ceil(n_tiles/2)] zero-initialised, self-resetting
  unsigned long long* trace;  // debug timeline (trace builds), else nullptr
  int trace_id;
  NextW next;         // weights of the GEMM that follows in the step (gemm_skinny.cuh)
};

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {  // 32 lanes x 8 columns
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}

// Epilogue: __device__ void operator()(int lane_row /*0..127*/, int tile /*packed 64-row tile*/, bool tile_valid,
//                                      int r /*row inside the tile*/, int rows, const float* acc /*[8]*/, float* scratch)
//           called by all 128 epilogue threads (it may shuffle and use `bar.sync 1, 128` with the 4 KB scratch).
template <class ALoader, class Epilogue>
__global__ void __launch_bounds__(SWP_THREADS, 2) gemm_swap_kernel(SwapCore g, ALoader aload, Epilogue epi) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* w_ring = smem;
  uint8_t* a_ring = smem + SWP_W_BYTES;
  uint64_t* full_a = reinterpret_cast<uint64_t*>(smem + SWP_BAR_OFFSET);
  uint64_t* empty_a = full_a + SWP_SA;
  uint64_t* full_w = empty_a + SWP_SA;
  uint64_t* empty_w = full_w + SWP_SW;
  uint64_t* tmem_full = empty_w + SWP_SW;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  float* s_scratch = reinterpret_cast<float*>(smem + SWP_SCRATCH_OFFSET);
  __shared__ int s_is_last;

  pdl_launch_dependents();

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) SK_STAMP(0);
  const int KB = g.k_blocks, KBU = KB / 2;
  const int n_groups = (g.n_tiles + 1) / 2;
  const int units = n_groups * KBU, G = gridDim.x, cta = blockIdx.x;
  const int u0 = sk_first_unit(cta, units, G), u1 = sk_first_unit(cta + 1, units, G);
  const int n_units = u1 - u0;

  if (tid == 0) {
    for (int s = 0; s < SWP_SA; ++s) mbar_init(&full_a[s], 32), mbar_init(&empty_a[s], 1);
    for (int s = 0; s < SWP_SW; ++s) mbar_init(&full_w[s], 1), mbar_init(&empty_w[s], 1);
    for (int b = 0; b < 2; ++b) mbar_init(&tmem_full[b], 1), mbar_init(&tmem_empty[b], 128);
    fence_barrier_init();
  }
  if (warp == 6) tmem_alloc(tmem_slot, 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (tid == 0) SK_STAMP(1);

  if (warp == 5) {
    // ------------------------------------------------------------------ weight producer (independent of prior kernels)
    // One thread gets ~2.7 bulk copies / us through the TMA unit (gemm_skinny.cuh), and a unit is four separate 8 KB blocks:
    // every ring slot has FOUR lanes, one per copy, so a unit's bytes are requested at once (the first unit of a launch lands
    // ~1 us earlier than with one lane issuing its four copies in turn).  Lane l serves slot l / 4 (units slot, slot + SW,
    // ...) and issues copy l % 4 = 2 * kk + tile.  A slot's lanes never run two ring revolutions ahead of the consumer (the
    // parity wait cannot tell 0 from 2 completions): every group simply owns one slot.
    constexpr int SWP_WLANES = 4 * SWP_SW;
    static_assert(SWP_WLANES <= 32, "four lanes per ring slot");
    if (lane < SWP_WLANES) {
      const uint64_t pol = l2_policy_evict_first();
      const uint8_t* wp = reinterpret_cast<const uint8_t*>(g.Wp);
      const int slot = lane >> 2, sub = lane & 3, kk = sub >> 1, tile = sub & 1;
      const unsigned peers = 0xFu << (slot * 4);
      for (int i = slot; i < n_units; i += SWP_SW) {
        const int u = u0 + i, grp = u / KBU, ku = u - grp * KBU;
        mbar_wait(&empty_w[slot], ((i / SWP_SW) & 1) ^ 1);
        const bool two = (2 * grp + 1) < g.n_tiles;  // the last group of an odd tile count has one tile
        if (sub == 0) mbar_arrive_expect_tx(&full_w[slot], two ? SWP_W_STAGE : SWP_W_STAGE / 2);
        __syncwarp(peers);  // the transaction count is armed before any of the unit's copies can complete
        if (tile == 0 || two)
          bulk_g2s_hint(w_ring + slot * SWP_W_STAGE + sub * SK_TILE_BYTES,
                        wp + ((size_t)(2 * grp + tile) * KB + 2 * ku + kk) * SK_TILE_BYTES, SK_TILE_BYTES, &full_w[slot], pol);
      }
      if (lane == 0) prefetch_next_weights(g.next, cta);
    }
  } else if (warp == 6) {
    // ------------------------------------------------------------------ MMA issuer: D[128 features, 16] += W * A^T
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc(GEMM_BM, 16, 1);
      int i = 0, f = 0, u = u0;
      while (u < u1) {
        const int grp = u / KBU, ku_begin = u - grp * KBU;
        const int ku_end = min(KBU, ku_begin + (u1 - u));
        const int b = f & 1;
        mbar_wait(&tmem_empty[b], ((f >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(b * 16);
        for (int ku = ku_begin; ku < ku_end; ++ku, ++i) {
          const int sa = i % SWP_SA, sw = i % SWP_SW;
          mbar_wait(&full_w[sw], (i / SWP_SW) & 1);
          if (i == 0) SK_STAMP(3);
          mbar_wait(&full_a[sa], (i / SWP_SA) & 1);
          if (i == 0) SK_STAMP(5);
          tc_fence_after();
          const uint32_t w_addr = smem_u32(w_ring + sw * SWP_W_STAGE), a_addr = smem_u32(a_ring + sa * SWP_A_STAGE);
#pragma unroll
          for (int kk = 0; kk < 2; ++kk) {
            const uint64_t wdesc = umma_desc_sw128(w_addr + kk * 2 * SK_TILE_BYTES);
            const uint64_t adesc = umma_desc_sw128(a_addr + kk * 1024);
#pragma unroll
            for (int k = 0; k < GEMM_BK / 16; ++k)
              umma_f16(tmem_d, wdesc + (uint64_t)(k * 2), adesc + (uint64_t)(k * 2), idesc,
                       (ku != ku_begin) || (kk != 0) || (k != 0));
          }
          umma_commit(&empty_a[sa]);
          umma_commit(&empty_w[sw]);
        }
        umma_commit(&tmem_full[b]);
        u += ku_end - ku_begin;
        ++f;
      }
      SK_STAMP(6);
    }
  } else if (warp == 4) {
    // ------------------------------------------------------------------ activation producer (cp.async, polling)
    pdl_wait_prior();
    if (lane == 0) SK_STAMP(4);
    const int rows = min(g.m_valid, 8);
    int issued = 0, arrived = 0;
    int ku = u0 % KBU;
    while (arrived < n_units) {
      bool can_issue = issued < n_units && (issued - arrived) < SWP_SA;
      if (can_issue) {
        uint64_t* eb = &empty_a[issued % SWP_SA];
        const uint32_t par = ((issued / SWP_SA) & 1) ^ 1;
        if (issued == arrived) mbar_wait(eb, par);
        else can_issue = mbar_test_wait(eb, par);
      }
      if (can_issue) {
        uint8_t* a_tile = a_ring + (issued % SWP_SA) * SWP_A_STAGE;
#pragma unroll
        for (int j = 0; j < 4; ++j) {  // 2 k-blocks x 8 rows x 8 chunks = 128 chunks, 4 per lane
          const int idx = j * 32 + lane;
          const int kk = idx >> 6, r = (idx >> 3) & 7, c = idx & 7;
          if (r < rows) {
            bool valid = true;
            const void* src = aload.src(r, 0, ku * 2 + kk, c, valid);
            cp_async16(a_tile + kk * 1024 + sw128_offset(r, c), src, valid ? 16u : 0u);
          }
        }
        cp_async_commit();
        ++issued;
        if (++ku == KBU) ku = 0;
      } else {
        switch (issued - arrived) {  // wait for the OLDEST pending group only
          case 1: cp_async_wait<0>(); break;
          case 2: cp_async_wait<1>(); break;
          case 3: cp_async_wait<2>(); break;
          default: cp_async_wait<3>(); break;
        }
        fence_proxy_async();
        mbar_arrive(&full_a[arrived % SWP_SA]);
        ++arrived;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps: thread == output feature
    pdl_wait_prior();
    const int rows = min(g.m_valid, 8);
    float4* ws = reinterpret_cast<float4*>(g.partial_ws);
    int f = 0, u = u0;
    while (u < u1) {
      const int grp = u / KBU, ku_begin = u - grp * KBU;
      const int ku_end = min(KBU, ku_begin + (u1 - u));
      const int b = f & 1;
      const bool complete = (ku_begin == 0 && ku_end == KBU);
      mbar_wait(&tmem_full[b], (f >> 1) & 1);
      if (tid == 0 && f == 0) SK_STAMP(7);
      tc_fence_after();
      float acc[8];
      tmem_ld8(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(b * 16), acc);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&tmem_empty[b]);
      const int tile = 2 * grp + (tid >> 6), r = tid & 63;
      const bool tile_valid = tile < g.n_tiles;
      if (complete) {
        epi(tid, tile, tile_valid, r, rows, acc, s_scratch);
      } else {
        const int slot = (u == u0) ? 0 : 1;  // a CTA has at most two partial fragments: its first and its last
        float4* mine = ws + ((size_t)cta * 2 + slot) * 256;
        mine[tid] = make_float4(acc[0], acc[1], acc[2], acc[3]);
        mine[128 + tid] = make_float4(acc[4], acc[5], acc[6], acc[7]);
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const int c_first = sk_owner(grp * KBU, units, G), c_last = sk_owner(grp * KBU + KBU - 1, units, G);
        if (tid == 0) {
          const int old = atomicAdd(&g.tickets[grp], 1);
          s_is_last = (old == c_last - c_first);
          if (s_is_last) g.tickets[grp] = 0;  // self-reset for the next launch / graph replay
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (s_is_last) {
          __threadfence();
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] = 0.f;
          for (int c0 = c_first; c0 <= c_last; c0 += SWP_MAXC) {  // all loads of a batch in flight, summed in CTA order
            float4 lo[SWP_MAXC], hi[SWP_MAXC];
#pragma unroll
            for (int j = 0; j < SWP_MAXC; ++j)
              if (c0 + j <= c_last) {
                const int sl = (sk_first_unit(c0 + j, units, G) / KBU == grp) ? 0 : 1;
                const float4* part = ws + ((size_t)(c0 + j) * 2 + sl) * 256;
                lo[j] = __ldcg(part + tid), hi[j] = __ldcg(part + 128 + tid);
              }
#pragma unroll
            for (int j = 0; j < SWP_MAXC; ++j)
              if (c0 + j <= c_last) {
                acc[0] += lo[j].x, acc[1] += lo[j].y, acc[2] += lo[j].z, acc[3] += lo[j].w;
                acc[4] += hi[j].x, acc[5] += hi[j].y, acc[6] += hi[j].z, acc[7] += hi[j].w;
              }
          }
          epi(tid, tile, tile_valid, r, rows, acc, s_scratch);
        }
      }
      u += ku_end - ku_begin;
      ++f;
    }
    if (tid == 0) SK_STAMP(8);
  }
  tc_fence_before();
  __syncthreads();
  if (tid == 0) SK_STAMP(9);
  if (warp == 6) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 32);
  }
}

template <class ALoader, class Epilogue>
cudaError_t launch_gemm_swap(const SwapCore& g, const ALoader& al, const Epilogue& epi, int grid, cudaStream_t st) {
  auto kern = gemm_swap_kernel<ALoader, Epilogue>;
  static bool configured = false;  // per template instantiation
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SWP_DYN_BYTES);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  const int units = ((g.n_tiles + 1) / 2) * (g.k_blocks / 2);
  grid = grid < 1 ? 1 : (grid > units ? units : grid);
  return launch_pdl(kern, dim3((unsigned)grid), dim3(SWP_THREADS), (size_t)SWP_DYN_BYTES, st, pdl_enabled(), g, al, epi);
}

}  // namespace zb
