// Weight-streaming GEMM of the decode step:  C[M <= 128, N] = A[M, K] * W[N, K]^T  (bf16 in, fp32 accumulate).
//
// At M = 2B <= 128 rows every projection of the decode step is bound by reading W once from HBM (intensity ~ M flop/B,
// far below the tensor ridge), so the kernel is built around ONE thing: every SM streams an equal, contiguous range of
// the packed weights with many bytes in flight, and nothing on the per-stage path waits for a round trip.
//
//   * Work unit = one 16 KB shared-memory stage of weights:
//       MODE_K2: two consecutive k-blocks of one 64-row n_tile (one contiguous 16 KB bulk copy)      -- M <= 32
//       MODE_T2: one k-block of two adjacent n_tiles (UMMA N = 128; the A k-block is reused twice)    -- M  > 32
//       MODE_T2K2: two k-blocks of two adjacent n_tiles (32 KB) -- M > 32, long streams (fc1, fc2, the hybrid's in_proj):
//                 a unit costs the MMA issuer and the activation producers ~0.2 us of barrier round trips besides its
//                 MMAs (per-unit in-kernel trace), so at 8-16 KB per unit that overhead, not HBM, set the pace
//       MODE_K1: one k-block of one n_tile (8 KB; fallback when K/64 is odd or the tile count is odd)
//       MODE_N16: eight consecutive k-blocks of one 16-row n_tile (weights packed with tile_n = 16)    -- M <= 8:
//                 N / 16 >= 128 independent column slices, every CTA runs the whole K of its slice, so no split-K
//                 fix-up, no partial sums and the fused epilogue (RoPE / residual) runs straight from TMEM
//     Units are numbered group-major (group = n_tile or n_tile pair) and CTA c of G owns units [c*U/G, (c+1)*U/G):
//     "stream-K" with a static, balanced partition.
//   * weight producer warp (one lane): 1-D TMA bulk copies (UBLKCP, evict-first) into a ring.  It does not depend on
//     the previous kernel, so under programmatic dependent launch it streams while the previous kernels still drain.
//   * A producer warp(s): cp.async gather of the A k-block(s) (functor, as in gemm_tc.cuh) into a short ring.  The loop
//     polls: it issues a stage whenever one is free and otherwise retires the oldest landed stage, so the round trip
//     "MMA done -> stage free -> load -> landed" is off the critical path.  Only A_ROWS (8 / 32 / 64 / 128) rows are staged;
//     the tensor core still reads a 128-row operand, the rows past A_ROWS are whatever follows in shared memory and
//     only ever reach TMEM lanes that nobody reads.
//   * MMA warp (one lane): tcgen05.mma M=128, two TMEM accumulator sets (ping-pong per fragment).
//   * warps 0-3 = epilogue: a CTA's range is cut into fragments (runs of units of one group).  A fragment that covers
//     the whole K runs the fused epilogue directly; otherwise its fp32 partial goes to the CTA's workspace slot and the
//     LAST contributor to arrive (ticket) sums all contributors in CTA order (bitwise reproducible) and runs the
//     epilogue.  Nobody waits for anybody: no co-residency requirement.
//   * M <= 32: 224 threads, <= 112 KB shared memory -> two CTAs per SM, i.e. the next GEMM of the step (launched early
//     by PDL) prefetches its weights next to the running one.
#pragma once
#include <cstdlib>

#include "common.cuh"
#include "gemm_tc.cuh"

namespace zb {

constexpr int SK_NT = 64;                         // weight rows of one packed tile
constexpr int SK_TILE_BYTES = SK_NT * 128;        // 8 KB packed (n_tile, k_block) image
constexpr int SK_PARTIAL_F4 = (SK_NT / 4) * 128;  // float4 per partial sub-tile (32 KB)
enum { MODE_K2 = 0, MODE_T2 = 1, MODE_K1 = 2, MODE_N16 = 3, MODE_T2K2 = 4 };
#ifndef SK_WLANES
#define SK_WLANES 4     // lanes of the weight-producer warp that issue bulk copies
#endif
#ifndef SK_SA128
#define SK_SA128 4      // activation ring depth at 128 rows (16 KB stages)
#endif
#ifndef SK_SW128_KB
#define SK_SW128_KB 128  // weight ring at 128 rows
#endif
#ifndef SK_SA128_T2K2
#define SK_SA128_T2K2 3  // activation ring depth at 128 rows with 32 KB stages (two k-blocks per unit)
#endif
#ifndef SK_T2K2_MIN_UNITS
#define SK_T2K2_MIN_UNITS 512  // 16 KB units of a GEMM (tile pairs x k-blocks) from which 32 KB units are used (measured: every
                               // projection of the decode step gains, out_proj with 512 included; 3.02 -> 2.97 ms at batch 64)
#endif
#ifndef SK_A_SLOW_START
#define SK_A_SLOW_START 1  // activation stages in flight before the first one has landed
#define SK_A_SLOW_GROW 2   // additional stages allowed per landed stage
#endif
#ifndef SK_SA64
#define SK_SA64 8       // activation ring depth at 33..64 rows (8 KB stages: the same 64 KB in flight as at 128 rows)
#endif

// Optional in-kernel timeline (build with ZB200_TRACE=1): globaltimer stamps per launch / CTA / event.
#ifdef SK_TRACE
constexpr int SKT_LAUNCHES = 64, SKT_CTAS = 160, SKT_EVENTS = 12;
__device__ __forceinline__ void sk_stamp(unsigned long long* trace, int launch, int ev) {
  if (trace && launch >= 0 && launch < SKT_LAUNCHES && blockIdx.x < SKT_CTAS) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    trace[((size_t)launch * SKT_CTAS + blockIdx.x) * SKT_EVENTS + ev] = t;
  }
}
#define SK_STAMP(ev) sk_stamp(g.trace, g.trace_id, ev)
// per-unit stamps of CTA 0 (launches 0..31): slot 32 + id, flat [unit < 320][6] = {w issued, w wait begins, w ready, a ready, MMAs issued, commits issued}
__device__ __forceinline__ void sk_ustamp(unsigned long long* trace, int launch, int unit, int what) {
  if (trace && launch >= 0 && launch < 32 && blockIdx.x == 0 && unit < 320) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    trace[(size_t)(32 + launch) * SKT_CTAS * SKT_EVENTS + unit * 6 + what] = t;
  }
}
#define SK_USTAMP(unit, what) sk_ustamp(g.trace, g.trace_id, unit, what)
#else
#define SK_STAMP(ev)
#define SK_USTAMP(unit, what)
#endif

// Epilogues with `static constexpr bool kDeferred = true` take every fragment raw through
//   partial(row, row_valid, n_tile, split, acc)        split = CTA index % n_splits (grid = n_groups * n_splits)
// and leave the reduction over splits to their consumer kernel (resid_ln_kernel): no workspace slots, no tickets.
template <class E, class = void>
struct sk_is_deferred { static constexpr bool value = false; };
template <class E>
struct sk_is_deferred<E, decltype((void)E::kDeferred)> { static constexpr bool value = E::kDeferred; };

// The GEMM that follows this one in the step (weights do not depend on anything): when a CTA has issued its last
// weight copy, its queue to HBM runs dry while the MMAs / epilogue / exit / next prologue go by.  It fills that bubble by
// pulling the first bytes that CTA `blockIdx.x` of the NEXT kernel will ask for into L2.  Both weight-streaming
// kernels number their work the same way: groups of `tg` tiles, units of `kg` k-blocks, CTA c owns units
// [c*U/G, (c+1)*U/G), and each tile of a group is contiguous along k in the packed layout.
struct NextW {
  const uint8_t* wp;  // nullptr: nothing to prefetch
  int n_tiles, k_blocks, grid, tg, kg;
  int bytes;          // per CTA
};
__host__ __device__ __forceinline__ int sk_first_unit(int c, int units, int grid);
__device__ __forceinline__ void prefetch_next_weights(const NextW& nx, int cta) {
  if (nx.wp == nullptr || cta >= nx.grid) return;
  const int KB = nx.k_blocks, KBU = KB / nx.kg, groups = (nx.n_tiles + nx.tg - 1) / nx.tg, units = groups * KBU;
  const int u0 = sk_first_unit(cta, units, nx.grid), u1 = sk_first_unit(cta + 1, units, nx.grid);
  const int grp = u0 / KBU, ku = u0 - grp * KBU;
  const int nu = min(u1 - u0, KBU - ku);  // units of the CTA's first fragment
  int per_tile = min(nu * nx.kg * 8192, nx.bytes / nx.tg);
  for (int j = 0; j < nx.tg; ++j) {
    const int tile = grp * nx.tg + j;
    if (tile >= nx.n_tiles) break;
    const uint8_t* p = nx.wp + ((size_t)tile * KB + (size_t)ku * nx.kg) * 8192;
    for (int off = 0; off < per_tile; off += 32768) bulk_prefetch_l2(p + off, (uint32_t)min(32768, per_tile - off));
  }
}

// KV pages of the NEXT attention kernel, pulled into L2 by the weight-producer lane once its last weight copy is out: the
// CTA's HBM queue would otherwise run dry through the MMAs' tail, the epilogue and the next kernel's ramp, and the decode
// attention is bound by exactly that HBM stream.  Item i = (sequence i / pages, page i % pages) of the layer's pages in use;
// CTA c takes items first_item + c * items_per_cta ...  A page is contiguous ([K|V][kv head][64][128] bf16).
struct KvPrefetch {
  const uint8_t* layer_base;  // pool + layer * num_pages * page_bytes; nullptr: nothing to prefetch
  const int* block_table;     // [n_seq][pages_per_seq]
  const int* pos_ptr;         // device-side position of the frame
  int pos_add, pages_per_seq, n_seq, page_bytes, first_item, items_per_cta;
};
__device__ __forceinline__ void prefetch_kv(const KvPrefetch& k, int cta) {
  if (k.layer_base == nullptr || k.items_per_cta <= 0) return;
  const int n_pages = min(k.pages_per_seq, (*k.pos_ptr + k.pos_add) / 64 + 1);
  const int total = k.n_seq * n_pages;
  for (int j = 0; j < k.items_per_cta; ++j) {
    const int item = k.first_item + cta * k.items_per_cta + j;
    if (item >= total) return;
    const int seq = item / n_pages, pg = item - seq * n_pages;
    const uint8_t* p = k.layer_base + (size_t)k.block_table[seq * k.pages_per_seq + pg] * k.page_bytes;
    for (int off = 0; off < k.page_bytes; off += 32768) bulk_prefetch_l2(p + off, (uint32_t)min(32768, k.page_bytes - off));
  }
}

struct SkinnyCore {
  const void* Wp;     // packed weights [n_tiles][k_blocks][8 KB]
  int n_tiles, k_blocks;
  int m_valid;        // rows of A (<= 128)
  float* partial_ws;  // [gridDim.x][2][TG][SK_PARTIAL_F4] float4
  int* tickets;       // [n_groups] zero-initialised, self-resetting
  int n_splits;       // deferred epilogues: CTAs per group (grid = n_groups * n_splits: every CTA is one fragment)
  unsigned long long* trace;  // debug timeline [64][160][12] (trace builds), else nullptr
  int trace_id;               // launch slot of the debug timeline
  NextW next;                 // weights of the GEMM that follows in the step
  KvPrefetch kvpf;            // KV pages of the attention kernel that follows in the step
};

template <int A_ROWS, int MODE>
struct SkinnyCfg {
  static constexpr int NTILE = (MODE == MODE_N16) ? 16 : SK_NT;  // weight rows of one packed tile
  static constexpr int TILE_BYTES = NTILE * 128;
  static constexpr int PARTIAL_F4 = (NTILE / 4) * 128;
  static constexpr int TG = (MODE == MODE_T2 || MODE == MODE_T2K2) ? 2 : 1;  // n_tiles per unit
  static constexpr int KG = (MODE == MODE_K2 || MODE == MODE_T2K2) ? 2 : (MODE == MODE_N16) ? 8 : 1;  // k-blocks per unit
  static_assert(MODE != MODE_N16 || A_ROWS <= 8, "16-column slices are the M <= 8 path");
  static constexpr bool BIG = A_ROWS >= 64;  // one CTA per SM: four activation-producer warps, 128 KB weight ring
  static constexpr int PROD_WARPS = BIG ? 4 : 1;
  static constexpr int PROD_THREADS = PROD_WARPS * 32;
  static constexpr int THREADS = (4 + PROD_WARPS + 2) * 32;
  static constexpr int MIN_BLOCKS = BIG ? 1 : 2;
  static constexpr int A_SUB_BYTES = A_ROWS * 128;
  static constexpr int A_STAGE_BYTES = KG * A_SUB_BYTES;
  static constexpr int W_STAGE_BYTES = TG * KG * TILE_BYTES;
  // The activation gather is a round trip to L2 per stage (~1.2 us under load), so its rate is ring bytes / latency: at
  // 64 rows a stage holds 8 KB and the ring must be twice as deep as at 128 rows to feed the same weight rate.
  static constexpr int SA = (MODE == MODE_T2K2) ? (A_ROWS >= 128 ? SK_SA128_T2K2 : BIG ? SK_SA64 / 2 : 4)
                            : (A_ROWS >= 128) ? SK_SA128 : BIG ? SK_SA64 : (A_ROWS <= 8 && MODE != MODE_N16) ? 8 : 4;
  static constexpr int SW = BIG ? (SK_SW128_KB * 1024 / W_STAGE_BYTES)
                                            : ((A_ROWS > 8 || MODE == MODE_N16 ? 64 : 80) * 1024 / W_STAGE_BYTES);
  static constexpr int A_BYTES = SA * A_STAGE_BYTES;
  static constexpr int W_BYTES = SW * W_STAGE_BYTES;
  static constexpr int BAR_OFFSET = A_BYTES + W_BYTES;
  static constexpr int N_BARS = 2 * SA + 2 * SW + 4;
  static constexpr int TOTAL = BAR_OFFSET + N_BARS * 8 + 16;
  static constexpr int DYN_BYTES = TOTAL + 1024;
  static constexpr uint32_t TMEM_COLS = 2 * TG * NTILE;  // 32, 128 or 256
  static_assert(W_BYTES >= GEMM_BM * 128, "the 128-row operand read must stay inside the allocation");
  static_assert(MIN_BLOCKS == 1 || DYN_BYTES <= 113 * 1024, "two CTAs per SM need <= 113 KB each");
};

// first unit of CTA c, and the CTA that owns unit u (inverse of the former)
__host__ __device__ __forceinline__ int sk_first_unit(int c, int units, int grid) { return (int)((long long)c * units / grid); }
__host__ __device__ __forceinline__ int sk_owner(int u, int units, int grid) {
  return (int)((((long long)u + 1) * grid + units - 1) / units) - 1;
}

// Order of the units inside a fragment.  Every CTA that walks the same k range reads the SAME activation lines from L2
// at the same moment (with one whole-K fragment per CTA -- fc1 at batch 64, the hybrid's in_proj -- that is every CTA of
// the grid on 64-128 lines: a handful of L2 slices serve all SMs while the rest idle, and the gather, not HBM, sets the
// pace).  So CTA c starts its fragment `rot` units in and wraps around: unit i of the CTA (u = u0 + i in memory order)
// is processed as k-unit  begin + (u - fragment start + rot) % length.  The weight producer, the activation producers
// and the MMA issuer all walk this order; the accumulation order of a tile therefore depends on the CTA index (fixed for
// a given shape and grid: results stay bitwise reproducible).
#ifndef SK_ROTATE
#define SK_ROTATE 0
#endif
__device__ __forceinline__ int sk_unit_ku(int u, int u0, int u1, int KBU, int cta, int& grp) {
  grp = u / KBU;
  const int g0 = grp * KBU;
#if SK_ROTATE
  const int fs = max(u0, g0), fe = min(u1, g0 + KBU), len = fe - fs;
  return fs - g0 + (u - fs + cta % len) % len;
#else
  (void)u0, (void)u1, (void)cta;
  return u - g0;
#endif
}

__device__ __forceinline__ bool mbar_test_wait(uint64_t* bar, uint32_t parity) {  // non-blocking
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

template <int A_ROWS, int MODE, class ALoader, class Epilogue>
__global__ void __launch_bounds__(SkinnyCfg<A_ROWS, MODE>::THREADS, SkinnyCfg<A_ROWS, MODE>::MIN_BLOCKS)
    gemm_skinny_kernel(SkinnyCore g, ALoader aload, Epilogue epi) {
  using C = SkinnyCfg<A_ROWS, MODE>;
  constexpr int TG = C::TG, KG = C::KG, SA = C::SA, SW = C::SW;
  constexpr int W_WARP = 4 + C::PROD_WARPS, MMA_WARP = W_WARP + 1;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_ring = smem;
  uint8_t* w_ring = smem + C::A_BYTES;
  uint64_t* full_a = reinterpret_cast<uint64_t*>(smem + C::BAR_OFFSET);
  uint64_t* empty_a = full_a + SA;
  uint64_t* full_w = empty_a + SA;
  uint64_t* empty_w = full_w + SW;
  uint64_t* tmem_full = empty_w + SW;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  __shared__ int s_is_last;

  pdl_launch_dependents();  // the next kernel of the step may start its own prologue / weight prefetch

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) SK_STAMP(0);
  const int KB = g.k_blocks;
  const int KBU = KB / KG;                 // units per group
  const int n_groups = g.n_tiles / TG;
  const int units = n_groups * KBU, G = gridDim.x, cta = blockIdx.x;
  const int u0 = sk_first_unit(cta, units, G), u1 = sk_first_unit(cta + 1, units, G);
  const int n_units = u1 - u0;

  if (tid == 0) {
    for (int s = 0; s < SA; ++s) {
      mbar_init(&full_a[s], C::PROD_THREADS);
      mbar_init(&empty_a[s], 1);
    }
    for (int s = 0; s < SW; ++s) {
      mbar_init(&full_w[s], 1);
      mbar_init(&empty_w[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full[b], 1);
      mbar_init(&tmem_empty[b], 128);
    }
    fence_barrier_init();
  }
  if (warp == MMA_WARP) tmem_alloc(tmem_slot, C::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (tid == 0) SK_STAMP(1);

  if (warp == W_WARP) {
    // ------------------------------------------------------------------ weight producer (independent of prior kernels)
    // One thread gets only ~2.7 bulk copies per microsecond through the TMA unit whatever their size
    // (scripts/exp_stream.cu: 22 GB/s with 8 KB copies, 44 GB/s with 16 KB), less than an SM's share of HBM.  SK_WLANES
    // lanes therefore issue in parallel, lane l owning units l, l + SK_WLANES, ...  A lane must never run two ring
    // revolutions ahead of the consumer (the parity wait cannot tell 0 from 2 completions): lanes <= slots guarantees it.
    static_assert(SK_WLANES / (TG == 2 ? 2 * KG : 1) <= SW, "units issued per round must not outnumber the ring slots");
    if (lane < SK_WLANES) {
      const uint64_t pol = l2_policy_evict_first();
      const uint8_t* wp = reinterpret_cast<const uint8_t*>(g.Wp);
      if constexpr (TG == 2) {
        // A unit is 2 * KG separate 8 KB blocks (stage layout [k-block of the unit][tile]: the two tiles of a k-block
        // form one N = 128 operand).  Each of its copies comes from a different lane, so a unit's bytes are requested at
        // once instead of one copy per ~0.3 us: lane l issues copy l % NCOPY of unit (iteration * UPI + l / NCOPY).
        constexpr int NCOPY = 2 * KG, UPI = SK_WLANES / NCOPY;
        static_assert(SK_WLANES % NCOPY == 0 && UPI >= 1, "lanes must be a multiple of the copies per unit");
        const int sub = lane % NCOPY, kk = sub >> 1, j = sub & 1, uo = lane / NCOPY;
        const unsigned peers = ((1u << NCOPY) - 1u) << (uo * NCOPY);
        for (int i = uo; i < n_units; i += UPI) {
          const int s = i % SW;
          int grp;
          const int ku = sk_unit_ku(u0 + i, u0, u1, KBU, cta, grp);
          mbar_wait(&empty_w[s], ((i / SW) & 1) ^ 1);
          if (sub == 0) mbar_arrive_expect_tx(&full_w[s], C::W_STAGE_BYTES);
          __syncwarp(peers);  // the transaction count is armed before any of the unit's copies can complete
          bulk_g2s_hint(w_ring + s * C::W_STAGE_BYTES + sub * C::TILE_BYTES,
                        wp + ((size_t)(2 * grp + j) * KB + (size_t)ku * KG + kk) * C::TILE_BYTES, C::TILE_BYTES, &full_w[s], pol);
          if (i == 0 && sub == 0) SK_STAMP(2);
          if (sub == 0) SK_USTAMP(i, 0);
        }
      } else {
        for (int i = lane; i < n_units; i += SK_WLANES) {  // a unit is contiguous in memory: one copy
          const int s = i % SW;
          int grp;
          const int ku = sk_unit_ku(u0 + i, u0, u1, KBU, cta, grp);
          mbar_wait(&empty_w[s], ((i / SW) & 1) ^ 1);
          mbar_arrive_expect_tx(&full_w[s], C::W_STAGE_BYTES);
          bulk_g2s_hint(w_ring + s * C::W_STAGE_BYTES, wp + ((size_t)grp * KBU + ku) * C::W_STAGE_BYTES, C::W_STAGE_BYTES,
                        &full_w[s], pol);
          if (i == 0) SK_STAMP(2);
          SK_USTAMP(i, 0);
        }
      }
      if (lane == 0) {
        SK_STAMP(10);
        prefetch_next_weights(g.next, cta);
        prefetch_kv(g.kvpf, cta);
      }
    }
  } else if (warp == MMA_WARP) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc(GEMM_BM, TG * C::NTILE, 1);
      int i = 0, f = 0, u = u0;
      while (u < u1) {
        const int grp = u / KBU, ku_begin = u - grp * KBU;
        const int ku_end = min(KBU, ku_begin + (u1 - u));
        const int b = f & 1;
        mbar_wait(&tmem_empty[b], ((f >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(b * TG * C::NTILE);
        for (int ku = ku_begin; ku < ku_end; ++ku, ++i) {
          const int sa = i % SA, sw = i % SW;
          SK_USTAMP(i, 1);
          mbar_wait(&full_w[sw], (i / SW) & 1);
          if (i == 0) SK_STAMP(3);
          SK_USTAMP(i, 2);
          mbar_wait(&full_a[sa], (i / SA) & 1);
          if (i == 0) SK_STAMP(5);
          SK_USTAMP(i, 3);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(a_ring + sa * C::A_STAGE_BYTES);
          const uint32_t w_addr = smem_u32(w_ring + sw * C::W_STAGE_BYTES);
#pragma unroll
          for (int kk = 0; kk < KG; ++kk) {
            const uint64_t adesc = umma_desc_sw128(a_addr + kk * C::A_SUB_BYTES);
            const uint64_t bdesc = umma_desc_sw128(w_addr + kk * TG * C::TILE_BYTES);
#ifndef SK_DBG_NOMMA
#pragma unroll
            for (int k = 0; k < GEMM_BK / 16; ++k)
              umma_f16(tmem_d, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc,
                       (ku != ku_begin) || (kk != 0) || (k != 0));  // (ku counts the fragment's units in processing order)
#else
            (void)adesc, (void)bdesc;
#endif
          }
          SK_USTAMP(i, 4);
#ifdef SK_DBG_ARRIVE  // what-if (with SK_DBG_NOMMA): plain arrives instead of tcgen05.commit
          mbar_arrive(&empty_a[sa]);
          mbar_arrive(&empty_w[sw]);
#else
          umma_commit(&empty_a[sa]);
          umma_commit(&empty_w[sw]);
#endif
          SK_USTAMP(i, 5);
        }
        umma_commit(&tmem_full[b]);
        u += ku_end - ku_begin;
        ++f;
      }
      SK_STAMP(6);
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ A producers (cp.async gather -> SW128 tiles)
    pdl_wait_prior();  // activations are written by the previous kernel
    const int ptid = tid - 128;
    if (ptid == 0) SK_STAMP(4);
    const int rows = min(g.m_valid, A_ROWS);
    constexpr int CHUNKS = KG * A_ROWS * 8;  // 16-byte chunks per stage
    constexpr int PER_THREAD = (CHUNKS + C::PROD_THREADS - 1) / C::PROD_THREADS;
    constexpr int MAXPEND = (SA < 8) ? SA : 8;
    int issued = 0, arrived = 0;
    while (arrived < n_units) {
      // Slow start: every CTA of the grid asks L2 for the same activations at the same moment, so with the whole ring
      // requested up front the FIRST stage lands only when the whole burst (64-96 KB per CTA, 10+ MB per grid: 2-3 us)
      // has been served.  The first stage goes out alone, then the ring fills as stages land.
      bool can_issue = issued < n_units && (issued - arrived) < min(MAXPEND, SK_A_SLOW_START + arrived * SK_A_SLOW_GROW);
      if (can_issue) {
        uint64_t* eb = &empty_a[issued % SA];
        const uint32_t par = ((issued / SA) & 1) ^ 1;
        if (issued == arrived) mbar_wait(eb, par);          // nothing in flight: block until a stage is free
        else can_issue = mbar_test_wait(eb, par);           // otherwise poll, and retire landed stages meanwhile
      }
      if (can_issue) {
        uint8_t* a_tile = a_ring + (issued % SA) * C::A_STAGE_BYTES;
        int grp_unused;
        const int ku = sk_unit_ku(u0 + issued, u0, u1, KBU, cta, grp_unused);
#pragma unroll
        for (int j = 0; j < PER_THREAD; ++j) {
          const int idx = j * C::PROD_THREADS + ptid;
          const int kk = idx / (A_ROWS * 8), rem = idx - kk * (A_ROWS * 8);
          const int r = rem >> 3, c = rem & 7;
#ifdef SK_DBG_NOA
          if (idx < CHUNKS && r < rows && issued == 0) {  // what-if: only the first stage is really loaded
#else
          if (idx < CHUNKS && r < rows) {
#endif
            bool valid = true;
            const void* src = aload.src(r, 0, ku * KG + kk, c, valid);
            cp_async16(a_tile + kk * C::A_SUB_BYTES + sw128_offset(r, c), src, valid ? 16u : 0u);
          }
        }
        cp_async_commit();
        ++issued;
      } else {
        switch (issued - arrived) {  // wait for the OLDEST pending group only
          case 1: cp_async_wait<0>(); break;
          case 2: cp_async_wait<1>(); break;
          case 3: cp_async_wait<2>(); break;
          case 4: cp_async_wait<3>(); break;
          case 5: cp_async_wait<4>(); break;
          case 6: cp_async_wait<5>(); break;
          case 7: cp_async_wait<6>(); break;
          default: cp_async_wait<7>(); break;
        }
        fence_proxy_async();
        mbar_arrive(&full_a[arrived % SA]);
        ++arrived;
      }
    }
    // ---- at 128 rows the four producer warps are done now: they take the second n_tile of the CTA's LAST fragment off
    // the epilogue warps (one warp per scheduler is latency-bound on the SwiGLU / store sequences)
    if constexpr (C::PROD_WARPS == 4 && TG == 2) {
      const int grp = (u1 - 1) / KBU, u_frag = max(u0, grp * KBU);
      const int ku_begin = u_frag - grp * KBU, ku_end = u1 - grp * KBU;
      const bool complete = (ku_begin == 0 && ku_end == KBU);
      if (sk_is_deferred<Epilogue>::value || complete) {
        const int f = grp - u0 / KBU, b = f & 1;
        const int hw = warp - 4, row = hw * 32 + lane;
        mbar_wait(&tmem_full[b], (f >> 1) & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(hw * 32) << 16) + (uint32_t)(b * TG * C::NTILE) + C::NTILE;
        float acc[C::NTILE];
        tmem_ld32(taddr, acc);
        tmem_ld32(taddr + 32, acc + 32);
        tmem_ld_wait();
        if constexpr (sk_is_deferred<Epilogue>::value)
          epi.partial_warp(reinterpret_cast<float*>(w_ring) + warp * 2048, lane, hw * 32, grp * TG + 1, cta % g.n_splits, acc);
        else
          epi(row, row < g.m_valid, 0, grp * TG + 1, acc);
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (row == TMEM lane)
    pdl_wait_prior();  // residual input, workspace and tickets belong to the previous kernels until they are done
    const int row = tid;
    const bool row_valid = row < g.m_valid;
    float4* ws = reinterpret_cast<float4*>(g.partial_ws);
    int f = 0, u = u0;
    while (u < u1) {
      const int grp = u / KBU, ku_begin = u - grp * KBU;
      const int ku_end = min(KBU, ku_begin + (u1 - u));
      const int b = f & 1;
      const bool complete = (ku_begin == 0 && ku_end == KBU);
      const int slot = (u == u0) ? 0 : 1;  // a CTA has at most two partial fragments: its first and its last
      float4* mine = ws + ((size_t)cta * 2 + slot) * TG * C::PARTIAL_F4;
      mbar_wait(&tmem_full[b], (f >> 1) & 1);
      if (tid == 0 && f == 0) SK_STAMP(7);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(b * TG * C::NTILE);
      float acc[C::NTILE];
      // the producer warps take tile 1 of the last fragment (see above)
      const bool helped = (C::PROD_WARPS == 4 && TG == 2) && (u + (ku_end - ku_begin) == u1) &&
                          (sk_is_deferred<Epilogue>::value || complete);
#pragma unroll
      for (int j = 0; j < TG; ++j) {
        if (helped && j == 1) break;
        if constexpr (C::NTILE == 16) {
          tmem_ld16(taddr, acc);
        } else {
          tmem_ld32(taddr + j * C::NTILE, acc);
          tmem_ld32(taddr + j * C::NTILE + 32, acc + 32);
        }
        tmem_ld_wait();
        if (j == TG - 1 || (helped && j == 0)) {
          tc_fence_before();
          mbar_arrive(&tmem_empty[b]);  // the issuer may start the next-but-one fragment into this accumulator set
        }
        if constexpr (sk_is_deferred<Epilogue>::value) {
          if constexpr (C::PROD_WARPS == 4) {
            if (helped)  // (the ring is idle: all MMAs of the CTA's last fragment have completed)
              epi.partial_warp(reinterpret_cast<float*>(w_ring) + warp * 2048, lane, warp * 32, grp * TG + j, cta % g.n_splits, acc);
            else
              epi.partial(row, row_valid, grp * TG + j, cta % g.n_splits, acc);
          } else {
            epi.partial(row, row_valid, grp * TG + j, cta % g.n_splits, acc);
          }
        } else if (complete) {
          epi(row, row_valid, 0, grp * TG + j, acc);
        } else if (row_valid) {
#pragma unroll
          for (int c = 0; c < C::NTILE / 4; ++c)
            mine[(j * (C::NTILE / 4) + c) * 128 + row] = make_float4(acc[4 * c], acc[4 * c + 1], acc[4 * c + 2], acc[4 * c + 3]);
        }
      }
      if (!sk_is_deferred<Epilogue>::value && !complete) {
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
          for (int j = 0; j < TG; ++j) {
            if (row_valid) {
#pragma unroll
              for (int c = 0; c < C::NTILE; ++c) acc[c] = 0.f;
              for (int cc = c_first; cc <= c_last; ++cc) {  // fixed order -> bitwise reproducible
                const int sl = (sk_first_unit(cc, units, G) / KBU == grp) ? 0 : 1;
                const float4* part = ws + (((size_t)cc * 2 + sl) * TG + j) * C::PARTIAL_F4;
#pragma unroll
                for (int c = 0; c < C::NTILE / 4; ++c) {
                  const float4 v = __ldcg(&part[c * 128 + row]);
                  acc[4 * c] += v.x, acc[4 * c + 1] += v.y, acc[4 * c + 2] += v.z, acc[4 * c + 3] += v.w;
                }
              }
            }
            if constexpr (!sk_is_deferred<Epilogue>::value) epi(row, row_valid, 0, grp * TG + j, acc);
          }
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
  if (warp == MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------ launch helpers
template <int A_ROWS, int MODE, class ALoader, class Epilogue>
cudaError_t launch_gemm_skinny_cfg(const SkinnyCore& g, const ALoader& al, const Epilogue& epi, int grid, cudaStream_t st) {
  using C = SkinnyCfg<A_ROWS, MODE>;
  auto kern = gemm_skinny_kernel<A_ROWS, MODE, ALoader, Epilogue>;
  static bool configured = false;  // per template instantiation
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::DYN_BYTES);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  const int units = (g.n_tiles / C::TG) * (g.k_blocks / C::KG);
  grid = grid < 1 ? 1 : (grid > units ? units : grid);
  return launch_pdl(kern, dim3((unsigned)grid), dim3(C::THREADS), (size_t)C::DYN_BYTES, st, pdl_enabled(), g, al, epi);
}

// M <= 8 with weights packed in 16-row tiles (g.n_tiles = N / 16, k_blocks % 8 == 0): one CTA per column slice.
template <class ALoader, class Epilogue>
cudaError_t launch_gemm_skinny_n16(const SkinnyCore& g, const ALoader& al, const Epilogue& epi, cudaStream_t st) {
  return launch_gemm_skinny_cfg<8, MODE_N16>(g, al, epi, g.n_tiles, st);
}

// Mode the launcher picks for (rows, n_tiles, k_blocks); prefer_k1 forces 8 KB single-tile units at M > 32 (smaller
// fix-up slots when a group is shared by several CTAs).
inline int skinny_mode(int m, int n_tiles, int k_blocks, bool prefer_k1 = false) {
  // An N <= 128 UMMA costs ~77 cycles whatever N is (scripts/exp_umma_rate.cu): with N = 64 (K2 / K1) the MMAs of a
  // 16 KB stage take 616 cycles, i.e. ~50 GB/s per SM at best; two tiles per MMA (T2) halve that.
  static const bool small_t2 = zb::tuning_env("ZB200_SMALL_T2") != nullptr;  // measured slower at batch 1-16: opt-in
  if (m <= 32 && !(small_t2 && n_tiles % 2 == 0 && !prefer_k1)) return (k_blocks % 2 == 0) ? MODE_K2 : MODE_K1;
  if (n_tiles % 2 != 0 || prefer_k1) return MODE_K1;
  // Every unit costs the MMA issuer and the activation producers ~0.2 us of barrier round trips on top of its MMAs
  // (in-kernel trace): long streams take two k-blocks of the tile pair per unit (32 KB of weights), short
  // ones keep 16 KB units so that a CTA's share still splits into a few pipeline stages.
  return (m > 32 && k_blocks % 2 == 0 && (int64_t)(n_tiles / 2) * k_blocks >= SK_T2K2_MIN_UNITS) ? MODE_T2K2 : MODE_T2;
}
inline int skinny_tg(int mode) { return (mode == MODE_T2 || mode == MODE_T2K2) ? 2 : 1; }
inline int skinny_kg(int mode) { return (mode == MODE_K2 || mode == MODE_T2K2) ? 2 : 1; }

// grid: number of CTAs wanted (clamped to the unit count).  Ticketed epilogues: ws must hold grid*2*tg 32 KB slots and
// tickets n_tiles ints.  Deferred epilogues: grid must be n_groups * splits with splits | units-per-group.
template <class ALoader, class Epilogue>
cudaError_t launch_gemm_skinny(const SkinnyCore& g, const ALoader& al, const Epilogue& epi, int grid, cudaStream_t st,
                               bool prefer_k1 = false) {
  const int mode = skinny_mode(g.m_valid, g.n_tiles, g.k_blocks, prefer_k1);
  if (g.m_valid <= 8)
    return mode == MODE_K2   ? launch_gemm_skinny_cfg<8, MODE_K2>(g, al, epi, grid, st)
           : mode == MODE_T2 ? launch_gemm_skinny_cfg<8, MODE_T2>(g, al, epi, grid, st)
                             : launch_gemm_skinny_cfg<8, MODE_K1>(g, al, epi, grid, st);
  if (g.m_valid <= 32)
    return mode == MODE_K2   ? launch_gemm_skinny_cfg<32, MODE_K2>(g, al, epi, grid, st)
           : mode == MODE_T2 ? launch_gemm_skinny_cfg<32, MODE_T2>(g, al, epi, grid, st)
                             : launch_gemm_skinny_cfg<32, MODE_K1>(g, al, epi, grid, st);
  if (g.m_valid <= 64)
    return mode == MODE_T2K2 ? launch_gemm_skinny_cfg<64, MODE_T2K2>(g, al, epi, grid, st)
           : mode == MODE_T2 ? launch_gemm_skinny_cfg<64, MODE_T2>(g, al, epi, grid, st)
                             : launch_gemm_skinny_cfg<64, MODE_K1>(g, al, epi, grid, st);
  return mode == MODE_T2K2 ? launch_gemm_skinny_cfg<128, MODE_T2K2>(g, al, epi, grid, st)
         : mode == MODE_T2 ? launch_gemm_skinny_cfg<128, MODE_T2>(g, al, epi, grid, st)
                           : launch_gemm_skinny_cfg<128, MODE_K1>(g, al, epi, grid, st);
}

// ---- grid planning for ticketed epilogues (shared by the engine and the single-GEMM hooks) ----------------------------
// Every CTA should be exactly one fragment, so grids are aligned to the group structure: whole groups per CTA when there
// are enough groups; otherwise each group is shared by at most 3 CTAs (the last one to arrive reduces, so the
// contributor count is the fix-up latency).  `sms` CTAs are resident at M > 32 (192 KB of shared memory each), 2 * sms at
// M <= 32.
struct SkPlan {
  int grid, splits;
  bool prefer_k1;
};
inline int big_m_ctas_per_sm() {  // experiment knob: CTAs per SM the planners assume at rows > 32
  static const int v = zb::tuning_env("ZB200_SK_CAP2") ? 2 : 1;
  return v;
}
inline SkPlan plan_ticketed(int sms, int rows, int n_tiles, int k_blocks) {
  const int cap = sms * (rows <= 32 ? 2 : big_m_ctas_per_sm());
  SkPlan p{cap, 1, false};
  int mode = skinny_mode(rows, n_tiles, k_blocks);
  int groups = n_tiles / skinny_tg(mode);
  if (groups * 2 <= cap && mode == MODE_T2) {  // few, shared groups: single-tile units keep the fix-up slots at 32 KB
    p.prefer_k1 = true;
    mode = MODE_K1, groups = n_tiles;
  }
  if (groups >= cap) {
    for (int r = 1; r <= 2; ++r)
      if (groups % r == 0 && groups / r <= cap) {
        p.grid = groups / r;
        return p;
      }
    p.grid = cap;  // unaligned stream-K: most groups end up with two contributors
  } else {
    static const int max_split = zb::tuning_env("ZB200_SK_MAXSPLIT") ? atoi(zb::tuning_env("ZB200_SK_MAXSPLIT")) : 3;
    const int s = std::max(1, std::min(max_split, cap / groups));
    p.grid = groups * s;
  }
  return p;
}

// Deferred epilogues (out_proj, fc2, the hybrid's out_proj): n_groups * splits CTAs, every CTA one fragment; the split
// sums are reduced by the consumer kernel (resid_ln_kernel / add_norm_kernel).  ws_floats: fp32 capacity of the partials.
inline SkPlan plan_deferred(int sms, int rows, int n_tiles, int k_blocks, int64_t ws_floats, int n) {
  const int cap = sms * (rows <= 32 ? 2 : big_m_ctas_per_sm());
  const int mode = skinny_mode(rows, n_tiles, k_blocks);
  const int groups = n_tiles / skinny_tg(mode), kbu = k_blocks / skinny_kg(mode);
  static const int max_split = zb::tuning_env("ZB200_DEF_MAXSPLIT") ? atoi(zb::tuning_env("ZB200_DEF_MAXSPLIT")) : 64;
  int s = 1;
  for (int c = 1; c <= kbu && c <= max_split; ++c)
    if (groups * c <= cap && kbu / c >= 2 && (int64_t)c * rows * n <= ws_floats) s = c;
  return SkPlan{groups * s, s, false};
}

}  // namespace zb
