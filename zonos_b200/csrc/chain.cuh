// Fused projection chain of the decode step: ONE persistent launch runs up to four weight-streaming GEMM phases with
// the reductions between them,
//     out_proj -> [split-K sum + residual + LayerNorm2] -> fc1 + SwiGLU -> fc2 -> [split-K sum + residual + next LayerNorm]
//              -> next layer's in_proj -> [split-K sum + RoPE + q / KV-page write]        (heads + CFG after the last layer)
// i.e. everything between two attention kernels (zonos/backbone/_torch.py:99-102,117-141,150-152; model.py:100-116).
//
// Why: as separate launches (gemm_skinny.cuh + resid_ln / qkv_finish kernels) every 2-25 us GEMM pays ~5 us of launch
// ramp, first-byte latency and tail with nothing co-resident to hide it (192 KB of rings per CTA), and the step spends
// ~54 us per layer on 18.6 us worth of weights.  Here one CTA per SM lives through all phases:
//   * the weight-producer lanes never stop: the unit list of all phases is one sequence, so while a phase's epilogue,
//     grid barrier and reduction run, the ring already holds the next phase's first weights (112-160 KB per SM);
//   * phases are ordered by monotonic counters in global memory (one `done` and one `fin` counter per phase; arrive =
//     fence + atomicAdd by one thread, wait = ld.acquire spin by one thread + a named barrier) -- all CTAs are
//     co-resident by construction (grid <= SM count, one CTA per SM), and every spin is bounded: on a timeout a sticky
//     error flag is raised and all later waits fall through, so a logic error ends in wrong numbers, not a hung GPU;
//   * the reductions that used to be kernels (resid_ln_kernel, qkv_finish_kernel) are run by CTA r for row r.
// Work partition inside a phase, epilogue arithmetic and rounding points are those of gemm_skinny.cuh / ops.cu.
#pragma once
#include "common.cuh"
#include "epilogues.cuh"
#include "gemm_skinny.cuh"
#include "kv.cuh"

namespace zb {

enum { CH_EPI_PARTIAL = 0, CH_EPI_SWIGLU = 1, CH_EPI_HEADS = 2 };
enum { CH_FIN_NONE = 0, CH_FIN_RESID_LN = 1, CH_FIN_QKV = 2 };
constexpr int CH_MAX_PHASES = 4;
constexpr int CH_SYNC_INTS = 2 * CH_MAX_PHASES;  // per launch: {done, fin} per phase
constexpr int CH_TRACE_EVENTS = 8;

struct ChainPhase {
  const uint8_t* W;  // packed weights [n_tiles][k_blocks][8 KB]
  int n_tiles, k_blocks;
  int grid;      // CTAs [0, grid) take units of this phase
  int n_splits;  // CH_EPI_PARTIAL: CTAs per tile pair (grid == pairs * n_splits); partial layout ws[split][row][ldn]
  int epi, fin;
  int a_pairs;   // A rows gathered as (cond, uncond) pairs (heads: ARowCfgPairs)
  const uint8_t* A;
  int a_row_bytes;
  int ldn;            // PARTIAL: columns of the output (N)
  bf16* h;            // SWIGLU: output [rows][ldh]
  int ldh;
  const bf16 *ln_w, *ln_b;  // FIN_RESID_LN: the norm that FOLLOWS the residual add
  int kv_layer;             // FIN_QKV: layer of the KV pool the k / v rows go to
};

struct ChainArgs {
  ChainPhase ph[CH_MAX_PHASES];
  int n_phases, rows;
  float* ws;       // split-K partials (deferred phases) / fix-up slots (ticketed phases): reused phase after phase
  int* tickets;    // [tile pairs] zero-initialised, self-resetting (ticketed phases)
  int* sync_mine;  // [CH_SYNC_INTS] zero on entry
  int* sync_prev;  // the previous chain launch's counters: cleared here (that launch has completed by then)
  int* err;        // sticky error flags (state[ST_ERROR])
  bf16 *x, *xn, *q;
  int d_model;
  float ln_eps;
  KvLayout kv;
  const float* rope;
  const int* pos_ptr;
  int n_heads;
  float* logits;
  int batch, n_q, vocab, tiles_per_head;
  float cfg_scale;
  unsigned long long* trace;  // optional [gridDim.x][CH_MAX_PHASES][CH_TRACE_EVENTS] globaltimer stamps
};

#ifndef CH_POLL_SLEEP
#define CH_POLL_SLEEP 0  // ns between polls of the fast path (experiment knob)
#endif
#ifndef CH_HELP
#define CH_HELP 1        // P warps take the second tile of a phase's last fragment (experiment knob)
#endif
#ifndef CH_SW128
#define CH_SW128 6       // weight-ring stages at 128 rows (experiment knob)
#endif
#ifndef CH_SA128
#define CH_SA128 4       // activation-ring stages at 128 rows (experiment knob)
#endif
template <int A_ROWS>
struct ChainCfg {
  // warps 0-3: epilogue group E; warps 4-7: A producers, then helper group P (tile 1 of a phase's last fragment, half of
  // every reduction); warp 8: weight producer; warp 9: MMA issuer.  Every variant has the four P warps: at few rows they
  // have little to gather, but the epilogues and reductions are the same size.
  static constexpr int PROD_WARPS = 4;
  static constexpr int PROD_THREADS = PROD_WARPS * 32;
  static constexpr int THREADS = (4 + PROD_WARPS + 2) * 32;
  static constexpr int A_STAGE_BYTES = A_ROWS * 128;
  static constexpr int W_STAGE_BYTES = 2 * SK_TILE_BYTES;  // one k-block of two adjacent n_tiles (UMMA N = 128)
  static constexpr int SA = (A_ROWS >= 128) ? CH_SA128 : 8;
  static constexpr int SW = (A_ROWS >= 128) ? CH_SW128 : 8;
  static constexpr int A_BYTES = SA * A_STAGE_BYTES;
  static constexpr int W_BYTES = SW * W_STAGE_BYTES;
  static constexpr int XPOSE_BYTES = (CH_HELP ? 8 : 4) * 8192;  // per-warp transpose buffers for coalesced partial stores (E and P warps)
  static constexpr int BAR_OFFSET = A_BYTES + W_BYTES + XPOSE_BYTES;
  static constexpr int N_BARS = 2 * SA + 2 * SW + 6;
  static constexpr int DYN_BYTES = BAR_OFFSET + N_BARS * 8 + 16 + 1024;
  static constexpr uint32_t TMEM_COLS = 256;  // two accumulator sets of 128 columns
  static_assert(A_BYTES % 1024 == 0 && W_BYTES >= GEMM_BM * 128, "operand tiles must be 1024-byte aligned / in bounds");
  static_assert(SK_WLANES <= SW, "weight-producer lanes must not outnumber the ring slots");
  static_assert(DYN_BYTES <= 227 * 1024, "shared memory budget");
};

__device__ __forceinline__ int ch_ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Bounded wait for a monotonic counter (one thread per CTA polls).  Sticky: once any wait of the launch timed out, the
// rest fall through.
__device__ __forceinline__ void ch_wait(const int* cnt, int target, int* err) {
  for (int it = 0; it < (1 << 22); ++it) {
    if (ch_ld_acquire(cnt) >= target) return;
    if (CH_POLL_SLEEP > 0) __nanosleep(CH_POLL_SLEEP);
  }
  if (*reinterpret_cast<volatile int*>(err) & 2) return;
  for (int it = 0; it < (1 << 22); ++it) {
    __nanosleep(100);
    if (ch_ld_acquire(cnt) >= target) return;
  }
  atomicOr(err, 2);
}
// Arrive after a CTA-wide (named) barrier: the barrier orders the group's writes before this thread's fence, the fence is
// cumulative at gpu scope (the cooperative-groups grid.sync pattern): one fence per CTA instead of one per thread.
__device__ __forceinline__ void ch_arrive(int* cnt) {
  __threadfence();
  atomicAdd(cnt, 1);
}
__device__ __forceinline__ void ch_stamp(const ChainArgs& a, int p, int ev) {
  if (a.trace) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    a.trace[((size_t)blockIdx.x * CH_MAX_PHASES + p) * CH_TRACE_EVENTS + ev] = t;
  }
}
__device__ __forceinline__ void ch_named_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// Units of phase `ph` owned by CTA `cta`: [u0, u1) of pairs * k_blocks units numbered pair-major.
struct ChainRange {
  int kbu, units, u0, u1;
};
__device__ __forceinline__ ChainRange ch_range(const ChainPhase& ph, int cta) {
  ChainRange r;
  r.kbu = ph.k_blocks;
  r.units = ((ph.n_tiles + 1) >> 1) * r.kbu;
  if (cta < ph.grid) {
    r.u0 = sk_first_unit(cta, r.units, ph.grid), r.u1 = sk_first_unit(cta + 1, r.units, ph.grid);
  } else {
    r.u0 = r.u1 = 0;
  }
  return r;
}

// ------------------------------------------------------------------------------------------------ reductions
// Both run on the 256 threads of the E and P groups of one CTA (named barrier 3) and issue ALL their global loads before
// the first use, so a row costs one L2 round trip, not one per column chunk and operand.
//
// Row `row` of:  o = bf16(sum_s ws[s][row][:]); x = bf16(x + o); xn = LayerNorm(x) * w + b     (resid_ln_kernel, ops.cu)
// d <= 4096, d % 256 == 0; thread t owns columns [8t, 8t + 8) (+ 2048 for d > 2048).
template <int MAXV>
__device__ __forceinline__ void ch_resid_ln_row_t(const ChainArgs& a, const ChainPhase& ph, int row, int t, float* red) {
  constexpr int BATCH = (MAXV == 1) ? 9 : 5;
  const int d = a.d_model, splits = ph.n_splits, rows = a.rows, lane = t & 31, warp = t >> 5;
  float o[MAXV][8];
  uint4 xu[MAXV], wu[MAXV], bu[MAXV];
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int col = (i * 256 + t) * 8;
    if (col < d) {
      xu[i] = __ldcg(reinterpret_cast<const uint4*>(a.x + (size_t)row * d + col));
      wu[i] = __ldg(reinterpret_cast<const uint4*>(ph.ln_w + col)), bu[i] = __ldg(reinterpret_cast<const uint4*>(ph.ln_b + col));
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) o[i][j] = 0.f;
  }
  for (int s0 = 0; s0 < splits; s0 += BATCH) {  // one round of independent loads for <= BATCH splits, added in split order
    float4 p0[MAXV][BATCH], p1[MAXV][BATCH];
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int col = (i * 256 + t) * 8;
#pragma unroll
      for (int s = 0; s < BATCH; ++s)
        if (col < d && s0 + s < splits) ldg256_cg(a.ws + ((size_t)(s0 + s) * rows + row) * d + col, p0[i][s], p1[i][s]);
    }
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int col = (i * 256 + t) * 8;
#pragma unroll
      for (int s = 0; s < BATCH; ++s)
        if (col < d && s0 + s < splits) {
          o[i][0] += p0[i][s].x, o[i][1] += p0[i][s].y, o[i][2] += p0[i][s].z, o[i][3] += p0[i][s].w;
          o[i][4] += p1[i][s].x, o[i][5] += p1[i][s].y, o[i][6] += p1[i][s].z, o[i][7] += p1[i][s].w;
        }
    }
  }
  float v[MAXV][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int col = (i * 256 + t) * 8;
    if (col < d) {
      const float2 x0 = unpack_bf16x2(xu[i].x), x1 = unpack_bf16x2(xu[i].y), x2 = unpack_bf16x2(xu[i].z), x3 = unpack_bf16x2(xu[i].w);
      const float xin[8] = {x0.x, x0.y, x1.x, x1.y, x2.x, x2.y, x3.x, x3.y};
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[i][j] = bf16_round(xin[j] + bf16_round(o[i][j]));
        sum += v[i][j];
      }
      *reinterpret_cast<uint4*>(a.x + (size_t)row * d + col) =
          make_uint4(pack_bf16x2(v[i][0], v[i][1]), pack_bf16x2(v[i][2], v[i][3]), pack_bf16x2(v[i][4], v[i][5]),
                     pack_bf16x2(v[i][6], v[i][7]));
    }
  }
  auto block_sum = [&](float x) {
    x = warp_sum(x);
    ch_named_sync(3, 256);
    if (lane == 0) red[warp] = x;
    ch_named_sync(3, 256);
    return ((red[0] + red[1]) + (red[2] + red[3])) + ((red[4] + red[5]) + (red[6] + red[7]));
  };
  const float mean = block_sum(sum) / (float)d;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i)
    if ((i * 256 + t) * 8 < d)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float c = v[i][j] - mean;
        sq += c * c;
      }
  const float rstd = 1.0f / sqrtf(block_sum(sq) / (float)d + a.ln_eps);
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int col = (i * 256 + t) * 8;
    if (col < d) {
      const uint32_t wv[4] = {wu[i].x, wu[i].y, wu[i].z, wu[i].w}, bv[4] = {bu[i].x, bu[i].y, bu[i].z, bu[i].w};
      uint32_t r[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 ww = unpack_bf16x2(wv[j]), bb = unpack_bf16x2(bv[j]);
        r[j] = pack_bf16x2((v[i][2 * j] - mean) * rstd * ww.x + bb.x, (v[i][2 * j + 1] - mean) * rstd * ww.y + bb.y);
      }
      *reinterpret_cast<uint4*>(a.xn + (size_t)row * d + col) = make_uint4(r[0], r[1], r[2], r[3]);
    }
  }
}
__device__ __forceinline__ void ch_resid_ln_row(const ChainArgs& a, const ChainPhase& ph, int row, int t, float* red) {
  if (a.d_model <= 2048) ch_resid_ln_row_t<1>(a, ph, row, t, red);
  else ch_resid_ln_row_t<2>(a, ph, row, t, red);
}

// Row `row` of the in_proj finish (qkv_finish_kernel, ops.cu): split sum -> bf16 -> interleaved RoPE on q and k -> q
// buffer / KV page rows.  Thread t owns columns [8t, 8t + 8), + 2048 per further round.
__device__ __forceinline__ void ch_qkv_row(const ChainArgs& a, const ChainPhase& ph, int row, int t) {
  const int n_cols = ph.ldn, splits = ph.n_splits, rows = a.rows, n_heads = a.n_heads;
  KvLayout kv = a.kv;
  kv.layer = ph.kv_layer;
  const int kv_heads = kv.n_kv_heads;
  const int pos = *a.pos_ptr;
  for (int col0 = t * 8; col0 < n_cols; col0 += 256 * 8) {
    const int head_all = col0 >> 7, within = col0 & 127;
    const bool rot = head_all < n_heads + kv_heads;
    float4 t01 = make_float4(0.f, 0.f, 0.f, 0.f), t23 = t01;
    if (rot) {
      const float4* cs = reinterpret_cast<const float4*>(a.rope + ((size_t)pos * 64 + (within >> 1)) * 2);
      t01 = __ldg(cs), t23 = __ldg(cs + 1);
    }
    float o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int s0 = 0; s0 < splits; s0 += 8) {
      float4 p0[8], p1[8];
#pragma unroll
      for (int s = 0; s < 8; ++s)
        if (s0 + s < splits) ldg256_cg(a.ws + ((size_t)(s0 + s) * rows + row) * n_cols + col0, p0[s], p1[s]);
#pragma unroll
      for (int s = 0; s < 8; ++s)
        if (s0 + s < splits) {
          o[0] += p0[s].x, o[1] += p0[s].y, o[2] += p0[s].z, o[3] += p0[s].w;
          o[4] += p1[s].x, o[5] += p1[s].y, o[6] += p1[s].z, o[7] += p1[s].w;
        }
    }
    float r[8];
    if (rot) {
      const float co[4] = {t01.x, t01.z, t23.x, t23.z}, si[4] = {t01.y, t01.w, t23.y, t23.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float u = bf16_round(o[2 * i]), w = bf16_round(o[2 * i + 1]);
        r[2 * i] = __fsub_rn(__fmul_rn(u, co[i]), __fmul_rn(w, si[i]));
        r[2 * i + 1] = __fadd_rn(__fmul_rn(w, co[i]), __fmul_rn(u, si[i]));
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) r[i] = o[i];
    }
    const uint4 v = make_uint4(pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]), pack_bf16x2(r[4], r[5]), pack_bf16x2(r[6], r[7]));
    if (head_all < n_heads) {
      *reinterpret_cast<uint4*>(a.q + (size_t)row * n_heads * 128 + head_all * 128 + within) = v;
    } else {
      const int is_v = head_all >= n_heads + kv_heads;
      const int head = head_all - n_heads - (is_v ? kv_heads : 0);
      reinterpret_cast<uint4*>(kv.row(row, pos, is_v, head))[(within >> 3) ^ (pos & 7)] = v;
    }
  }
}

// ------------------------------------------------------------------------------------------------ the kernel
// Fused epilogue of one 64-column tile held by a warp (lane == TMEM lane == row within the warp's 32-row quarter).
__device__ __forceinline__ void ch_tile_epilogue(const ChainArgs& a, const ChainPhase& ph, int quarter, int lane, int n_tile,
                                                 int split, float* acc, float* xpose_w) {
  const int row = quarter * 32 + lane;
  const bool row_valid = row < a.rows;
  if (ph.epi == CH_EPI_PARTIAL) {
    const EpiPartial epi{a.ws, a.rows, ph.ldn};
    epi.partial_warp(xpose_w, lane, quarter * 32, n_tile, split, acc);
  } else if (ph.epi == CH_EPI_SWIGLU) {
    const EpiSwiGLU<SK_NT> epi{ph.h, ph.ldh};
    epi(row, row_valid, 0, n_tile, acc);
  } else {
    const EpiHeadsCfg<SK_NT> epi{a.logits, a.batch, a.n_q, a.vocab, a.tiles_per_head, a.cfg_scale};
    epi(row, row_valid, 0, n_tile, acc);
  }
}
// A fragment [ku_begin, ku_end) of tile pair `pair` that ends the CTA's range of a phase is "helped": the P warps take its
// second tile.  Only fragments whose epilogue runs straight from TMEM qualify (deferred partials, or a complete K).
__device__ __forceinline__ bool ch_helped(const ChainPhase& ph, int tiles, bool is_last, bool complete) {
  return CH_HELP && tiles == 2 && is_last && (ph.epi == CH_EPI_PARTIAL || complete);
}

template <int A_ROWS>
__global__ void __launch_bounds__(ChainCfg<A_ROWS>::THREADS, 1) chain_kernel(const __grid_constant__ ChainArgs a) {
  using C = ChainCfg<A_ROWS>;
  constexpr int SA = C::SA, SW = C::SW;
  constexpr int W_WARP = 4 + C::PROD_WARPS, MMA_WARP = W_WARP + 1;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_ring = smem;
  uint8_t* w_ring = smem + C::A_BYTES;
  float* xpose = reinterpret_cast<float*>(smem + C::A_BYTES + C::W_BYTES);
  uint64_t* full_a = reinterpret_cast<uint64_t*>(smem + C::BAR_OFFSET);
  uint64_t* empty_a = full_a + SA;
  uint64_t* full_w = empty_a + SA;
  uint64_t* empty_w = full_w + SW;
  uint64_t* tmem_full = empty_w + SW;
  uint64_t* tmem_empty = tmem_full + 2;    // E group has read the accumulator set
  uint64_t* tmem_empty_h = tmem_empty + 2;  // P group has read its tile of a helped fragment
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_h + 2);
  __shared__ int s_is_last;
  __shared__ float s_red[8];

  pdl_launch_dependents();

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, cta = blockIdx.x;
  if (tid == 0) {
    ch_stamp(a, 0, 7);
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
      mbar_init(&tmem_empty_h[b], 128);
    }
    fence_barrier_init();
  }
  if (warp == MMA_WARP) tmem_alloc(tmem_slot, C::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == W_WARP) {
    // ---------------------------------------------------------------- weight producer: one unit sequence over all phases
    // Unit gi (global index over the CTA's units of all phases) is issued by lane gi % SK_WLANES, so a lane's successive
    // units are exactly SK_WLANES <= SW apart and it can never be two ring revolutions ahead of the consumer (the parity
    // wait cannot tell 0 from 2 completions).
    if (lane < SK_WLANES) {
      const uint64_t pol = l2_policy_evict_first();
      int gi = 0;
      for (int p = 0; p < a.n_phases; ++p) {
        const ChainPhase& ph = a.ph[p];
        const ChainRange r = ch_range(ph, cta);
        const int KB = ph.k_blocks, n_tiles = ph.n_tiles;
        const uint8_t* wbase = ph.W;
        for (int u = r.u0; u < r.u1; ++u, ++gi) {
          if (gi % SK_WLANES != lane) continue;
          const int s = gi % SW;
          const int pair = u / r.kbu, ku = u - pair * r.kbu;
          const int tiles = min(2, n_tiles - 2 * pair);
          mbar_wait(&empty_w[s], ((gi / SW) & 1) ^ 1);
          mbar_arrive_expect_tx(&full_w[s], (uint32_t)(tiles * SK_TILE_BYTES));
          uint8_t* dst = w_ring + s * C::W_STAGE_BYTES;
          const uint8_t* src = wbase + ((size_t)(2 * pair) * KB + ku) * SK_TILE_BYTES;
          bulk_g2s_hint(dst, src, SK_TILE_BYTES, &full_w[s], pol);
          if (tiles == 2) bulk_g2s_hint(dst + SK_TILE_BYTES, src + (size_t)KB * SK_TILE_BYTES, SK_TILE_BYTES, &full_w[s], pol);
        }
      }
    }
  } else if (warp == MMA_WARP) {
    // ---------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      int gi = 0, f = 0;
      uint32_t hpend = 0, hpar = 0;  // bit b: set b's previous fragment was helped (await the P group's read too) / its parity
      for (int p = 0; p < a.n_phases; ++p) {
        const ChainPhase& ph = a.ph[p];
        const ChainRange r = ch_range(ph, cta);
        int u = r.u0;
        while (u < r.u1) {
          const int pair = u / r.kbu, ku_begin = u - pair * r.kbu;
          const int ku_end = min(r.kbu, ku_begin + (r.u1 - u));
          const int tiles = min(2, ph.n_tiles - 2 * pair);
          const uint32_t idesc = umma_idesc(GEMM_BM, tiles * SK_NT, 1);
          const int b = f & 1;
          mbar_wait(&tmem_empty[b], ((f >> 1) & 1) ^ 1);
          if ((hpend >> b) & 1) {
            mbar_wait(&tmem_empty_h[b], (hpar >> b) & 1);
            hpar ^= 1u << b, hpend &= ~(1u << b);
          }
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + (uint32_t)(b * 2 * SK_NT);
          for (int ku = ku_begin; ku < ku_end; ++ku, ++gi) {
            const int sa = gi % SA, sw = gi % SW;
            mbar_wait(&full_w[sw], (gi / SW) & 1);
            if (ku == ku_begin && u == r.u0) ch_stamp(a, p, 0);
            mbar_wait(&full_a[sa], (gi / SA) & 1);
            if (ku == ku_begin && u == r.u0) ch_stamp(a, p, 2);
            tc_fence_after();
            const uint64_t adesc = umma_desc_sw128(smem_u32(a_ring + sa * C::A_STAGE_BYTES));
            const uint64_t bdesc = umma_desc_sw128(smem_u32(w_ring + sw * C::W_STAGE_BYTES));
#pragma unroll
            for (int k = 0; k < GEMM_BK / 16; ++k)
              umma_f16(tmem_d, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (ku != ku_begin) || (k != 0));
            umma_commit(&empty_a[sa]);
            umma_commit(&empty_w[sw]);
          }
          umma_commit(&tmem_full[b]);
          if (ch_helped(ph, tiles, u + (ku_end - ku_begin) == r.u1, ku_begin == 0 && ku_end == r.kbu)) hpend |= 1u << b;
          u += ku_end - ku_begin;
          ++f;
        }
      }
    }
  } else if (warp >= 4) {
    // ---------------------------------------------------------------- P group: A producers, then helpers
    pdl_wait_prior();  // the first phase's activations come from the previous kernel
    const int ptid = tid - 128, quarter = warp - 4;
    const int rows = min(a.rows, A_ROWS);
    constexpr int CHUNKS = A_ROWS * 8;  // 16-byte chunks per stage
    constexpr int PER_THREAD = (CHUNKS + C::PROD_THREADS - 1) / C::PROD_THREADS;
    constexpr int MAXPEND = (SA < 8) ? SA : 8;
    int a_base = 0, f_base = 0;
    for (int p = 0; p < a.n_phases; ++p) {
      const ChainPhase& ph = a.ph[p];
      const ChainRange r = ch_range(ph, cta);
      const int n_units = r.u1 - r.u0;
      if (n_units > 0) {
        if (p > 0) {  // this phase's A is produced inside the launch: wait for the reduction (or the epilogues) before it
          if (ptid == 0) {
            const ChainPhase& pv = a.ph[p - 1];
            if (pv.fin != CH_FIN_NONE) ch_wait(&a.sync_mine[2 * (p - 1) + 1], a.rows, a.err);
            else ch_wait(&a.sync_mine[2 * (p - 1)], pv.grid, a.err);
          }
          ch_named_sync(2, C::PROD_THREADS);
        }
        if (ptid == 0) ch_stamp(a, p, 1);
        // Per-phase invariants of this thread's chunks.  Chunk j of the stage is row 16 j + r0, 16-byte column c: its
        // swizzled destination is dst0 + 2048 j (the XOR term only sees r0 & 7) and its source src0 + j * step, so three
        // registers describe all of them.  They are pinned with empty asm statements: left alone the compiler re-derives
        // them per chunk from dynamically indexed kernel parameters, and the gather becomes the bottleneck of the phase.
        uint64_t abase = reinterpret_cast<uint64_t>(ph.A);
        uint32_t src0, dst0, step;
        {
          const uint32_t arb = (uint32_t)ph.a_row_bytes, r0 = (uint32_t)ptid >> 3, c = (uint32_t)ptid & 7u;
          const bool pairs = ph.a_pairs != 0;
          const uint32_t g0 = pairs ? ((r0 & 1u) ? (uint32_t)a.batch + (r0 >> 1) : (r0 >> 1)) : r0;
          src0 = g0 * arb + c * 16u;
          step = (pairs ? 8u : 16u) * arb;
          dst0 = sw128_offset(r0, c);
        }
        asm volatile("" : "+l"(abase), "+r"(src0), "+r"(dst0), "+r"(step));
        const int r0i = ptid >> 3;
        int issued = 0, arrived = 0;
        int ku = r.u0 % r.kbu;
        while (arrived < n_units) {
          bool can_issue = issued < n_units && (issued - arrived) < MAXPEND;
          if (can_issue) {
            const int gi = a_base + issued;
            uint64_t* eb = &empty_a[gi % SA];
            const uint32_t par = ((gi / SA) & 1) ^ 1;
            if (issued == arrived) mbar_wait(eb, par);
            else can_issue = mbar_test_wait(eb, par);
          }
          if (can_issue) {
            uint8_t* a_tile = a_ring + ((a_base + issued) % SA) * C::A_STAGE_BYTES;
            const uint8_t* src_k = reinterpret_cast<const uint8_t*>(abase) + (size_t)ku * 128 + src0;
            uint8_t* dst_k = a_tile + dst0;
#pragma unroll
            for (int j = 0; j < PER_THREAD; ++j)
              if (j * C::PROD_THREADS + ptid < CHUNKS && j * 16 + r0i < rows) cp_async16(dst_k + j * 2048, src_k + (size_t)j * step, 16u);
            cp_async_commit();
            ++issued;
            if (++ku == r.kbu) ku = 0;
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
            mbar_arrive(&full_a[(a_base + arrived) % SA]);
            ++arrived;
          }
        }
        // ---- every A stage of the phase has landed: take the second tile of the CTA's last fragment off the E group
        const int pair_last = (r.u1 - 1) / r.kbu, u_frag = max(r.u0, pair_last * r.kbu);
        const int kb_ = u_frag - pair_last * r.kbu, ke_ = r.u1 - pair_last * r.kbu;
        const int tiles = min(2, ph.n_tiles - 2 * pair_last);
        const int n_frag = pair_last - r.u0 / r.kbu + 1;
        if (ch_helped(ph, tiles, true, kb_ == 0 && ke_ == r.kbu)) {
          const int f = f_base + n_frag - 1, b = f & 1;
          mbar_wait(&tmem_full[b], (f >> 1) & 1);
          tc_fence_after();
          const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(b * 2 * SK_NT) + SK_NT;
          float acc[SK_NT];
          tmem_ld32(taddr, acc);
          tmem_ld32(taddr + 32, acc + 32);
          tmem_ld_wait();
          tc_fence_before();
          mbar_arrive(&tmem_empty_h[b]);
          ch_tile_epilogue(a, ph, quarter, lane, pair_last * 2 + 1, cta % ph.n_splits, acc, xpose + warp * 2048);
          ch_named_sync(3, 256);  // with the E group: all outputs of the CTA are written, E's thread 0 arrives for both
        }
        a_base += n_units, f_base += n_frag;
      }
      if (ph.fin != CH_FIN_NONE && cta < a.rows) {  // half of row `cta`'s reduction
        ch_named_sync(3, 256);  // E's thread 0 has seen every CTA's arrival
        if (ph.fin == CH_FIN_RESID_LN) ch_resid_ln_row(a, ph, cta, tid, s_red);
        else ch_qkv_row(a, ph, cta, tid);
        ch_named_sync(3, 256);
      }
    }
  } else {
    // ---------------------------------------------------------------- E group: epilogues (row == TMEM lane) + reductions
    pdl_wait_prior();  // residual stream, workspace, tickets and counters belong to the previous kernels until they are done
    if (cta == 0 && tid < CH_SYNC_INTS && a.sync_prev) a.sync_prev[tid] = 0;
    const int row = tid;
    const bool row_valid = row < a.rows;
    float4* ws4 = reinterpret_cast<float4*>(a.ws);
    int f = 0;
    for (int p = 0; p < a.n_phases; ++p) {
      const ChainPhase& ph = a.ph[p];
      const ChainRange r = ch_range(ph, cta);
      const int KBU = r.kbu, units = r.units, G = ph.grid;
      bool phase_helped = false;
      int u = r.u0;
      while (u < r.u1) {
        const int pair = u / KBU, ku_begin = u - pair * KBU;
        const int ku_end = min(KBU, ku_begin + (r.u1 - u));
        const int tiles = min(2, ph.n_tiles - 2 * pair);
        const int b = f & 1;
        const bool complete = (ku_begin == 0 && ku_end == KBU);
        const bool is_last = (u + (ku_end - ku_begin) == r.u1);
        const bool helped = ch_helped(ph, tiles, is_last, complete);
        const bool direct = (ph.epi == CH_EPI_PARTIAL) || complete;
        const int n_mine = helped ? 1 : tiles;
        phase_helped = helped;
        const int slot = (u == r.u0) ? 0 : 1;  // a CTA has at most two partial fragments per phase: its first and its last
        float4* mine = ws4 + ((size_t)cta * 2 + slot) * 2 * SK_PARTIAL_F4;
        mbar_wait(&tmem_full[b], (f >> 1) & 1);
        if (tid == 0 && is_last) ch_stamp(a, p, 3);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(b * 2 * SK_NT);
        float acc[SK_NT];
        for (int j = 0; j < n_mine; ++j) {
          tmem_ld32(taddr + j * SK_NT, acc);
          tmem_ld32(taddr + j * SK_NT + 32, acc + 32);
          tmem_ld_wait();
          if (j == n_mine - 1) {
            tc_fence_before();
            mbar_arrive(&tmem_empty[b]);  // the issuer may start the next-but-one fragment into this accumulator set
          }
          if (direct) {
            ch_tile_epilogue(a, ph, warp, lane, pair * 2 + j, cta % ph.n_splits, acc, xpose + warp * 2048);
          } else if (row_valid) {
#pragma unroll
            for (int c = 0; c < SK_NT / 4; ++c)
              mine[(j * (SK_NT / 4) + c) * 128 + row] = make_float4(acc[4 * c], acc[4 * c + 1], acc[4 * c + 2], acc[4 * c + 3]);
          }
        }
        if (!direct) {  // ticketed fix-up: the last contributor sums in CTA order
          __threadfence();
          ch_named_sync(1, 128);
          const int c_first = sk_owner(pair * KBU, units, G), c_last = sk_owner(pair * KBU + KBU - 1, units, G);
          if (tid == 0) {
            const int old = atomicAdd(&a.tickets[pair], 1);
            s_is_last = (old == c_last - c_first);
            if (s_is_last) a.tickets[pair] = 0;  // self-reset for the next phase / launch / graph replay
          }
          ch_named_sync(1, 128);
          if (s_is_last) {
            __threadfence();
            for (int j = 0; j < tiles; ++j) {
#pragma unroll
              for (int c = 0; c < SK_NT; ++c) acc[c] = 0.f;
              if (row_valid) {
                for (int cc = c_first; cc <= c_last; ++cc) {  // fixed order -> bitwise reproducible
                  const int sl = (sk_first_unit(cc, units, G) / KBU == pair) ? 0 : 1;
                  const float4* part = ws4 + (((size_t)cc * 2 + sl) * 2 + j) * SK_PARTIAL_F4;
#pragma unroll
                  for (int c = 0; c < SK_NT / 4; ++c) {
                    const float4 v = __ldcg(&part[c * 128 + row]);
                    acc[4 * c] += v.x, acc[4 * c + 1] += v.y, acc[4 * c + 2] += v.z, acc[4 * c + 3] += v.w;
                  }
                }
              }
              ch_tile_epilogue(a, ph, warp, lane, pair * 2 + j, 0, acc, xpose + warp * 2048);
            }
          }
          ch_named_sync(1, 128);  // s_is_last is rewritten by the next fragment
        }
        u += ku_end - ku_begin;
        ++f;
      }
      if (r.u1 > r.u0) {  // every output of this CTA for the phase is written: arrive on the phase's `done` counter
        if (phase_helped) ch_named_sync(3, 256);
        else ch_named_sync(1, 128);
        if (tid == 0) {
          ch_arrive(&a.sync_mine[2 * p]);
          ch_stamp(a, p, 4);
        }
      }
      if (ph.fin != CH_FIN_NONE && cta < a.rows) {  // CTA r reduces row r once all CTAs of the phase have arrived
        if (tid == 0) {
          ch_wait(&a.sync_mine[2 * p], ph.grid, a.err);
          ch_stamp(a, p, 5);
        }
        ch_named_sync(3, 256);
        if (ph.fin == CH_FIN_RESID_LN) ch_resid_ln_row(a, ph, cta, tid, s_red);
        else ch_qkv_row(a, ph, cta, tid);
        ch_named_sync(3, 256);
        if (tid == 0) {
          ch_arrive(&a.sync_mine[2 * p + 1]);
          ch_stamp(a, p, 6);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------ launch
template <int A_ROWS>
cudaError_t launch_chain_cfg(const ChainArgs& a, int grid, cudaStream_t st) {
  using C = ChainCfg<A_ROWS>;
  auto kern = chain_kernel<A_ROWS>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::DYN_BYTES);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  return launch_pdl(kern, dim3((unsigned)grid), dim3(C::THREADS), (size_t)C::DYN_BYTES, st, pdl_enabled(), a);
}
// grid: CTAs of the launch (<= SM count: all CTAs must be co-resident; every phase's `grid` must be <= it).
inline cudaError_t launch_chain(const ChainArgs& a, int grid, cudaStream_t st) {
  if (a.rows <= 8) return launch_chain_cfg<8>(a, grid, st);
  if (a.rows <= 32) return launch_chain_cfg<32>(a, grid, st);
  return launch_chain_cfg<128>(a, grid, st);
}

// Phase planning (host).  PARTIAL phases: pairs * splits CTAs, every CTA one fragment of one tile pair (n_tiles even).
inline void chain_plan_partial(ChainPhase& ph, int sms, int rows, int64_t ws_floats) {
  const int pairs = ph.n_tiles / 2, kbu = ph.k_blocks;
  int s = 1;
  for (int c = 1; c <= kbu && c <= 64; ++c)
    if (pairs * c <= sms && kbu / c >= 2 && (int64_t)c * rows * ph.ldn <= ws_floats) s = c;
  ph.n_splits = s, ph.grid = pairs * s;
}
// Ticketed phases: whole pairs per CTA when there are enough pairs, else each pair shared by <= 3 CTAs.
inline void chain_plan_ticketed(ChainPhase& ph, int sms) {
  const int pairs = (ph.n_tiles + 1) / 2;
  ph.n_splits = 1;
  if (pairs >= sms) {
    ph.grid = sms;  // unaligned stream-K: most pairs end up with two contributors
    for (int r = 1; r <= 2; ++r)
      if (pairs % r == 0 && pairs / r <= sms) {
        ph.grid = pairs / r;
        break;
      }
  } else {
    ph.grid = pairs * std::max(1, std::min(3, sms / pairs));
  }
}

}  // namespace zb
