// Persistent implicit-GEMM convolution of the DAC decoder on tcgen05 (one CTA per SM, static tile schedule).
//
//   out[t, n] = sum_cb sum_tap  X[t + shift(tap), cb*64 .. cb*64+63] . W[n][tap][cb*64 ..]        (fp16 in, fp32 accumulate)
//
//   * The activation "halo" tile of a channel block -- rows t0+smin .. t0+smin+127+span of X[b] -- is fetched ONCE by
//     a 3-D TMA tensor copy (box 64 ch x halo rows x 1 sequence, SWIZZLE_128B, rows outside [0,T) zero-filled by the
//     TMA unit: that is the convolution's zero padding, also at sequence boundaries inside the batch).  Every tap
//     reads it through a UMMA descriptor that starts `shift(tap) - smin` rows (x128 bytes) further down: the 128-byte
//     swizzle is a function of the absolute shared-memory address, so a row-shifted start is still a valid operand
//     (checked on the hardware by scripts/exp_umma_rowshift.cu).  A dilated k=7 conv therefore moves its input
//     through L2->SM once instead of seven times and needs no gather arithmetic at all.
//   * Weights are the pack.cu images ([n_tile][k_block = tap*cblocks + cb][NT*128 B]), streamed by 1-D bulk copies
//     through a ring; when the whole filter fits next to the activation ring (the 96- and 192-channel 1x1 convs, the
//     96-channel k=7 convs) it is loaded once per CTA and stays resident for all of the CTA's tiles.
//   * Two TMEM accumulators (2 x 256 columns): the epilogue of tile i (8 warps: bias, residual, fp16 rounding, Snake,
//     stores) overlaps the MMAs of tile i+1 (1 issuing thread) and the copies of tile i+2 (2 producer threads).
//
// Warp roles (352 threads): warps 0-7 epilogue (warp w: TMEM lanes 32*(w%4).., 16-column chunks of parity w/4),
// warp 8 activation TMA, warp 9 MMA issuer (+ TMEM alloc), warp 10 weight copies.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace zb {

typedef __half f16;

#ifndef DC_TE_WARPS
#define DC_TE_WARPS 12
#endif
// Epilogue variants (template MODE).  The epilogue is bound by the latency of its own arithmetic, so warps per scheduler
// are what speeds it up: 12 epilogue warps wherever the registers allow it.
//   DC_DIRECT_RESID  direct global loads / stores, residual rows held in registers (165 registers): 8 warps (fallback for
//                    channel counts the TMA boxes do not divide)
//   DC_TMA           residual / raw / activated boxes through the TMA unit: DC_TE_WARPS warps
//   DC_DIRECT        direct stores, no residual input (k = 7 convs, transposed convs): 12 warps
// Three more warps follow the epilogue warps: activation TMA, MMA issuer (+ TMEM alloc), weight copies.
constexpr int DC_DIRECT_RESID = 0, DC_TMA = 1, DC_DIRECT = 2;
template <int MODE>
struct DcCfg {
  static constexpr int EW = MODE == DC_TMA ? DC_TE_WARPS : (MODE == DC_DIRECT ? 12 : 8);
  static constexpr int THREADS = (EW + 3) * 32;
};
constexpr int DC_MAX_SA = 4;    // activation ring depth
constexpr int DC_PF_TILES = 3;  // L2 prefetch distance of the activation boxes, in tiles of this CTA
constexpr int DC_MAX_SW = 40;   // weight ring depth / resident k-blocks
constexpr int DC_SMEM_BYTES = 219 * 1024;  // dynamic; + 7.5 KB static epilogue constants < 227 KB
constexpr int DC_BAR_BYTES = (2 * DC_MAX_SA + 2 * DC_MAX_SW + 4) * 8 + 16;
// TMA epilogue (template flag TE): every epilogue warp owns a ring of DC_E_RS slots (residual box in, raw result out,
// in place) and DC_E_AS slots for the Snake-activated result; a slot is one 32-row x 32-channel fp16 box (SWIZZLE_64B).
constexpr int DC_E_RS = 3, DC_E_AS = 2, DC_E_SLOT = 2048;
constexpr int DC_E_WARP_BYTES = (DC_E_RS + DC_E_AS) * DC_E_SLOT;
constexpr int DC_E_BYTES = DC_TE_WARPS * DC_E_WARP_BYTES;
constexpr int DC_E_BAR_BYTES = DC_TE_WARPS * DC_E_RS * 8;
constexpr int DC_E_EXTRA = DC_E_BAR_BYTES + DC_E_BYTES;  // barriers + slots

struct DacConvArgs {
  const uint8_t* Wp;       // packed weights
  int n_tiles;             // N tiles of NT columns
  int tpc, n_groups;       // NT tiles per CTA tile (tpc*NT <= 256), ceil(n_tiles / tpc)
  int cblocks, taps;       // 64-channel blocks of the input, filter taps
  int shift0, shift_step;  // time shift of tap i = shift0 + i*shift_step
  int smin;                // smallest shift
  int halo_rows;           // 128 + (largest - smallest shift)
  int a_stage_bytes;       // halo_rows*128 rounded up to 1024
  int w_stage_bytes;       // tpc*NT*128
  int sa, sw;              // ring depths actually used
  int w_resident;          // 1: taps*cblocks <= sw and n_groups == 1 -> weights loaded once per CTA
  int tiles_per_seq, m_rows, total_tiles;
  // epilogue
  const float* bias;       // [n_total]
  const f16* resid;        // optional [B][T_out][pitch_out]  (plain convs only)
  f16* out_raw;            // optional
  f16* out_act;            // optional: Snake with the consumer's alpha
  const float* alpha;      // [C_out]
  const float* inv_alpha;  // [C_out] = 1/(alpha + 1e-9)
  int T_out, pitch_out, C_out, n_total;
  int up, pad;             // transposed conv: stride and padding (up == 0: plain conv)
  float* wav;              // last conv (direct epilogue only): tanh of the fp16-rounded result as fp32 [B][T_out][C_out]
  long long* trace;        // SK_TRACE builds: [grid][8] cycle counters
};

// 3-D tiled TMA load (SASS: UTMALDG), completion counted in bytes on an mbarrier.  Coordinates are signed: elements
// outside the tensor are written as zeros.
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
      : "memory");
}

// 3-D tiled TMA store shared -> global (SASS: UTMASTG), bulk-group completion; elements outside the tensor are dropped.
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, uint32_t smem_src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(map)),
               "r"(smem_src), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void bulk_commit_group() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_group_read() {  // at most N groups still reading their shared-memory source
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_group_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// Snake1d (modeling_dac.py:94-99): x + 1/(alpha+1e-9) * sin(alpha*x)^2 = x + h - h*cos(k*x) with k = 2*alpha and
// h = 0.5/(alpha+1e-9) (both precomputed per channel): one multiply, the SFU cosine (which reduces its argument by
// taking the fraction of k*x/2pi; abs. error ~5e-7), one add and one FMA.  The result is rounded to fp16 by the caller:
// the absolute error h*5e-7 is below half an fp16 ulp for every |result| > 1e-3.
__device__ __forceinline__ float snake_cos(float x, float k, float h) { return fmaf(-h, __cosf(k * x), x + h); }

// Eight columns of one output row: + bias and fp16 rounding (the conv's fp16 output under autocast), + residual in
// fp16 (HADD2: the fp32 sum of two fp16 numbers rounds to the same fp16 value), Snake in fp32, fp16 again.
__device__ __forceinline__ void dc_group8(const float* acc, const float4& b0, const float4& b1, const float4& k0,
                                          const float4& k1, const float4& h0, const float4& h1, bool has_resid,
                                          const uint4& resid, bool has_act, uint4& raw, uint4& act) {
  const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
  const float kk[8] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w};
  const float hh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
  __half2* rw = reinterpret_cast<__half2*>(&raw);
  __half2* ac = reinterpret_cast<__half2*>(&act);
  const __half2* rs = reinterpret_cast<const __half2*>(&resid);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    __half2 v = __floats2half2_rn(acc[2 * i] + bb[2 * i], acc[2 * i + 1] + bb[2 * i + 1]);
    if (has_resid) v = __hadd2(v, rs[i]);
    rw[i] = v;
    if (has_act) {
      const float2 f = __half22float2(v);
      ac[i] = __floats2half2_rn(snake_cos(f.x, kk[2 * i], hh[2 * i]), snake_cos(f.y, kk[2 * i + 1], hh[2 * i + 1]));
    }
  }
}

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// Same box, L2 only (SASS: UTMAPF): pulls a later tile's activations out of HBM while the ring is still occupied.
__device__ __forceinline__ void tma_prefetch_3d(const CUtensorMap* map, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global.tile [%0, {%1, %2, %3}];" ::"l"(reinterpret_cast<uint64_t>(map)),
               "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}

#ifdef SK_TRACE
#define DC_TIMED_WAIT(counter, bar, parity) \
  do {                                      \
    const long long _t0 = clock64();        \
    mbar_wait(bar, parity);                 \
    counter += clock64() - _t0;             \
  } while (0)
#else
#define DC_TIMED_WAIT(counter, bar, parity) mbar_wait(bar, parity)
#endif
#ifdef SK_TRACE
#define DC_TIMED_WAIT_RELAXED(counter, bar, parity) \
  do {                                              \
    const long long _t0 = clock64();                \
    mbar_wait_relaxed(bar, parity);                 \
    counter += clock64() - _t0;                     \
  } while (0)
#else
#define DC_TIMED_WAIT_RELAXED(counter, bar, parity) mbar_wait_relaxed(bar, parity)
#endif

__device__ __forceinline__ void dc_tile_coords(const DacConvArgs& a, int tile, int& m_tile, int& group) {
  m_tile = tile / a.n_groups;
  group = tile - m_tile * a.n_groups;
}

// TE: epilogue traffic through the TMA unit (plain convs: residual boxes in, raw / activated boxes out); the maps are
// [batch][T_out][C_out] fp16 with a 32 x 32 x 1 box.  TE == false: direct global loads / stores from the epilogue threads.
template <int NT, int MODE>
__global__ void __launch_bounds__(DcCfg<MODE>::THREADS, 1)
dac_conv_kernel(const __grid_constant__ CUtensorMap xmap, const __grid_constant__ CUtensorMap rmap,
                const __grid_constant__ CUtensorMap rawmap, const __grid_constant__ CUtensorMap actmap, const DacConvArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_ring = smem;
  uint8_t* w_ring = smem + a.sa * a.a_stage_bytes;
  // (TE) epilogue slots right after the rings (all ring stages are multiples of 1 KB, so the slots stay 1 KB aligned),
  // then the barriers
  uint8_t* e_buf = w_ring + a.sw * a.w_stage_bytes;                                   // TE: [warp][slots]
  uint64_t* a_full = reinterpret_cast<uint64_t*>(e_buf + (MODE == DC_TMA ? DC_E_BYTES : 0));
  uint64_t* a_empty = a_full + DC_MAX_SA;
  uint64_t* w_full = a_empty + DC_MAX_SA;
  uint64_t* w_empty = w_full + DC_MAX_SW;
  uint64_t* t_full = w_empty + DC_MAX_SW;
  uint64_t* t_empty = t_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);
  uint64_t* e_full = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(a_full) + DC_BAR_BYTES);  // TE: [warp][DC_E_RS]

  // per-column epilogue constants of the current column group (two copies: tiles alternate when n_groups > 1)
  __shared__ __align__(16) float s_bias[2][256], s_alpha[2][256], s_inv[2][256];
  __shared__ int2 s_grp[2][32];  // per 8 columns: (time offset r - pad of the transposed conv's phase, output channel)

  constexpr int EW = DcCfg<MODE>::EW;
  constexpr bool TE = MODE == DC_TMA, RESID = MODE == DC_DIRECT_RESID;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kb_per_tile = a.cblocks * a.taps;

  if (tid == 0) {
    for (int s = 0; s < DC_MAX_SA; ++s) mbar_init(&a_full[s], 1), mbar_init(&a_empty[s], 1);
    for (int s = 0; s < DC_MAX_SW; ++s) mbar_init(&w_full[s], 1), mbar_init(&w_empty[s], 1);
    for (int s = 0; s < 2; ++s) mbar_init(&t_full[s], 1), mbar_init(&t_empty[s], EW);
    if constexpr (TE)
      for (int s = 0; s < EW * DC_E_RS; ++s) mbar_init(&e_full[s], 1);
    fence_barrier_init();
  }
  if (warp == EW + 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == EW) {
    // ---------------------------------------------------------------- activation halo tiles: one TMA per channel block
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&xmap)) : "memory");
      int s = 0;
      uint32_t empty_parity = 1;  // ring slot and the parity of its "empty" barrier (flips each time the ring wraps)
      [[maybe_unused]] long long tr_wait0 = 0, tr_start = clock64();
      // The ring holds at most one or two tiles' worth of halo boxes when the weights are resident, far less than the
      // HBM latency-bandwidth product: boxes of the tiles DC_PF_TILES ahead are prefetched into L2.
      auto prefetch_tile = [&](int tile) {
        if (tile >= a.total_tiles) return;
        int m_tile, group;
        dc_tile_coords(a, tile, m_tile, group);
        if (a.n_groups > 1 && group != 0) return;  // the other column groups of an m-tile find it in L2 anyway
        const int b = m_tile / a.tiles_per_seq, t0 = (m_tile - b * a.tiles_per_seq) * 128;
        for (int cb = 0; cb < a.cblocks; ++cb) tma_prefetch_3d(&xmap, cb * 64, t0 + a.smin, b);
      };
      for (int k = 1; k < DC_PF_TILES; ++k) prefetch_tile(blockIdx.x + k * gridDim.x);
      for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x) {
        prefetch_tile(tile + DC_PF_TILES * gridDim.x);
        int m_tile, group;
        dc_tile_coords(a, tile, m_tile, group);
        const int b = m_tile / a.tiles_per_seq, t0 = (m_tile - b * a.tiles_per_seq) * 128;
        for (int cb = 0; cb < a.cblocks; ++cb) {
          DC_TIMED_WAIT(tr_wait0, &a_empty[s], empty_parity);
          mbar_arrive_expect_tx(&a_full[s], a.halo_rows * 128);
          tma_load_3d(a_ring + s * a.a_stage_bytes, &xmap, cb * 64, t0 + a.smin, b, &a_full[s]);
          if (++s == a.sa) s = 0, empty_parity ^= 1;
        }
      }
#ifdef SK_TRACE
      if (a.trace) a.trace[blockIdx.x * 8 + 6] = tr_wait0, a.trace[blockIdx.x * 8 + 7] = clock64() - tr_start;
#endif
    }
  } else if (warp == EW + 2) {
    // ---------------------------------------------------------------- weight k-blocks, in the order the MMAs use them
    // Lane jt copies NT tile jt of every stage: one thread only gets a few bulk copies per microsecond through the
    // TMA unit (scripts/exp_stream.cu), a 256-column stage is four of them.
    if (lane < 4) {
      int s = 0;
      uint32_t empty_parity = 1;  // ring slot and the parity of its "empty" barrier (flips each time the ring wraps)
      const size_t tile_stride = (size_t)kb_per_tile * (NT * 128);  // bytes between consecutive NT tiles of one k-block
      for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x) {
        const int group = a.n_groups == 1 ? 0 : tile % a.n_groups;
        const int n_tile0 = group * a.tpc;
        const int tiles_here = min(a.tpc, a.n_tiles - n_tile0);
        const uint32_t stage_tx = tiles_here * NT * 128;
        const uint8_t* wg = a.Wp + (size_t)n_tile0 * tile_stride;
        for (int cb = 0; cb < a.cblocks; ++cb) {
          const uint8_t* src = wg + (size_t)cb * (NT * 128);  // k-block index = tap*cblocks + cb
          for (int tap = 0; tap < a.taps; ++tap, src += (size_t)a.cblocks * (NT * 128)) {
            if (!a.w_resident) mbar_wait(&w_empty[s], empty_parity);
            if (lane == 0) mbar_arrive_expect_tx(&w_full[s], stage_tx);
            if (lane < tiles_here)
              bulk_g2s(w_ring + s * a.w_stage_bytes + lane * NT * 128, src + lane * tile_stride, NT * 128, &w_full[s]);
            if (++s == a.sw) s = 0, empty_parity ^= 1;
          }
        }
        if (a.w_resident) break;  // the whole filter is in shared memory now and stays
      }
    }
  } else if (warp == EW + 1) {
    // ---------------------------------------------------------------- MMA issuer
    // The whole warp runs the loops (warp-uniform control flow keeps the descriptors in uniform registers and the
    // issue path short: the single issuing thread is the pacing resource of the small-N layers); one elected lane
    // issues the MMAs and commits.
    {
      int sa = 0, sw = 0, it = 0;
      uint32_t a_parity = 0, w_parity = 0;
      const uint32_t a_ring_addr = smem_u32(a_ring), w_ring_addr = smem_u32(w_ring);
      const uint32_t tap0_off = (a.shift0 - a.smin) * 128;
      const int tap_step = a.shift_step * 128;
      const bool leader = elect_one();
      [[maybe_unused]] long long tr_te = 0, tr_af = 0, tr_wf = 0, tr_start = clock64();
      for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
        const int group = a.n_groups == 1 ? 0 : tile % a.n_groups;
        const int tiles_here = min(a.tpc, a.n_tiles - group * a.tpc);
        const uint32_t idesc = umma_idesc(128, tiles_here * NT, 0);
        const int buf = it & 1;
        DC_TIMED_WAIT(tr_te, &t_empty[buf], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * 256;
        uint32_t accumulate = 0;
        for (int cb = 0; cb < a.cblocks; ++cb) {
          DC_TIMED_WAIT(tr_af, &a_full[sa], a_parity);
          uint32_t a_addr = a_ring_addr + sa * a.a_stage_bytes + tap0_off;
          for (int tap = 0; tap < a.taps; ++tap, a_addr += tap_step) {
            if (!a.w_resident || it == 0) DC_TIMED_WAIT(tr_wf, &w_full[sw], w_parity);
            tc_fence_after();
            if (leader) {
              const uint64_t adesc = umma_desc_sw128(a_addr);
              const uint64_t bdesc = umma_desc_sw128(w_ring_addr + sw * a.w_stage_bytes);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                umma_f16(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, accumulate);
                accumulate = 1;
              }
              if (!a.w_resident) umma_commit(&w_empty[sw]);
            }
            __syncwarp();
            if (++sw == a.sw) sw = 0, w_parity ^= 1;
          }
          if (leader) umma_commit(&a_empty[sa]);
          __syncwarp();
          if (++sa == a.sa) sa = 0, a_parity ^= 1;
        }
        if (leader) umma_commit(&t_full[buf]);
        __syncwarp();
      }
#ifdef SK_TRACE
      if (a.trace && leader) {
        long long* o = a.trace + blockIdx.x * 8;
        o[0] = tr_te, o[1] = tr_af, o[2] = tr_wf, o[3] = clock64() - tr_start;
      }
#endif
    }
  } else if constexpr (TE) {
    // ---------------------------------------------------------------- epilogue through the TMA unit (plain convs)
    // Warp w owns rows 32*(w%4).. of every tile and every (EW/4)-th 32-channel block of the CTA's block sequence (tiles
    // in order, blocks in order inside a tile; every tile has NB blocks).  Per block: the residual box arrives in a
    // ring slot two blocks ahead, the thread (= row) reads its 64 bytes, writes the raw result back in place and the
    // activated one to a second slot; one lane stores both boxes with the TMA unit and refills the slot whose store
    // has finished reading.  No global access goes through the LSU.
    constexpr int STRIDE = EW / 4;
    const int q = warp & 3, par = warp >> 2;
    const uint32_t my = smem_u32(e_buf + warp * DC_E_WARP_BYTES);
    uint64_t* my_full = e_full + warp * DC_E_RS;
    const bool has_resid = a.resid != nullptr, has_raw = a.out_raw != nullptr, has_act = a.out_act != nullptr;
    const uint32_t row_off = lane * 64, sw_x = (lane >> 1) & 3;  // SWIZZLE_64B: 16-byte chunk ^= (row >> 1) & 3
    const int NB = a.tpc * (NT / 32);                            // 32-channel blocks per tile (all column groups alike)
    const int n_my_tiles = (a.total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    struct Pos {
      int b, t, n0;
    };
    auto tile_pos = [&](int it_) {  // box origin of this warp's rows in the CTA's it_-th tile
      int m_tile, group;
      dc_tile_coords(a, blockIdx.x + it_ * gridDim.x, m_tile, group);
      Pos p;
      p.b = m_tile / a.tiles_per_seq;
      p.t = (m_tile - p.b * a.tiles_per_seq) * 128 + q * 32;
      p.n0 = group * a.tpc * NT;
      return p;
    };
    // prefetch cursor (lane 0 only): block n_issued of this warp = (pf_it, pf_cbk)
    int pf_it = 0, pf_cbk = par, n_issued = 0;
    while (pf_cbk >= NB) pf_cbk -= NB, ++pf_it;
    Pos pf_pos = tile_pos(pf_it < n_my_tiles ? pf_it : 0);
    auto issue_next_load = [&]() {  // lane 0
      if (has_resid && pf_it < n_my_tiles) {
        const int slot = n_issued % DC_E_RS;
        mbar_arrive_expect_tx(&my_full[slot], DC_E_SLOT);
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                my + slot * DC_E_SLOT),
            "l"(reinterpret_cast<uint64_t>(&rmap)), "r"(pf_pos.n0 + pf_cbk * 32), "r"(pf_pos.t), "r"(pf_pos.b),
            "r"(smem_u32(&my_full[slot]))
            : "memory");
      }
      ++n_issued, pf_cbk += STRIDE;
      if (pf_cbk >= NB) {
        while (pf_cbk >= NB) pf_cbk -= NB, ++pf_it;
        if (pf_it < n_my_tiles) pf_pos = tile_pos(pf_it);
      }
    };
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&rmap)) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&rawmap)) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&actmap)) : "memory");
      for (int k = 0; k < DC_E_RS - 1; ++k) issue_next_load();
    }
    int cbk = par, idx = 0;  // this warp's next block is block `cbk` of tile `it` (cbk may run past NB: next tile)
    [[maybe_unused]] long long tr_tf = 0, tr_start = clock64();
    for (int it = 0; it < n_my_tiles; ++it) {
      const int buf = it & 1;
      const int cbuf = a.n_groups == 1 ? 0 : buf;
      const Pos pos = tile_pos(it);
      if (a.n_groups > 1 || it == 0) {
        const int col = tid, n = pos.n0 + col;
        if (col < NB * 32) {
          const bool ok = n < a.n_total;
          s_bias[cbuf][col] = ok ? __ldg(a.bias + n) : 0.f;
          s_alpha[cbuf][col] = ok && has_act ? 2.f * __ldg(a.alpha + n) : 0.f;
          s_inv[cbuf][col] = ok && has_act ? 0.5f * __ldg(a.inv_alpha + n) : 0.f;
        }
        asm volatile("bar.sync 1, %0;" ::"r"(EW * 32) : "memory");
      }
      DC_TIMED_WAIT(tr_tf, &t_full[buf], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * 256 + ((uint32_t)(q * 32) << 16);
      if (cbk >= NB) {  // no block of this tile is this warp's
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&t_empty[buf]);
      }
      for (; cbk < NB; cbk += STRIDE, ++idx) {
        const uint32_t rs = my + (idx % DC_E_RS) * DC_E_SLOT + row_off;
        const uint32_t as = my + (DC_E_RS + idx % DC_E_AS) * DC_E_SLOT + row_off;
        float acc[32];
        tmem_ld32(taddr + cbk * 32, acc);
        // per-column constants of the block's first 8 columns (in flight with the TMEM load and the residual wait)
        const float* cb = &s_bias[cbuf][cbk * 32];
        const float* ck = &s_alpha[cbuf][cbk * 32];
        const float* ch = &s_inv[cbuf][cbk * 32];
        uint4 res[4];
        if (has_resid) {
          mbar_wait(&my_full[idx % DC_E_RS], (idx / DC_E_RS) & 1);
#pragma unroll
          for (int g = 0; g < 4; ++g) res[g] = lds128(rs + ((uint32_t)(g ^ sw_x) << 4));
        }
        tmem_ld_wait();
        if (cbk + STRIDE >= NB) {  // last TMEM read of this warp in this tile: hand the accumulator back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&t_empty[buf]);
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const float4 b0 = *reinterpret_cast<const float4*>(cb + g * 8), b1 = *reinterpret_cast<const float4*>(cb + g * 8 + 4);
          const float4 k0 = *reinterpret_cast<const float4*>(ck + g * 8), k1 = *reinterpret_cast<const float4*>(ck + g * 8 + 4);
          const float4 h0 = *reinterpret_cast<const float4*>(ch + g * 8), h1 = *reinterpret_cast<const float4*>(ch + g * 8 + 4);
          uint4 raw, act;
          dc_group8(acc + g * 8, b0, b1, k0, k1, h0, h1, has_resid, res[g], has_act, raw, act);
          const uint32_t off = (uint32_t)(g ^ sw_x) << 4;
          if (has_raw) sts128(rs + off, raw);
          if (has_act) sts128(as + off, act);
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          const int c0 = pos.n0 + cbk * 32;
          if (has_raw) tma_store_3d(&rawmap, rs - row_off, c0, pos.t, pos.b);
          if (has_act) tma_store_3d(&actmap, as - row_off, c0, pos.t, pos.b);
          bulk_commit_group();
          bulk_wait_group_read<1>();  // the previous block's boxes have left shared memory: its slots are free
          issue_next_load();
        }
        __syncwarp();
      }
      cbk -= NB;
    }
    if (lane == 0) bulk_wait_group_all();
#ifdef SK_TRACE
    if (a.trace && tid == 0) a.trace[blockIdx.x * 8 + 4] = tr_tf, a.trace[blockIdx.x * 8 + 5] = clock64() - tr_start;
#endif
    tc_fence_before();
  } else {
    // ---------------------------------------------------------------- epilogue: TMEM -> bias/residual/Snake -> fp16 rows
    constexpr int STRIDE = EW / 4;  // warps per row quarter: warp (q, par) takes the 16-column chunks par, par+STRIDE, ...
    const int q = warp & 3, par = warp >> 2;
    const int row = q * 32 + lane;
    const bool has_resid = RESID && a.resid != nullptr && a.up == 0;
    // a 16-column chunk of a row is one aligned 32-byte piece of the output when the channel count is a multiple of 16
    const bool wide = (a.C_out & 15) == 0 && (a.pitch_out & 15) == 0;
    int it = 0;
    [[maybe_unused]] long long tr_tf = 0, tr_start = clock64();
    for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
      int m_tile, group;
      dc_tile_coords(a, tile, m_tile, group);
      const int n0 = group * a.tpc * NT;
      const int n_cols = min(a.tpc, a.n_tiles - group * a.tpc) * NT;
      const int b = m_tile / a.tiles_per_seq, j = (m_tile - b * a.tiles_per_seq) * 128 + row;
      const bool row_valid = j < a.m_rows;
      const int buf = it & 1;
      const int n_chunks = n_cols >> 4;
      const int n_mine = (n_chunks - par + STRIDE - 1) / STRIDE;
      // the residual row is streamed from HBM: issue this thread's loads before waiting for the accumulator
      uint4 rbuf[RESID ? 16 : 2];
      if (has_resid && row_valid) {
        const f16* rrow = a.resid + ((size_t)b * a.T_out + j) * a.pitch_out + n0;
#pragma unroll
        for (int c = 0; c < (RESID ? 8 : 0); ++c) {
          const int ch = 2 * c + par;
          if (ch < n_chunks && n0 + ch * 16 < a.n_total) {
            if (wide) {
              ldg256_stream(rrow + ch * 16, rbuf[2 * c], rbuf[2 * c + 1]);
            } else {
              rbuf[2 * c] = __ldcs(reinterpret_cast<const uint4*>(rrow + ch * 16));
              if (n0 + ch * 16 + 8 < a.n_total) rbuf[2 * c + 1] = __ldcs(reinterpret_cast<const uint4*>(rrow + ch * 16 + 8));
            }
          }
        }
      }
      // per-column constants of this tile's column group -> shared memory (the L1 is a few KB next to 200+ KB of
      // rings and is flushed by the streaming stores, so per-use global loads would each go to L2)
      const int cbuf = a.n_groups == 1 ? 0 : buf;
      if (a.n_groups > 1 || it == 0) {
        const int col = tid, n = n0 + col;
        if (col < n_cols) {
          const bool ok = n < a.n_total;
          int co = n, r = 0;
          if (a.up > 0 && ok) r = n / a.C_out, co = n - r * a.C_out;
          s_bias[cbuf][col] = ok ? __ldg(a.bias + n) : 0.f;
          s_alpha[cbuf][col] = ok && a.out_act ? 2.f * __ldg(a.alpha + co) : 0.f;       // k = 2*alpha    (snake_cos)
          s_inv[cbuf][col] = ok && a.out_act ? 0.5f * __ldg(a.inv_alpha + co) : 0.f;    // h = 0.5/(alpha+1e-9)
          if ((col & 7) == 0) s_grp[cbuf][col >> 3] = make_int2(ok ? (a.up > 0 ? r - a.pad : 0) : (1 << 30), co);
        }
        asm volatile("bar.sync 1, %0;" ::"r"(EW * 32) : "memory");
      }
      DC_TIMED_WAIT(tr_tf, &t_full[buf], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * 256 + ((uint32_t)(q * 32) << 16);
      if (n_mine <= 0) {  // (the 16-column tile of the last conv: one warp per row quarter has work)
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&t_empty[buf]);
        continue;
      }
      float acc[2][16];
      tmem_ld16(taddr + par * 16, acc[0]);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        if (c >= n_mine) break;
        const int ch = STRIDE * c + par;
        tmem_ld_wait();
        if (c + 1 < n_mine) {
          tmem_ld16(taddr + (ch + STRIDE) * 16, acc[(c + 1) & 1]);  // in flight while this chunk is processed
        } else {  // last TMEM read of this warp: hand the accumulator back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&t_empty[buf]);
        }
        if (!row_valid) continue;
        uint4 o_raw[2], o_act[2];
        size_t offs[2];
        bool live[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int col = ch * 16 + h * 8;
          const int2 grp = s_grp[cbuf][col >> 3];
          const int t = a.up > 0 ? a.up * j + grp.x : j + grp.x;
          live[h] = t >= 0 && t < a.T_out;
          offs[h] = ((size_t)b * a.T_out + (live[h] ? t : 0)) * a.pitch_out + grp.y;
          if (!live[h]) continue;
          const float4 b0 = *reinterpret_cast<const float4*>(&s_bias[cbuf][col]);
          const float4 b1 = *reinterpret_cast<const float4*>(&s_bias[cbuf][col + 4]);
          const float4 k0 = *reinterpret_cast<const float4*>(&s_alpha[cbuf][col]);
          const float4 k1 = *reinterpret_cast<const float4*>(&s_alpha[cbuf][col + 4]);
          const float4 h0 = *reinterpret_cast<const float4*>(&s_inv[cbuf][col]);
          const float4 h1 = *reinterpret_cast<const float4*>(&s_inv[cbuf][col + 4]);
          dc_group8(&acc[c & 1][h * 8], b0, b1, k0, k1, h0, h1, has_resid, rbuf[RESID ? 2 * c + h : 0], a.out_act != nullptr,
                    o_raw[h], o_act[h]);
        }
        if (wide && live[0] && live[1]) {  // (both halves are in one phase / row: C_out % 16 == 0)
          if (a.out_raw) stg256(a.out_raw + offs[0], o_raw[0], o_raw[1]);
          if (a.out_act) stg256(a.out_act + offs[0], o_act[0], o_act[1]);
        } else {
#pragma unroll
          for (int h = 0; h < 2; ++h)
            if (live[h]) {
              if (a.wav) {  // modeling_dac.py:436-437: tanh of the (fp16) conv output
                const __half2* rh = reinterpret_cast<const __half2*>(&o_raw[h]);
                float w8[8];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const float2 f = __half22float2(rh[i]);
                  w8[2 * i] = tanhf(f.x), w8[2 * i + 1] = tanhf(f.y);
                }
                float4* dst = reinterpret_cast<float4*>(a.wav + offs[h]);
                dst[0] = make_float4(w8[0], w8[1], w8[2], w8[3]);
                dst[1] = make_float4(w8[4], w8[5], w8[6], w8[7]);
              }
              if (a.out_raw) *reinterpret_cast<uint4*>(a.out_raw + offs[h]) = o_raw[h];
              if (a.out_act) *reinterpret_cast<uint4*>(a.out_act + offs[h]) = o_act[h];
            }
        }
      }
    }
#ifdef SK_TRACE
    if (a.trace && tid == 0) a.trace[blockIdx.x * 8 + 4] = tr_tf, a.trace[blockIdx.x * 8 + 5] = clock64() - tr_start;
#endif
    tc_fence_before();
  }
  __syncthreads();
  if (warp == EW + 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace zb
