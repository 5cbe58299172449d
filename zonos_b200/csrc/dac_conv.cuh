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

constexpr int DC_THREADS = 352;
constexpr int DC_EPI_THREADS = 256;
constexpr int DC_MAX_SA = 4;    // activation ring depth
constexpr int DC_MAX_SW = 16;   // weight ring depth / resident k-blocks
constexpr int DC_SMEM_BYTES = 220 * 1024;
constexpr int DC_BAR_BYTES = (2 * DC_MAX_SA + 2 * DC_MAX_SW + 4) * 8 + 16;

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

// Snake1d (modeling_dac.py:94-99): x + 1/(alpha+1e-9) * sin(alpha*x)^2.  sin^2 has period pi: reduce alpha*x to
// [-pi/2, pi/2] with a two-constant Cody-Waite step, then the SFU sine (abs. error ~5e-7 there) -- the result is
// rounded to fp16 by the caller, so this is well inside half an fp16 ulp of the libm path and costs ~8 instructions.
__device__ __forceinline__ float snake_fast(float x, float alpha, float inv_alpha) {
  const float y = alpha * x;
  const float k = rintf(y * 0.318309886183790672f);
  float r = fmaf(k, -3.14159274101257324f, y);    // pi rounded to fp32
  r = fmaf(k, 8.74227765734758577e-8f, r);        // fp32(pi) - pi
  const float s = __sinf(r);
  return fmaf(inv_alpha, s * s, x);
}

__device__ __forceinline__ void dc_tile_coords(const DacConvArgs& a, int tile, int& m_tile, int& group) {
  m_tile = tile / a.n_groups;
  group = tile - m_tile * a.n_groups;
}

template <int NT>
__global__ void __launch_bounds__(DC_THREADS, 1)
dac_conv_kernel(const __grid_constant__ CUtensorMap xmap, const DacConvArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_ring = smem;
  uint8_t* w_ring = smem + a.sa * a.a_stage_bytes;
  uint64_t* a_full = reinterpret_cast<uint64_t*>(w_ring + a.sw * a.w_stage_bytes);
  uint64_t* a_empty = a_full + DC_MAX_SA;
  uint64_t* w_full = a_empty + DC_MAX_SA;
  uint64_t* w_empty = w_full + DC_MAX_SW;
  uint64_t* t_full = w_empty + DC_MAX_SW;
  uint64_t* t_empty = t_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kb_per_tile = a.cblocks * a.taps;

  if (tid == 0) {
    for (int s = 0; s < DC_MAX_SA; ++s) mbar_init(&a_full[s], 1), mbar_init(&a_empty[s], 1);
    for (int s = 0; s < DC_MAX_SW; ++s) mbar_init(&w_full[s], 1), mbar_init(&w_empty[s], 1);
    for (int s = 0; s < 2; ++s) mbar_init(&t_full[s], 1), mbar_init(&t_empty[s], DC_EPI_THREADS / 32);
    fence_barrier_init();
  }
  if (warp == 9) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 8) {
    // ---------------------------------------------------------------- activation halo tiles: one TMA per channel block
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&xmap)) : "memory");
      int g = 0;  // running stage counter
      for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x) {
        int m_tile, group;
        dc_tile_coords(a, tile, m_tile, group);
        const int b = m_tile / a.tiles_per_seq, t0 = (m_tile - b * a.tiles_per_seq) * 128;
        for (int cb = 0; cb < a.cblocks; ++cb, ++g) {
          const int s = g % a.sa;
          mbar_wait(&a_empty[s], ((g / a.sa) & 1) ^ 1);
          mbar_arrive_expect_tx(&a_full[s], a.halo_rows * 128);
          tma_load_3d(a_ring + s * a.a_stage_bytes, &xmap, cb * 64, t0 + a.smin, b, &a_full[s]);
        }
      }
    }
  } else if (warp == 10) {
    // ---------------------------------------------------------------- weight k-blocks, in the order the MMAs use them
    if (lane == 0) {
      int g = 0;
      bool first = true;
      for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x) {
        if (a.w_resident && !first) break;
        first = false;
        int m_tile, group;
        dc_tile_coords(a, tile, m_tile, group);
        const int n_tile0 = group * a.tpc;
        const int tiles_here = min(a.tpc, a.n_tiles - n_tile0);
        for (int cb = 0; cb < a.cblocks; ++cb)
          for (int tap = 0; tap < a.taps; ++tap, ++g) {
            const int s = g % a.sw;
            if (!a.w_resident) mbar_wait(&w_empty[s], ((g / a.sw) & 1) ^ 1);
            mbar_arrive_expect_tx(&w_full[s], tiles_here * NT * 128);
            const int kb = tap * a.cblocks + cb;
            for (int jt = 0; jt < tiles_here; ++jt)
              bulk_g2s(w_ring + s * a.w_stage_bytes + jt * NT * 128,
                       a.Wp + ((size_t)(n_tile0 + jt) * kb_per_tile + kb) * (NT * 128), NT * 128, &w_full[s]);
          }
      }
    }
  } else if (warp == 9) {
    // ---------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      int ga = 0, gw = 0, it = 0;
      for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
        int m_tile, group;
        dc_tile_coords(a, tile, m_tile, group);
        const int tiles_here = min(a.tpc, a.n_tiles - group * a.tpc);
        const uint32_t idesc = umma_idesc(128, tiles_here * NT, 0);
        const int buf = it & 1;
        mbar_wait(&t_empty[buf], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * 256;
        uint32_t accumulate = 0;
        for (int cb = 0; cb < a.cblocks; ++cb, ++ga) {
          const int sa = ga % a.sa;
          mbar_wait(&a_full[sa], (ga / a.sa) & 1);
          const uint32_t a_addr = smem_u32(a_ring + sa * a.a_stage_bytes);
          for (int tap = 0; tap < a.taps; ++tap, ++gw) {
            const int sw = gw % a.sw;
            if (!a.w_resident || it == 0) mbar_wait(&w_full[sw], (gw / a.sw) & 1);
            tc_fence_after();
            const uint64_t adesc = umma_desc_sw128(a_addr + (a.shift0 + tap * a.shift_step - a.smin) * 128);
            const uint64_t bdesc = umma_desc_sw128(smem_u32(w_ring + sw * a.w_stage_bytes));
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              umma_f16(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, accumulate);
              accumulate = 1;
            }
            if (!a.w_resident) umma_commit(&w_empty[sw]);
          }
          umma_commit(&a_empty[sa]);
        }
        umma_commit(&t_full[buf]);
      }
    }
  } else {
    // ---------------------------------------------------------------- epilogue: TMEM -> bias/residual/Snake -> fp16 rows
    const int q = warp & 3, par = warp >> 2;
    const int row = q * 32 + lane;
    int it = 0;
    for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
      int m_tile, group;
      dc_tile_coords(a, tile, m_tile, group);
      const int n0 = group * a.tpc * NT;
      const int n_cols = min(a.tpc, a.n_tiles - group * a.tpc) * NT;
      const int b = m_tile / a.tiles_per_seq, j = (m_tile - b * a.tiles_per_seq) * 128 + row;
      const bool row_valid = j < a.m_rows;
      const int buf = it & 1;
      const int n_chunks = n_cols >> 4;
      // the residual row is streamed from HBM: issue this thread's loads before waiting for the accumulator
      uint4 rbuf[16];
      const bool has_resid = a.resid != nullptr && a.up == 0;
      if (has_resid && row_valid) {
        const f16* rrow = a.resid + ((size_t)b * a.T_out + j) * a.pitch_out + n0;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const int ch = 2 * c + par;
          if (ch < n_chunks && n0 + ch * 16 < a.n_total) {
            rbuf[2 * c] = __ldcs(reinterpret_cast<const uint4*>(rrow + ch * 16));
            rbuf[2 * c + 1] = __ldcs(reinterpret_cast<const uint4*>(rrow + ch * 16 + 8));
          }
        }
      }
      mbar_wait(&t_full[buf], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * 256 + ((uint32_t)(q * 32) << 16);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const int ch = 2 * c + par;
        if (ch >= n_chunks) break;
        float acc[16];
        tmem_ld16(taddr + ch * 16, acc);
        tmem_ld_wait();
        if (c == 7 || ch + 2 >= n_chunks) {  // last TMEM read of this warp: hand the accumulator back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&t_empty[buf]);
        }
        if (!row_valid) continue;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int n = n0 + ch * 16 + h * 8;
          if (n >= a.n_total) break;
          int t = j, co = n;
          if (a.up > 0) {
            const int r = n / a.C_out;
            co = n - r * a.C_out;
            t = a.up * j + r - a.pad;
          }
          if (t < 0 || t >= a.T_out) continue;
          const size_t off = ((size_t)b * a.T_out + t) * a.pitch_out + co;
          const float4 b0 = __ldg(reinterpret_cast<const float4*>(a.bias + n));
          const float4 b1 = __ldg(reinterpret_cast<const float4*>(a.bias + n + 4));
          const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
          float v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = __half2float(__float2half_rn(acc[h * 8 + i] + bb[i]));
          if (has_resid) {
            const uint4 ru = rbuf[2 * c + h];
            const __half2* rh = reinterpret_cast<const __half2*>(&ru);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float2 rf = __half22float2(rh[i]);
              v[2 * i] = __half2float(__float2half_rn(v[2 * i] + rf.x));
              v[2 * i + 1] = __half2float(__float2half_rn(v[2 * i + 1] + rf.y));
            }
          }
          if (a.out_raw) {
            uint4 o;
            __half2* oh = reinterpret_cast<__half2*>(&o);
#pragma unroll
            for (int i = 0; i < 4; ++i) oh[i] = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
            *reinterpret_cast<uint4*>(a.out_raw + off) = o;
          }
          if (a.out_act) {
            const float4 a0 = __ldg(reinterpret_cast<const float4*>(a.alpha + co));
            const float4 a1 = __ldg(reinterpret_cast<const float4*>(a.alpha + co + 4));
            const float4 i0 = __ldg(reinterpret_cast<const float4*>(a.inv_alpha + co));
            const float4 i1 = __ldg(reinterpret_cast<const float4*>(a.inv_alpha + co + 4));
            const float al[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float ia[8] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w};
            uint4 o;
            __half2* oh = reinterpret_cast<__half2*>(&o);
#pragma unroll
            for (int i = 0; i < 4; ++i)
              oh[i] = __floats2half2_rn(snake_fast(v[2 * i], al[2 * i], ia[2 * i]),
                                        snake_fast(v[2 * i + 1], al[2 * i + 1], ia[2 * i + 1]));
            *reinterpret_cast<uint4*>(a.out_act + off) = o;
          }
        }
      }
      if (n_chunks <= par) {  // (cannot happen: n_cols >= 32) keep the barrier protocol whole anyway
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

}  // namespace zb
