// Mamba2 mixer, single decode step (the hybrid backbone, zonos/backbone/_mamba_ssm.py:18-31 builds its
// layers with mamba_ssm's create_block; the arithmetic restated here is mamba_ssm 2.2.4 `Mamba2.step`:
// causal_conv1d_update -> selective_state_update -> RMSNormGated).  The projections around these kernels are the
// weight-streaming tcgen05 GEMMs of the transformer path.
//
// All three kernels are HBM-bound; the state update dominates: the SSM state of one (sequence, head) is P x N bf16
// (64 x 128 = 16 KB), read once and written once per token and layer -- 2 MB per sequence and layer at the hybrid's shape.
#include "../../include/zonos_b200.h"
#include "common.cuh"

#include <cstdlib>

namespace zb {

__device__ __forceinline__ float silu_f(float v) { return v / (1.f + __expf(-v)); }

// ---------------------------------------------------------------------------------------------- causal_conv1d_update
// state [rows][C][4] bf16 (oldest first): shift by one, append x[c] = zxbcdt[row][col0 + c], out = silu(sum_k s[k] w[c][k] + b)
__global__ void __launch_bounds__(256) mamba2_conv_update_kernel(bf16* __restrict__ state, const bf16* __restrict__ zxbcdt,
                                                                 int ld, int col0, const bf16* __restrict__ w,
                                                                 const bf16* __restrict__ bias, bf16* __restrict__ out,
                                                                 int channels) {
  pdl_launch_dependents();
  const int c = blockIdx.x * blockDim.x + threadIdx.x, row = blockIdx.y;
  if (c >= channels) return;
  const uint2 wu = __ldg(reinterpret_cast<const uint2*>(w + (size_t)c * 4));
  const float b = bias ? __bfloat162float(bias[c]) : 0.f;
  pdl_wait_prior();
  uint2* sp = reinterpret_cast<uint2*>(state + ((size_t)row * channels + c) * 4);
  const uint2 su = *sp;
  const bf16 xin = zxbcdt[(size_t)row * ld + col0 + c];
  const float2 s01 = unpack_bf16x2(su.x), s23 = unpack_bf16x2(su.y);
  const float2 w01 = unpack_bf16x2(wu.x), w23 = unpack_bf16x2(wu.y);
  const float x = __bfloat162float(xin);
  // window after the shift: s1, s2, s3, x
  const float acc = s01.y * w01.x + s23.x * w01.y + s23.y * w23.x + x * w23.y + b;
  *sp = make_uint2(pack_bf16x2(s01.y, s23.x), pack_bf16x2(s23.y, x));
  out[(size_t)row * channels + c] = __float2bfloat16_rn(silu_f(acc));
}

// ---------------------------------------------------------------------------------------------- selective_state_update
// grid (head, row), 256 threads.  h[p][n] = h[p][n] * exp(dt*A) + dt * x[p] * B[n];  y[p] = sum_n h[p][n] * C[n] + D * x[p]
// with dt = softplus(dt_raw + dt_bias), A = -exp(A_log).  Lane layout: 16 lanes x 8 states cover one p row (256 B), a
// warp covers two rows per 16-byte access; each lane keeps its 8 B / C values in registers.
template <int N>
__global__ void __launch_bounds__(256) mamba2_ssd_step_kernel(bf16* __restrict__ ssm_state, const bf16* __restrict__ xbc,
                                                              int ld_xbc, const bf16* __restrict__ zxbcdt, int ld, int dt_col0,
                                                              const float* __restrict__ A_log, const float* __restrict__ Dp,
                                                              const float* __restrict__ dt_bias, bf16* __restrict__ y,
                                                              int n_heads, int head_dim, int n_groups) {
  static_assert(N == 128, "one p row = 16 lanes x 8 states");
  pdl_launch_dependents();
  const int h = blockIdx.x, row = blockIdx.y, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int psub = lane >> 4, n0 = (lane & 15) * 8;
  const int d_inner = n_heads * head_dim, grp = h / (n_heads / n_groups);
  const float A = -__expf(A_log[h]), Dh = Dp[h], dtb = dt_bias[h];
  bf16* st = ssm_state + ((size_t)row * n_heads + h) * (size_t)head_dim * N;
  // the state does not depend on the previous kernel of the step: fetch this thread's rows before the dependency resolves
  constexpr int MAXI = 8;  // head_dim <= 128
  uint4 hs[MAXI];
#pragma unroll
  for (int i = 0; i < MAXI; ++i) {
    const int p = 2 * (warp + 8 * i) + psub;
    if (p < head_dim) hs[i] = *reinterpret_cast<const uint4*>(st + (size_t)p * N + n0);
  }
  pdl_wait_prior();
  const bf16* xr = xbc + (size_t)row * ld_xbc;
  const uint4 bu = *reinterpret_cast<const uint4*>(xr + d_inner + grp * N + n0);
  const uint4 cu = *reinterpret_cast<const uint4*>(xr + d_inner + n_groups * N + grp * N + n0);
  float Bv[8], Cv[8];
  {
    const uint32_t bw[4] = {bu.x, bu.y, bu.z, bu.w}, cw[4] = {cu.x, cu.y, cu.z, cu.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 b2 = unpack_bf16x2(bw[j]), c2 = unpack_bf16x2(cw[j]);
      Bv[2 * j] = b2.x, Bv[2 * j + 1] = b2.y, Cv[2 * j] = c2.x, Cv[2 * j + 1] = c2.y;
    }
  }
  const float dtr = __bfloat162float(zxbcdt[(size_t)row * ld + dt_col0 + h]) + dtb;
  const float dt = dtr <= 20.f ? log1pf(expf(dtr)) : dtr;
  const float dA = expf(dt * A);
#pragma unroll
  for (int i = 0; i < MAXI; ++i) {
    const int p = 2 * (warp + 8 * i) + psub;
    if (p >= head_dim) break;  // (warp-uniform up to psub: head_dim is even)
    const float x = __bfloat162float(xr[h * head_dim + p]);
    const float dtx = dt * x;
    const uint32_t hw[4] = {hs[i].x, hs[i].y, hs[i].z, hs[i].w};
    uint32_t ow[4];
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 h2 = unpack_bf16x2(hw[j]);
      const float a = fmaf(h2.x, dA, dtx * Bv[2 * j]), b = fmaf(h2.y, dA, dtx * Bv[2 * j + 1]);
      acc = fmaf(a, Cv[2 * j], acc);
      acc = fmaf(b, Cv[2 * j + 1], acc);
      ow[j] = pack_bf16x2(a, b);
    }
    *reinterpret_cast<uint4*>(st + (size_t)p * N + n0) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((lane & 15) == 0) y[(size_t)row * d_inner + h * head_dim + p] = __float2bfloat16_rn(fmaf(Dh, x, acc));
  }
}

// Persistent form of the step kernel for head_dim 64 (the Mamba2 default): the grid is a multiple of the SM count, each CTA
// walks (row, head) tiles with a grid stride and fetches the NEXT tile's 16 KB of state while it updates the current one,
// so the HBM stream never drains between tiles (the one-tile-per-CTA form measured 0.66 of the HBM peak at 64 rows).
template <int N>
__global__ void __launch_bounds__(256) mamba2_ssd_step_p64_kernel(bf16* __restrict__ ssm_state, const bf16* __restrict__ xbc,
                                                                  int ld_xbc, const bf16* __restrict__ zxbcdt, int ld,
                                                                  int dt_col0, const float* __restrict__ A_log,
                                                                  const float* __restrict__ Dp, const float* __restrict__ dt_bias,
                                                                  bf16* __restrict__ y, int n_heads, int n_groups, int n_tiles) {
  static_assert(N == 128, "one p row = 16 lanes x 8 states");
  constexpr int P = 64, MAXI = 4;
  pdl_launch_dependents();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int psub = lane >> 4, n0 = (lane & 15) * 8;
  const int d_inner = n_heads * P;
  int tile = blockIdx.x;
  if (tile >= n_tiles) return;
  uint4 hs[MAXI], hn[MAXI];
  {
    const bf16* st = ssm_state + (size_t)tile * P * N;
#pragma unroll
    for (int i = 0; i < MAXI; ++i) hs[i] = *reinterpret_cast<const uint4*>(st + (size_t)(2 * (warp + 8 * i) + psub) * N + n0);
  }
  pdl_wait_prior();
  for (; tile < n_tiles; tile += gridDim.x) {
    const int row = tile / n_heads, h = tile - row * n_heads, grp = h / (n_heads / n_groups);
    const int nxt = tile + gridDim.x;
    if (nxt < n_tiles) {
      const bf16* sn = ssm_state + (size_t)nxt * P * N;
#pragma unroll
      for (int i = 0; i < MAXI; ++i) hn[i] = *reinterpret_cast<const uint4*>(sn + (size_t)(2 * (warp + 8 * i) + psub) * N + n0);
    }
    bf16* st = ssm_state + (size_t)tile * P * N;
    const bf16* xr = xbc + (size_t)row * ld_xbc;
    const uint4 bu = *reinterpret_cast<const uint4*>(xr + d_inner + grp * N + n0);
    const uint4 cu = *reinterpret_cast<const uint4*>(xr + d_inner + n_groups * N + grp * N + n0);
    float xv[MAXI];
#pragma unroll
    for (int i = 0; i < MAXI; ++i) xv[i] = __bfloat162float(xr[h * P + 2 * (warp + 8 * i) + psub]);
    const float A = -__expf(__ldg(A_log + h)), Dh = __ldg(Dp + h);
    const float dtr = __bfloat162float(zxbcdt[(size_t)row * ld + dt_col0 + h]) + __ldg(dt_bias + h);
    const float dt = dtr <= 20.f ? log1pf(expf(dtr)) : dtr;
    const float dA = expf(dt * A);
    float Bv[8], Cv[8];
    {
      const uint32_t bw[4] = {bu.x, bu.y, bu.z, bu.w}, cw[4] = {cu.x, cu.y, cu.z, cu.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 b2 = unpack_bf16x2(bw[j]), c2 = unpack_bf16x2(cw[j]);
        Bv[2 * j] = b2.x, Bv[2 * j + 1] = b2.y, Cv[2 * j] = c2.x, Cv[2 * j + 1] = c2.y;
      }
    }
#pragma unroll
    for (int i = 0; i < MAXI; ++i) {
      const int p = 2 * (warp + 8 * i) + psub;
      const float dtx = dt * xv[i];
      const uint32_t hw[4] = {hs[i].x, hs[i].y, hs[i].z, hs[i].w};
      uint32_t ow[4];
      float acc = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 h2 = unpack_bf16x2(hw[j]);
        const float a = fmaf(h2.x, dA, dtx * Bv[2 * j]), b = fmaf(h2.y, dA, dtx * Bv[2 * j + 1]);
        acc = fmaf(a, Cv[2 * j], acc);
        acc = fmaf(b, Cv[2 * j + 1], acc);
        ow[j] = pack_bf16x2(a, b);
      }
      *reinterpret_cast<uint4*>(st + (size_t)p * N + n0) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if ((lane & 15) == 0) y[(size_t)row * d_inner + h * P + p] = __float2bfloat16_rn(fmaf(Dh, xv[i], acc));
    }
#pragma unroll
    for (int i = 0; i < MAXI; ++i) hs[i] = hn[i];
  }
}

// ---------------------------------------------------------------------------------------------- prefill: conv over a sequence
// Rows are sequence-major (row = seq * seq_len + t, the engine's prefill layout).  Thread = (sequence, channel): the 4-tap
// window lives in registers over the whole sequence; the final window is the conv state the decode steps continue from.
__global__ void __launch_bounds__(256) mamba2_conv_seq_kernel(bf16* __restrict__ state, const bf16* __restrict__ zxbcdt, int ld,
                                                              int col0, const bf16* __restrict__ w, const bf16* __restrict__ bias,
                                                              bf16* __restrict__ out, int seq_len, int channels, int use_init) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x, seq = blockIdx.y;
  if (c >= channels) return;
  const uint2 wu = __ldg(reinterpret_cast<const uint2*>(w + (size_t)c * 4));
  const float2 w01 = unpack_bf16x2(wu.x), w23 = unpack_bf16x2(wu.y);
  const float b = bias ? __bfloat162float(bias[c]) : 0.f;
  uint2* sp = reinterpret_cast<uint2*>(state + ((size_t)seq * channels + c) * 4);
  float s1 = 0.f, s2 = 0.f, s3 = 0.f;  // the three most recent inputs (oldest first)
  if (use_init) {
    const uint2 su = *sp;
    s1 = unpack_bf16x2(su.x).y;
    const float2 s23 = unpack_bf16x2(su.y);
    s2 = s23.x, s3 = s23.y;
  }
  float s0 = 0.f;
  const bf16* xin = zxbcdt + (size_t)seq * seq_len * ld + col0 + c;
  bf16* o = out + (size_t)seq * seq_len * channels + c;
  for (int t = 0; t < seq_len; ++t) {
    const float x = __bfloat162float(xin[(size_t)t * ld]);
    const float acc = s1 * w01.x + s2 * w01.y + s3 * w23.x + x * w23.y + b;
    o[(size_t)t * channels] = __float2bfloat16_rn(silu_f(acc));
    s0 = s1, s1 = s2, s2 = s3, s3 = x;
  }
  *sp = make_uint2(pack_bf16x2(s0, s1), pack_bf16x2(s2, s3));
}

// ---------------------------------------------------------------------------------------------- prefill: SSD scan
// Same lane layout as the step kernel; the state of (sequence, head) stays in fp32 registers over the whole sequence
// (4 rows x 8 states per thread) and is written once at the end, so the scan moves only x / B / C / dt in and y out.
// Tokens are staged 16 at a time in shared memory with cp.async (double-buffered: chunk i+1 arrives while chunk i is
// scanned), dt / exp(dt*A) are computed once per token by 16 threads, y leaves as coalesced 128-byte rows per chunk.
// (Sequential in t: the prefix is <= a few hundred tokens; a chunked SSD form on the tensor cores is what a long-prompt
// prefill would need.)
constexpr int SCAN_TC = 16;
template <int N>
__global__ void __launch_bounds__(256) mamba2_ssd_scan_kernel(bf16* __restrict__ ssm_state, const bf16* __restrict__ xbc,
                                                              int ld_xbc, const bf16* __restrict__ zxbcdt, int ld, int dt_col0,
                                                              const float* __restrict__ A_log, const float* __restrict__ Dp,
                                                              const float* __restrict__ dt_bias, bf16* __restrict__ y,
                                                              int seq_len, int n_heads, int head_dim, int n_groups, int use_init) {
  static_assert(N == 128, "one p row = 16 lanes x 8 states");
  constexpr int MAXI = 4, PMAX = 64;  // head_dim <= 64 in this kernel (32 fp32 state registers per thread)
  __shared__ __align__(16) bf16 sB[2][SCAN_TC][N], sC[2][SCAN_TC][N], sX[2][SCAN_TC][PMAX], sY[SCAN_TC][PMAX];
  __shared__ float sDt[2][SCAN_TC], sDa[2][SCAN_TC];
  const int h = blockIdx.x, seq = blockIdx.y, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int psub = lane >> 4, n0 = (lane & 15) * 8;
  const int d_inner = n_heads * head_dim, grp = h / (n_heads / n_groups);
  const float A = -__expf(A_log[h]), Dh = Dp[h], dtb = dt_bias[h];
  bf16* st = ssm_state + ((size_t)seq * n_heads + h) * (size_t)head_dim * N;
  float hs[MAXI][8];
#pragma unroll
  for (int i = 0; i < MAXI; ++i) {
    const int p = 2 * (warp + 8 * i) + psub;
#pragma unroll
    for (int j = 0; j < 8; ++j) hs[i][j] = 0.f;
    if (use_init && p < head_dim) {
      const uint4 u = *reinterpret_cast<const uint4*>(st + (size_t)p * N + n0);
      const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(uw[j]);
        hs[i][2 * j] = f.x, hs[i][2 * j + 1] = f.y;
      }
    }
  }
  const size_t row0 = (size_t)seq * seq_len;
  const int boff = d_inner + grp * N, coff = d_inner + n_groups * N + grp * N, xoff = h * head_dim;
  const int xchunks = head_dim / 8;  // 16-byte pieces of a token's x
  // stage chunk `ck` into buffer `buf`: per token 16 pieces of B, 16 of C, head_dim/8 of x
  auto stage = [&](int ck, int buf) {
    const int t0 = ck * SCAN_TC, per_tok = 32 + xchunks;
    for (int i = tid; i < SCAN_TC * per_tok; i += 256) {
      const int tt = i / per_tok, piece = i - tt * per_tok;
      if (t0 + tt >= seq_len) continue;
      const bf16* xr = xbc + (row0 + t0 + tt) * ld_xbc;
      if (piece < 16) cp_async16(&sB[buf][tt][piece * 8], xr + boff + piece * 8);
      else if (piece < 32) cp_async16(&sC[buf][tt][(piece - 16) * 8], xr + coff + (piece - 16) * 8);
      else cp_async16(&sX[buf][tt][(piece - 32) * 8], xr + xoff + (piece - 32) * 8);
    }
    cp_async_commit();
  };
  auto load_dt = [&](int ck) -> float {  // threads 0..15: the raw dt of their token in chunk ck
    const int t = ck * SCAN_TC + tid;
    return (tid < SCAN_TC && t < seq_len) ? __bfloat162float(zxbcdt[(row0 + t) * ld + dt_col0 + h]) : 0.f;
  };
  const int n_chunks = (seq_len + SCAN_TC - 1) / SCAN_TC;
  stage(0, 0);
  float dt_raw = load_dt(0);
  for (int ck = 0; ck < n_chunks; ++ck) {
    const int buf = ck & 1;
    if (tid < SCAN_TC) {
      const float dts = dt_raw + dtb;
      const float dt = dts <= 20.f ? log1pf(expf(dts)) : dts;
      sDt[buf][tid] = dt, sDa[buf][tid] = expf(dt * A);
    }
    if (ck + 1 < n_chunks) {
      stage(ck + 1, buf ^ 1);
      dt_raw = load_dt(ck + 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();  // chunk ck is in shared memory; the previous chunk's y rows have been written out
    const int nt = min(SCAN_TC, seq_len - ck * SCAN_TC);
    for (int tt = 0; tt < nt; ++tt) {
      const uint4 bu = *reinterpret_cast<const uint4*>(&sB[buf][tt][n0]);
      const uint4 cu = *reinterpret_cast<const uint4*>(&sC[buf][tt][n0]);
      float Bv[8], Cv[8];
      {
        const uint32_t bw[4] = {bu.x, bu.y, bu.z, bu.w}, cw[4] = {cu.x, cu.y, cu.z, cu.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 b2 = unpack_bf16x2(bw[j]), c2 = unpack_bf16x2(cw[j]);
          Bv[2 * j] = b2.x, Bv[2 * j + 1] = b2.y, Cv[2 * j] = c2.x, Cv[2 * j + 1] = c2.y;
        }
      }
      const float dt = sDt[buf][tt], dA = sDa[buf][tt];
#pragma unroll
      for (int i = 0; i < MAXI; ++i) {
        const int p = 2 * (warp + 8 * i) + psub;
        if (p >= head_dim) break;
        const float x = __bfloat162float(sX[buf][tt][p]);
        const float dtx = dt * x;
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          hs[i][j] = fmaf(hs[i][j], dA, dtx * Bv[j]);
          acc = fmaf(hs[i][j], Cv[j], acc);
        }
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if ((lane & 15) == 0) sY[tt][p] = __float2bfloat16_rn(fmaf(Dh, x, acc));
      }
    }
    __syncthreads();
    for (int i = tid; i < nt * xchunks; i += 256) {  // y rows of the chunk: head_dim*2 contiguous bytes per token
      const int tt = i / xchunks, piece = i - tt * xchunks;
      *reinterpret_cast<uint4*>(y + (row0 + ck * SCAN_TC + tt) * d_inner + xoff + piece * 8) =
          *reinterpret_cast<const uint4*>(&sY[tt][piece * 8]);
    }
  }
#pragma unroll
  for (int i = 0; i < MAXI; ++i) {
    const int p = 2 * (warp + 8 * i) + psub;
    if (p < head_dim)
      *reinterpret_cast<uint4*>(st + (size_t)p * N + n0) =
          make_uint4(pack_bf16x2(hs[i][0], hs[i][1]), pack_bf16x2(hs[i][2], hs[i][3]), pack_bf16x2(hs[i][4], hs[i][5]),
                     pack_bf16x2(hs[i][6], hs[i][7]));
  }
}

// ---------------------------------------------------------------------------------------------- RMSNormGated
// out = rmsnorm_g(y * silu(z)) * w over each of n_groups channel groups (norm_before_gate = False); z = zxbcdt[:, :d_inner]
__global__ void __launch_bounds__(256) mamba2_gated_rmsnorm_kernel(const bf16* __restrict__ y, const bf16* __restrict__ zxbcdt,
                                                                   int ld, const bf16* __restrict__ w, bf16* __restrict__ out,
                                                                   int d_inner, int n_groups, float eps) {
  pdl_launch_dependents();
  __shared__ float red[8];
  __shared__ float bcast;
  const int row = blockIdx.x / n_groups, g = blockIdx.x % n_groups, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gs = d_inner / n_groups;
  constexpr int MAXV = 4;  // group size <= 8192
  uint4 wu[MAXV];
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = (i * 256 + tid) * 8;
    if (c < gs) wu[i] = __ldg(reinterpret_cast<const uint4*>(w + g * gs + c));
  }
  pdl_wait_prior();
  float v[MAXV][8];
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = (i * 256 + tid) * 8;
    if (c < gs) {
      const uint4 yu = *reinterpret_cast<const uint4*>(y + (size_t)row * d_inner + g * gs + c);
      const uint4 zu = *reinterpret_cast<const uint4*>(zxbcdt + (size_t)row * ld + g * gs + c);
      const uint32_t yw[4] = {yu.x, yu.y, yu.z, yu.w}, zw[4] = {zu.x, zu.y, zu.z, zu.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 y2 = unpack_bf16x2(yw[j]), z2 = unpack_bf16x2(zw[j]);
        v[i][2 * j] = y2.x * silu_f(z2.x), v[i][2 * j + 1] = y2.y * silu_f(z2.y);
        sq += v[i][2 * j] * v[i][2 * j] + v[i][2 * j + 1] * v[i][2 * j + 1];
      }
    }
  }
  sq = warp_sum(sq);
  if (lane == 0) red[warp] = sq;
  __syncthreads();
  if (tid == 0) {
    float a = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) a += red[k];
    bcast = 1.0f / sqrtf(a / (float)gs + eps);
  }
  __syncthreads();
  const float rstd = bcast;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = (i * 256 + tid) * 8;
    if (c < gs) {
      const uint32_t ww[4] = {wu[i].x, wu[i].y, wu[i].z, wu[i].w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 w2 = unpack_bf16x2(ww[j]);
        o[j] = pack_bf16x2(v[i][2 * j] * rstd * w2.x, v[i][2 * j + 1] * rstd * w2.y);
      }
      *reinterpret_cast<uint4*>(out + (size_t)row * d_inner + g * gs + c) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}


// ---------------------------------------------------------------------------------------------- fused add + norm (prenorm)
// mamba_ssm Block with fused_add_norm (layer_norm_fn, prenorm=True, residual_in_fp32=True; _mamba_ssm.py:18-31,46-57):
//   residual <- x + residual (fp32, written back; residual = x for the first layer),  out = norm(residual) * w (+ b)  in bf16
// rms != 0: RMSNorm (no mean subtraction, bias optional), else LayerNorm.  One CTA per row, fp32 statistics.
// With `ws` (deferred split-K GEMM, EpiPartial layout [splits][rows][d] fp32) the kernel is also the GEMM's reduction:
//   x = bf16(sum_s ws[s][row][:])   (splits added in order: the nn.Linear output, rounded like the eager reference)
__global__ void __launch_bounds__(256) add_norm_kernel(const bf16* __restrict__ x, float* __restrict__ residual, int has_residual,
                                                       const bf16* __restrict__ w, const bf16* __restrict__ b,
                                                       bf16* __restrict__ out, int d, float eps, int rms,
                                                       const float* __restrict__ ws, int splits, int rows) {
  pdl_launch_dependents();
  __shared__ float red[8];
  __shared__ float bcast;
  const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int MAXV = 4;  // d <= 8192
  uint4 wu[MAXV], bu[MAXV];
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = (i * 256 + tid) * 8;
    if (c < d) {
      wu[i] = __ldg(reinterpret_cast<const uint4*>(w + c));
      bu[i] = b ? __ldg(reinterpret_cast<const uint4*>(b + c)) : make_uint4(0, 0, 0, 0);
    }
  }
  pdl_wait_prior();
  float v[MAXV][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = (i * 256 + tid) * 8;
    if (c < d) {
      uint4 xu;
      if (ws != nullptr) {
        float o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        constexpr int BATCH = 12;  // independent loads in flight (the planner uses <= 12 splits here: one round trip)
        for (int s0 = 0; s0 < splits; s0 += BATCH) {
          float4 a0[BATCH], a1[BATCH];
#pragma unroll
          for (int sp = 0; sp < BATCH; ++sp)
            if (s0 + sp < splits) ldg256_cg(reinterpret_cast<const float4*>(ws + ((size_t)(s0 + sp) * rows + row) * d + c), a0[sp], a1[sp]);
#pragma unroll
          for (int sp = 0; sp < BATCH; ++sp)
            if (s0 + sp < splits) {
              o[0] += a0[sp].x, o[1] += a0[sp].y, o[2] += a0[sp].z, o[3] += a0[sp].w;
              o[4] += a1[sp].x, o[5] += a1[sp].y, o[6] += a1[sp].z, o[7] += a1[sp].w;
            }
        }
        xu = make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
      } else {
        xu = *reinterpret_cast<const uint4*>(x + (size_t)row * d + c);
      }
      const uint32_t xw[4] = {xu.x, xu.y, xu.z, xu.w};
      float4* rp = reinterpret_cast<float4*>(residual + (size_t)row * d + c);
      float4 r0 = make_float4(0.f, 0.f, 0.f, 0.f), r1 = r0;
      if (has_residual) r0 = rp[0], r1 = rp[1];
      const float rr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 x2 = unpack_bf16x2(xw[j]);
        v[i][2 * j] = x2.x + rr[2 * j], v[i][2 * j + 1] = x2.y + rr[2 * j + 1];
        sum += v[i][2 * j] + v[i][2 * j + 1];
      }
      rp[0] = make_float4(v[i][0], v[i][1], v[i][2], v[i][3]);
      rp[1] = make_float4(v[i][4], v[i][5], v[i][6], v[i][7]);
    }
  }
  auto block_sum = [&](float t) {
    t = warp_sum(t);
    __syncthreads();
    if (lane == 0) red[warp] = t;
    __syncthreads();
    if (tid == 0) {
      float a = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) a += red[k];
      bcast = a;
    }
    __syncthreads();
    return bcast;
  };
  const float mean = rms ? 0.f : block_sum(sum) / (float)d;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i)
    if ((i * 256 + tid) * 8 < d)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float t = v[i][j] - mean;
        sq += t * t;
      }
  const float rstd = 1.0f / sqrtf(block_sum(sq) / (float)d + eps);
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = (i * 256 + tid) * 8;
    if (c < d) {
      const uint32_t ww[4] = {wu[i].x, wu[i].y, wu[i].z, wu[i].w}, bw[4] = {bu[i].x, bu[i].y, bu[i].z, bu[i].w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 w2 = unpack_bf16x2(ww[j]), b2 = unpack_bf16x2(bw[j]);
        o[j] = pack_bf16x2((v[i][2 * j] - mean) * rstd * w2.x + b2.x, (v[i][2 * j + 1] - mean) * rstd * w2.y + b2.y);
      }
      *reinterpret_cast<uint4*>(out + (size_t)row * d + c) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}


// ---------------------------------------------------------------------------------------------- hybrid attention / MLP glue
// The hybrid's attention layers are mamba_ssm MHA modules: Wqkv output [q (H*128) | k (Hkv*128) | v (Hkv*128)], rotary on
// q and k in the NON-interleaved form (pairs (i, i + 64) of a head), cos / sin cached in the activation dtype (bf16).
// This kernel splits a row of the projection into q, k, v and rotates q and k at position pos0 + (row % seq_len); the
// paged-KV write and the attention itself are the transformer path's kernels.  (In a C-side engine this is one more
// epilogue of the in_proj GEMM, or -- with the q/k weight rows permuted at pack time -- the existing interleaved one.)
__global__ void __launch_bounds__(256) rope_split_kernel(const bf16* __restrict__ qkv, const float* __restrict__ inv_freq,
                                                         bf16* __restrict__ q, bf16* __restrict__ k, bf16* __restrict__ v,
                                                         int seq_len, int pos0, const int* __restrict__ pos_ptr, int n_heads,
                                                         int n_kv_heads) {
  pdl_launch_dependents();
  const int row = blockIdx.x, tid = threadIdx.x;
  const int ld = (n_heads + 2 * n_kv_heads) * 128;
  pdl_wait_prior();
  const float pos = (float)((pos_ptr ? *pos_ptr : pos0) + row % seq_len);  // pos_ptr: device-side position (graph replay)
  const bf16* src = qkv + (size_t)row * ld;
  const int n_rot = (n_heads + n_kv_heads) * 64;  // rotated pairs of the row
  for (int i = tid; i < n_rot; i += 256) {
    const int head = i >> 6, j = i & 63;
    const float ang = pos * inv_freq[j];
    const float c = bf16_round(cosf(ang)), sn = bf16_round(sinf(ang));
    const float x1 = __bfloat162float(src[head * 128 + j]), x2 = __bfloat162float(src[head * 128 + j + 64]);
    const bf16 o1 = __float2bfloat16_rn(x1 * c - x2 * sn), o2 = __float2bfloat16_rn(x1 * sn + x2 * c);
    bf16* dst = head < n_heads ? q + (size_t)row * n_heads * 128 + head * 128
                               : k + (size_t)row * n_kv_heads * 128 + (head - n_heads) * 128;
    dst[j] = o1, dst[j + 64] = o2;
  }
  const bf16* vs = src + (n_heads + n_kv_heads) * 128;
  for (int i = tid; i < n_kv_heads * 128; i += 256) v[(size_t)row * n_kv_heads * 128 + i] = vs[i];
}

// mamba_ssm GatedMLP between fc1 and fc2: y, gate = fc1(x).chunk(2, -1); out = y * silu(gate)
__global__ void __launch_bounds__(256) gated_mlp_act_kernel(const bf16* __restrict__ a, bf16* __restrict__ out, int d_ff,
                                                            long long total) {
  pdl_launch_dependents();
  pdl_wait_prior();
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const long long r = i / d_ff;
    const int c = (int)(i - r * d_ff);
    const float y = __bfloat162float(a[r * 2 * d_ff + c]), g = __bfloat162float(a[r * 2 * d_ff + d_ff + c]);
    out[i] = __float2bfloat16_rn(y * (g / (1.f + expf(-g))));
  }
}

cudaError_t launch_add_norm(const bf16* x, float* residual, int has_residual, const bf16* w, const bf16* b, bf16* out, int rows,
                            int d, float eps, int rms, const float* ws, int splits, cudaStream_t st) {
  return launch_pdl(add_norm_kernel, dim3((unsigned)rows), dim3(256), 0, st, pdl_enabled(), x, residual, has_residual, w, b, out, d,
                    eps, rms, ws, splits, rows);
}

}  // namespace zb

using namespace zb;

extern "C" int zb200_mamba2_conv_update(void* conv_state, const void* zxbcdt, int ld, int col0, const void* w, const void* bias,
                                        void* out, int rows, int channels, int d_conv, void* stream) {
  ZB_REQUIRE(conv_state && zxbcdt && w && out && rows > 0 && channels > 0, "mamba2_conv_update: bad arguments");
  ZB_REQUIRE(d_conv == 4, "mamba2_conv_update: d_conv=%d (only the Mamba2 default 4 is built)", d_conv);
  ZB_REQUIRE(col0 >= 0 && col0 + channels <= ld, "mamba2_conv_update: columns [%d, %d) exceed the row pitch %d", col0, col0 + channels, ld);
  ZB_CHECK_CUDA(launch_pdl(mamba2_conv_update_kernel, dim3((unsigned)ceil_div(channels, 256), (unsigned)rows), dim3(256), 0,
                           (cudaStream_t)stream, pdl_enabled(), (bf16*)conv_state, (const bf16*)zxbcdt, ld, col0, (const bf16*)w,
                           (const bf16*)bias, (bf16*)out, channels));
  return ZB200_OK;
}

extern "C" int zb200_mamba2_ssd_step(void* ssm_state, const void* xbc, int ld_xbc, const void* zxbcdt, int ld, int dt_col0,
                                     const float* A_log, const float* D, const float* dt_bias, void* y, int rows, int n_heads,
                                     int head_dim, int d_state, int n_groups, void* stream) {
  ZB_REQUIRE(ssm_state && xbc && zxbcdt && A_log && D && dt_bias && y && rows > 0, "mamba2_ssd_step: bad arguments");
  ZB_REQUIRE(d_state == 128, "mamba2_ssd_step: d_state=%d (only the Mamba2 default 128 is built)", d_state);
  ZB_REQUIRE(head_dim > 0 && head_dim <= 128 && head_dim % 8 == 0 && n_heads > 0 && n_groups > 0 && n_heads % n_groups == 0,
             "mamba2_ssd_step: unsupported head layout H=%d P=%d G=%d", n_heads, head_dim, n_groups);
  ZB_REQUIRE(ld_xbc >= n_heads * head_dim + 2 * n_groups * d_state && ld_xbc % 8 == 0 && dt_col0 + n_heads <= ld,
             "mamba2_ssd_step: row pitches too small");
  static const bool one_tile = zb::tuning_env("ZB200_SSD_ONE_TILE") != nullptr;  // A/B switch
  if (head_dim == 64 && !one_tile) {
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int n_tiles = rows * n_heads;
    static const int per_sm = zb::tuning_env("ZB200_SSD_CTAS_PER_SM") ? atoi(zb::tuning_env("ZB200_SSD_CTAS_PER_SM")) : 6;
    const int grid = n_tiles < sms * per_sm ? n_tiles : sms * per_sm;
    ZB_CHECK_CUDA(launch_pdl(mamba2_ssd_step_p64_kernel<128>, dim3((unsigned)grid), dim3(256), 0, (cudaStream_t)stream, pdl_enabled(),
                             (bf16*)ssm_state, (const bf16*)xbc, ld_xbc, (const bf16*)zxbcdt, ld, dt_col0, A_log, D, dt_bias,
                             (bf16*)y, n_heads, n_groups, n_tiles));
    return ZB200_OK;
  }
  ZB_CHECK_CUDA(launch_pdl(mamba2_ssd_step_kernel<128>, dim3((unsigned)n_heads, (unsigned)rows), dim3(256), 0, (cudaStream_t)stream,
                           pdl_enabled(), (bf16*)ssm_state, (const bf16*)xbc, ld_xbc, (const bf16*)zxbcdt, ld, dt_col0, A_log, D,
                           dt_bias, (bf16*)y, n_heads, head_dim, n_groups));
  return ZB200_OK;
}

extern "C" int zb200_mamba2_gated_rmsnorm(const void* y, const void* zxbcdt, int ld, const void* w, void* out, int rows,
                                          int d_inner, int n_groups, float eps, void* stream) {
  ZB_REQUIRE(y && zxbcdt && w && out && rows > 0 && d_inner > 0 && n_groups > 0, "mamba2_gated_rmsnorm: bad arguments");
  ZB_REQUIRE(d_inner % n_groups == 0 && (d_inner / n_groups) % 8 == 0 && d_inner / n_groups <= 8192 && ld >= d_inner && ld % 8 == 0,
             "mamba2_gated_rmsnorm: unsupported shape d_inner=%d groups=%d ld=%d", d_inner, n_groups, ld);
  ZB_CHECK_CUDA(launch_pdl(mamba2_gated_rmsnorm_kernel, dim3((unsigned)(rows * n_groups)), dim3(256), 0, (cudaStream_t)stream,
                           pdl_enabled(), (const bf16*)y, (const bf16*)zxbcdt, ld, (const bf16*)w, (bf16*)out, d_inner, n_groups,
                           eps));
  return ZB200_OK;
}

extern "C" int zb200_mamba2_conv_seq(void* conv_state, const void* zxbcdt, int ld, int col0, const void* w, const void* bias,
                                     void* out, int n_seq, int seq_len, int channels, int d_conv, int use_initial_state,
                                     void* stream) {
  ZB_REQUIRE(conv_state && zxbcdt && w && out && n_seq > 0 && seq_len > 0 && channels > 0, "mamba2_conv_seq: bad arguments");
  ZB_REQUIRE(d_conv == 4, "mamba2_conv_seq: d_conv=%d (only the Mamba2 default 4 is built)", d_conv);
  ZB_REQUIRE(col0 >= 0 && col0 + channels <= ld, "mamba2_conv_seq: columns [%d, %d) exceed the row pitch %d", col0, col0 + channels, ld);
  mamba2_conv_seq_kernel<<<dim3((unsigned)ceil_div(channels, 256), (unsigned)n_seq), 256, 0, (cudaStream_t)stream>>>(
      (bf16*)conv_state, (const bf16*)zxbcdt, ld, col0, (const bf16*)w, (const bf16*)bias, (bf16*)out, seq_len, channels,
      use_initial_state);
  ZB_CHECK_LAUNCH("mamba2_conv_seq_kernel");
  return ZB200_OK;
}

extern "C" int zb200_mamba2_ssd_scan(void* ssm_state, const void* xbc, int ld_xbc, const void* zxbcdt, int ld, int dt_col0,
                                     const float* A_log, const float* D, const float* dt_bias, void* y, int n_seq, int seq_len,
                                     int n_heads, int head_dim, int d_state, int n_groups, int use_initial_state, void* stream) {
  ZB_REQUIRE(ssm_state && xbc && zxbcdt && A_log && D && dt_bias && y && n_seq > 0 && seq_len > 0, "mamba2_ssd_scan: bad arguments");
  ZB_REQUIRE(d_state == 128, "mamba2_ssd_scan: d_state=%d (only the Mamba2 default 128 is built)", d_state);
  ZB_REQUIRE(head_dim > 0 && head_dim <= 64 && head_dim % 8 == 0 && n_heads > 0 && n_groups > 0 && n_heads % n_groups == 0,
             "mamba2_ssd_scan: unsupported head layout H=%d P=%d G=%d (P <= 64)", n_heads, head_dim, n_groups);
  ZB_REQUIRE(ld_xbc >= n_heads * head_dim + 2 * n_groups * d_state && ld_xbc % 8 == 0 && dt_col0 + n_heads <= ld,
             "mamba2_ssd_scan: row pitches too small");
  mamba2_ssd_scan_kernel<128><<<dim3((unsigned)n_heads, (unsigned)n_seq), 256, 0, (cudaStream_t)stream>>>(
      (bf16*)ssm_state, (const bf16*)xbc, ld_xbc, (const bf16*)zxbcdt, ld, dt_col0, A_log, D, dt_bias, (bf16*)y, seq_len, n_heads,
      head_dim, n_groups, use_initial_state);
  ZB_CHECK_LAUNCH("mamba2_ssd_scan_kernel");
  return ZB200_OK;
}

extern "C" int zb200_add_norm(const void* x, float* residual, int has_residual, const void* w, const void* b, void* out, int rows,
                              int d, float eps, int rms, void* stream) {
  ZB_REQUIRE(x && residual && w && out && rows > 0, "add_norm: bad arguments");
  ZB_REQUIRE(d > 0 && d % 8 == 0 && d <= 8192, "add_norm: d=%d must be a multiple of 8 (<= 8192)", d);
  ZB_CHECK_CUDA(launch_add_norm((const bf16*)x, residual, has_residual, (const bf16*)w, (const bf16*)b, (bf16*)out, rows, d, eps, rms,
                                nullptr, 0, (cudaStream_t)stream));
  return ZB200_OK;
}

extern "C" int zb200_rope_split(const void* qkv, const float* inv_freq, void* q, void* k, void* v, int rows, int seq_len, int pos0,
                                const int* pos_ptr, int n_heads, int n_kv_heads, void* stream) {
  ZB_REQUIRE(qkv && inv_freq && q && k && v && rows > 0 && seq_len > 0 && rows % seq_len == 0 && pos0 >= 0 && n_heads > 0 &&
                 n_kv_heads > 0,
             "rope_split: bad arguments");
  ZB_CHECK_CUDA(launch_pdl(rope_split_kernel, dim3((unsigned)rows), dim3(256), 0, (cudaStream_t)stream, pdl_enabled(),
                           (const bf16*)qkv, inv_freq, (bf16*)q, (bf16*)k, (bf16*)v, seq_len, pos0, pos_ptr, n_heads, n_kv_heads));
  return ZB200_OK;
}

extern "C" int zb200_gated_mlp_act(const void* fc1_out, void* out, int rows, int d_ff, void* stream) {
  ZB_REQUIRE(fc1_out && out && rows > 0 && d_ff > 0, "gated_mlp_act: bad arguments");
  const long long total = (long long)rows * d_ff;
  const int grid = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  ZB_CHECK_CUDA(launch_pdl(gated_mlp_act_kernel, dim3((unsigned)grid), dim3(256), 0, (cudaStream_t)stream, pdl_enabled(),
                           (const bf16*)fc1_out, (bf16*)out, d_ff, total));
  return ZB200_OK;
}
