// Small kernels of the decode step (embedding sum, LayerNorm) and single-kernel test hooks of the C ABI.
#include "../../include/zonos_b200.h"
#include "common.cuh"
#include "epilogues.cuh"
#include "gemm_tc.cuh"
#include "ops.cuh"

namespace zb {

// ------------------------------------------------------------------------------------------------ LayerNorm
// One warp per row, row held in registers (d <= 4096), fp32 two-pass statistics, affine, bf16 out.
template <int MAX_VEC>
__global__ void layernorm_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w, const bf16* __restrict__ b,
                                 bf16* __restrict__ out, int rows, int d, float eps, int in_row_mul, int in_row_add) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const int nvec = d / 256;  // uint4 (8 bf16) per lane
  const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)(warp * in_row_mul + in_row_add) * d);
  float v[MAX_VEC][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < MAX_VEC; ++i) {
    if (i < nvec) {
      const uint4 u = xr[i * 32 + lane];
      const float2 a = unpack_bf16x2(u.x), bb = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), dd = unpack_bf16x2(u.w);
      v[i][0] = a.x, v[i][1] = a.y, v[i][2] = bb.x, v[i][3] = bb.y, v[i][4] = c.x, v[i][5] = c.y, v[i][6] = dd.x, v[i][7] = dd.y;
#pragma unroll
      for (int j = 0; j < 8; ++j) sum += v[i][j];
    }
  }
  const float mean = warp_sum(sum) / (float)d;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < MAX_VEC; ++i)
    if (i < nvec)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float t = v[i][j] - mean;
        sq += t * t;
      }
  const float rstd = 1.0f / sqrtf(warp_sum(sq) / (float)d + eps);
  uint4* orow = reinterpret_cast<uint4*>(out + (size_t)warp * d);
  const uint4* wr = reinterpret_cast<const uint4*>(w);
  const uint4* br = reinterpret_cast<const uint4*>(b);
#pragma unroll
  for (int i = 0; i < MAX_VEC; ++i) {
    if (i < nvec) {
      const uint4 wu = __ldg(wr + i * 32 + lane), bu = __ldg(br + i * 32 + lane);
      const uint32_t wv[4] = {wu.x, wu.y, wu.z, wu.w}, bv[4] = {bu.x, bu.y, bu.z, bu.w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 ww = unpack_bf16x2(wv[j]), bb = unpack_bf16x2(bv[j]);
        o[j] = pack_bf16x2((v[i][2 * j] - mean) * rstd * ww.x + bb.x, (v[i][2 * j + 1] - mean) * rstd * ww.y + bb.y);
      }
      orow[i * 32 + lane] = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

cudaError_t launch_layernorm(const bf16* x, const bf16* w, const bf16* b, bf16* out, int rows, int d, float eps,
                             int in_row_mul, int in_row_add, cudaStream_t st) {
  const int warps_per_block = 4;
  const int blocks = ceil_div(rows, warps_per_block);
  if (d <= 2048)
    layernorm_kernel<8><<<blocks, warps_per_block * 32, 0, st>>>(x, w, b, out, rows, d, eps, in_row_mul, in_row_add);
  else
    layernorm_kernel<16><<<blocks, warps_per_block * 32, 0, st>>>(x, w, b, out, rows, d, eps, in_row_mul, in_row_add);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ embedding sum
// grid (S, B): hidden[b*out_stride_b + out_off + s] = sum_q emb[q][ids[b,q,col0+s]], adds rounded to bf16 in q order;
// the row is written twice when dup_off > 0 (cond and uncond copies, model.py:141 / :196-199).
__global__ void embed_sum_kernel(const int32_t* __restrict__ ids, int ids_stride_b, int ids_stride_q, const int* col_ptr,
                                 int col_add, const bf16* __restrict__ emb, bf16* __restrict__ out, int n_q, int vocab,
                                 int d, int out_stride_b, int out_off, int dup_off) {
  const int s = blockIdx.x, b = blockIdx.y;
  const int col = (col_ptr ? *col_ptr : 0) + col_add + s;
  __shared__ int sid[ZB200_MAX_CODEBOOKS];
  if (threadIdx.x < n_q) {
    int id = ids[(size_t)b * ids_stride_b + (size_t)threadIdx.x * ids_stride_q + col];
    sid[threadIdx.x] = min(max(id, 0), vocab - 1);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < d / 8; i += blockDim.x) {
    float acc[8];
#pragma unroll
    for (int q = 0; q < ZB200_MAX_CODEBOOKS; ++q) {
      if (q < n_q) {
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(emb + ((size_t)q * vocab + sid[q]) * d) + i);
        const float2 a = unpack_bf16x2(u.x), bb = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), dd = unpack_bf16x2(u.w);
        const float e[8] = {a.x, a.y, bb.x, bb.y, c.x, c.y, dd.x, dd.y};
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = (q == 0) ? e[j] : bf16_round(acc[j] + e[j]);
      }
    }
    const uint4 o = make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]), pack_bf16x2(acc[4], acc[5]),
                               pack_bf16x2(acc[6], acc[7]));
    const size_t r0 = (size_t)b * out_stride_b + out_off + s;
    reinterpret_cast<uint4*>(out + r0 * d)[i] = o;
    if (dup_off > 0) reinterpret_cast<uint4*>(out + (r0 + dup_off) * d)[i] = o;
  }
}

cudaError_t launch_embed_sum(const int32_t* ids, int ids_stride_b, int ids_stride_q, const int* col_ptr, int col_add,
                             int n_cols, const bf16* emb, bf16* out, int batch, int n_q, int vocab, int d,
                             int out_stride_b, int out_off, int dup_off, cudaStream_t st) {
  dim3 grid(n_cols, batch);
  embed_sum_kernel<<<grid, 256, 0, st>>>(ids, ids_stride_b, ids_stride_q, col_ptr, col_add, emb, out, n_q, vocab, d,
                                         out_stride_b, out_off, dup_off);
  return cudaGetLastError();
}

}  // namespace zb

using namespace zb;

extern "C" int zb200_layernorm(const void* x, const void* w, const void* b, void* out, int rows, int d, float eps,
                               void* stream) {
  ZB_REQUIRE(rows >= 0 && d > 0 && d % 256 == 0 && d <= 4096, "layernorm: d=%d must be a multiple of 256, <= 4096", d);
  if (rows == 0) return ZB200_OK;
  ZB_CHECK_CUDA(launch_layernorm((const bf16*)x, (const bf16*)w, (const bf16*)b, (bf16*)out, rows, d, eps, 1, 0,
                                 (cudaStream_t)stream));
  return ZB200_OK;
}

extern "C" int zb200_embed_sum(const int32_t* ids, int ids_stride_b, int ids_stride_q, int n_cols, const void* emb,
                               void* out, int batch, int n_q, int vocab, int d, int out_stride_b, int out_off,
                               int dup_rows_offset, void* stream) {
  ZB_REQUIRE(n_q > 0 && n_q <= ZB200_MAX_CODEBOOKS && d % 8 == 0 && batch >= 0 && n_cols >= 0, "embed_sum: bad shape");
  if (batch == 0 || n_cols == 0) return ZB200_OK;
  ZB_CHECK_CUDA(launch_embed_sum(ids, ids_stride_b, ids_stride_q, nullptr, 0, n_cols, (const bf16*)emb, (bf16*)out, batch,
                                 n_q, vocab, d, out_stride_b, out_off, dup_rows_offset, (cudaStream_t)stream));
  return ZB200_OK;
}

extern "C" int zb200_gemm_bf16(const void* a, int m, int lda, const void* w_packed, int n, int k, void* out, int ldo,
                               int splits, float* ws, int* tickets, void* stream) {
  ZB_REQUIRE(m > 0 && n > 0 && k > 0 && k % 64 == 0 && lda % 8 == 0, "gemm: bad shape m=%d n=%d k=%d lda=%d", m, n, k, lda);
  ZB_REQUIRE(splits >= 1 && splits <= k / 64, "gemm: splits=%d out of range", splits);
  ZB_REQUIRE(splits == 1 || (ws && tickets), "gemm: split-K needs workspace and tickets");
  constexpr int NT = 64;
  GemmCore g{};
  g.Wp = w_packed;
  g.k_blocks = k / 64;
  g.kb_per_split = ceil_div(g.k_blocks, splits);
  g.splits = ceil_div(g.k_blocks, g.kb_per_split);
  g.m_valid = m;
  g.partial_ws = ws;
  g.tickets = tickets;
  g.stream_weights = m <= 128;
  ARowMajor al{(const uint8_t*)a, (size_t)lda * 2};
  EpiStoreBf16<NT> epi{(bf16*)out, ldo, n};
  ZB_CHECK_CUDA((launch_gemm_tc<NT, 6, 1>(g, al, epi, ceil_div(n, NT), ceil_div(m, GEMM_BM), (cudaStream_t)stream)));
  return ZB200_OK;
}
