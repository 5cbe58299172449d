// Small kernels of the decode step (embedding sum, LayerNorm) and single-kernel test hooks of the C ABI.
#include "../../include/zonos_b200.h"
#include "common.cuh"
#include "epilogues.cuh"
#include "epilogues_t.cuh"
#include "gemm_skinny.cuh"
#include "gemm_swap.cuh"
#include "gemm_big.cuh"
#include "gemm_tc.cuh"
#include "ops.cuh"

unsigned long long* g_trace_buf = nullptr;  // debug timeline (zb200_debug_trace); also stamped by the engine's GEMMs
int g_trace_next = 0;

namespace zb {

// ------------------------------------------------------------------------------------------------ LayerNorm
// One warp per row, row held in registers (d <= 4096), fp32 two-pass statistics, affine, bf16 out.
template <int MAX_VEC>
__global__ void layernorm_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w, const bf16* __restrict__ b,
                                 bf16* __restrict__ out, int rows, int d, float eps, int in_row_mul, int in_row_add) {
  pdl_launch_dependents();
  pdl_wait_prior();
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
  const dim3 grid(blocks), block(warps_per_block * 32);
  if (d <= 2048)
    return launch_pdl(layernorm_kernel<8>, grid, block, 0, st, pdl_enabled(), x, w, b, out, rows, d, eps, in_row_mul, in_row_add);
  return launch_pdl(layernorm_kernel<16>, grid, block, 0, st, pdl_enabled(), x, w, b, out, rows, d, eps, in_row_mul, in_row_add);
}

// ------------------------------------------------------------------------------------------------ residual + LayerNorm
// Consumer of a deferred split-K GEMM (EpiPartial): one CTA per row,
//   o = bf16(sum_s ws[s][row][:])  (splits added in order, fp32)      -- the nn.Linear output, rounded like the reference
//   x = bf16(x + o)                                                  -- residual add (_torch.py:100-101)
//   xn = LayerNorm(x) * w + b                                        -- the NEXT norm (norm2, next layer's norm, or norm_f)
// which replaces the separate residual epilogue + LayerNorm kernel pair by one pass over the row.
constexpr int RLN_THREADS = 256;
constexpr int RLN_BATCH = 12;  // splits summed per round of loads (the planners use <= 12 splits: one round trip)
__global__ void __launch_bounds__(RLN_THREADS) resid_ln_kernel(bf16* __restrict__ x, const float* __restrict__ ws, int splits,
                                                               int rows, const bf16* __restrict__ w, const bf16* __restrict__ b,
                                                               bf16* __restrict__ xn, int d, float eps,
                                                               unsigned long long* trace, int trace_id) {
  pdl_launch_dependents();
#ifdef SK_TRACE
  if (threadIdx.x == 0) sk_stamp(trace, trace_id, 0);
#endif
  __shared__ float red[RLN_THREADS / 32], red2[RLN_THREADS / 32];
  const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int MAXV = 2;  // d <= 4096
  // the LayerNorm parameters do not depend on the previous kernel: fetch them before the dependency resolves
  uint4 wu[MAXV], bu[MAXV];
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int col = (i * RLN_THREADS + tid) * 8;
    if (col < d) wu[i] = __ldg(reinterpret_cast<const uint4*>(w + col)), bu[i] = __ldg(reinterpret_cast<const uint4*>(b + col));
  }
  pdl_wait_prior();
#ifdef SK_TRACE
  if (threadIdx.x == 0) sk_stamp(trace, trace_id, 4);
#endif
  float v[MAXV][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int col = (i * RLN_THREADS + tid) * 8;
    if (col < d) {
      float o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int s0 = 0; s0 < splits; s0 += RLN_BATCH) {  // up to 24 independent loads in flight, added in split order
        float4 a[RLN_BATCH], c[RLN_BATCH];
#pragma unroll
        for (int s = 0; s < RLN_BATCH; ++s)
          if (s0 + s < splits) {
            const float4* p = reinterpret_cast<const float4*>(ws + ((size_t)(s0 + s) * rows + row) * d + col);
            ldg256_cg(p, a[s], c[s]);  // 32 contiguous bytes per lane: one 256-bit load
          }
#pragma unroll
        for (int s = 0; s < RLN_BATCH; ++s)
          if (s0 + s < splits) {
            o[0] += a[s].x, o[1] += a[s].y, o[2] += a[s].z, o[3] += a[s].w;
            o[4] += c[s].x, o[5] += c[s].y, o[6] += c[s].z, o[7] += c[s].w;
          }
      }
      const uint4 xu = *reinterpret_cast<const uint4*>(x + (size_t)row * d + col);
      const float2 x0 = unpack_bf16x2(xu.x), x1 = unpack_bf16x2(xu.y), x2 = unpack_bf16x2(xu.z), x3 = unpack_bf16x2(xu.w);
      const float xin[8] = {x0.x, x0.y, x1.x, x1.y, x2.x, x2.y, x3.x, x3.y};
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[i][j] = bf16_round(xin[j] + bf16_round(o[j]));
        sum += v[i][j];
      }
      *reinterpret_cast<uint4*>(x + (size_t)row * d + col) =
          make_uint4(pack_bf16x2(v[i][0], v[i][1]), pack_bf16x2(v[i][2], v[i][3]), pack_bf16x2(v[i][4], v[i][5]),
                     pack_bf16x2(v[i][6], v[i][7]));
    }
  }
#ifdef SK_TRACE
  if (threadIdx.x == 0) sk_stamp(trace, trace_id, 5);
#endif
  // one barrier per reduction: each of the two uses its own slots, and every thread adds the warp sums in the same order
  auto block_sum = [&](float t, float* slots) {
    t = warp_sum(t);
    if (lane == 0) slots[warp] = t;
    __syncthreads();
    float a = 0.f;
#pragma unroll
    for (int k = 0; k < RLN_THREADS / 32; ++k) a += slots[k];
    return a;
  };
  const float mean = block_sum(sum, red) / (float)d;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i)
    if ((i * RLN_THREADS + tid) * 8 < d)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float t = v[i][j] - mean;
        sq += t * t;
      }
  const float rstd = 1.0f / sqrtf(block_sum(sq, red2) / (float)d + eps);
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int col = (i * RLN_THREADS + tid) * 8;
    if (col < d) {
      const uint32_t wv[4] = {wu[i].x, wu[i].y, wu[i].z, wu[i].w}, bv[4] = {bu[i].x, bu[i].y, bu[i].z, bu[i].w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 ww = unpack_bf16x2(wv[j]), bb = unpack_bf16x2(bv[j]);
        o[j] = pack_bf16x2((v[i][2 * j] - mean) * rstd * ww.x + bb.x, (v[i][2 * j + 1] - mean) * rstd * ww.y + bb.y);
      }
      *reinterpret_cast<uint4*>(xn + (size_t)row * d + col) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
#ifdef SK_TRACE
  __syncthreads();
  if (threadIdx.x == 0) sk_stamp(trace, trace_id, 9);
#endif
}

cudaError_t launch_resid_ln(bf16* x, const float* ws, int splits, int rows, const bf16* w, const bf16* b, bf16* xn, int d,
                            float eps, cudaStream_t st, unsigned long long* trace, int trace_id) {
  return launch_pdl(resid_ln_kernel, dim3(rows), dim3(RLN_THREADS), 0, st, pdl_enabled(), x, ws, splits, rows, w, b, xn, d, eps,
                    trace, trace_id);
}

// ------------------------------------------------------------------------------------------------ in_proj finish
// Consumer of the deferred split-K in_proj GEMM at M > 32 (EpiPartial): sums the splits in order, rounds to bf16 (the
// nn.Linear output), applies the interleaved-pair RoPE to q and k (_torch.py:18-30) and writes q and the K/V cache row
// (_torch.py:33-49) -- the work of EpiQkvRope, spread over rows x columns/8 threads instead of the last CTA of a tile.
__global__ void qkv_finish_kernel(const float* __restrict__ ws, int splits, int rows, int n_cols, bf16* __restrict__ q_out,
                                  KvLayout kv, const float* __restrict__ rope, const int* pos_ptr, int n_heads) {
  pdl_launch_dependents();
  pdl_wait_prior();
  const int row = blockIdx.x, col0 = threadIdx.x * 8;
  if (col0 >= n_cols) return;
  float o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int s0 = 0; s0 < splits; s0 += 8) {
    float4 a[8], c[8];
#pragma unroll
    for (int s = 0; s < 8; ++s)
      if (s0 + s < splits) {
        const float4* p = reinterpret_cast<const float4*>(ws + ((size_t)(s0 + s) * rows + row) * n_cols + col0);
        ldg256_cg(p, a[s], c[s]);
      }
#pragma unroll
    for (int s = 0; s < 8; ++s)
      if (s0 + s < splits) {
        o[0] += a[s].x, o[1] += a[s].y, o[2] += a[s].z, o[3] += a[s].w;
        o[4] += c[s].x, o[5] += c[s].y, o[6] += c[s].z, o[7] += c[s].w;
      }
  }
  const int pos = *pos_ptr;
  const int head_all = col0 >> 7, within = col0 & 127;
  const int kv_heads = kv.n_kv_heads;
  float r[8];
  if (head_all < n_heads + kv_heads) {
    const float4* cs = reinterpret_cast<const float4*>(rope + ((size_t)pos * 64 + (within >> 1)) * 2);
    const float4 t01 = __ldg(cs), t23 = __ldg(cs + 1);
    const float co[4] = {t01.x, t01.z, t23.x, t23.z}, si[4] = {t01.y, t01.w, t23.y, t23.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float a = bf16_round(o[2 * i]), b = bf16_round(o[2 * i + 1]);
      r[2 * i] = __fsub_rn(__fmul_rn(a, co[i]), __fmul_rn(b, si[i]));
      r[2 * i + 1] = __fadd_rn(__fmul_rn(b, co[i]), __fmul_rn(a, si[i]));
    }
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = o[i];
  }
  const uint4 v = make_uint4(pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]), pack_bf16x2(r[4], r[5]), pack_bf16x2(r[6], r[7]));
  if (head_all < n_heads) {
    *reinterpret_cast<uint4*>(q_out + (size_t)row * n_heads * 128 + head_all * 128 + within) = v;
  } else {
    const int is_v = head_all >= n_heads + kv_heads;
    const int head = head_all - n_heads - (is_v ? kv_heads : 0);
    reinterpret_cast<uint4*>(kv.row(row, pos, is_v, head))[(within >> 3) ^ (pos & 7)] = v;
  }
}

cudaError_t launch_qkv_finish(const float* ws, int splits, int rows, int n_cols, bf16* q_out, const KvLayout& kv,
                              const float* rope, const int* pos_ptr, int n_heads, cudaStream_t st) {
  return launch_pdl(qkv_finish_kernel, dim3(rows), dim3((n_cols / 8 + 31) / 32 * 32), 0, st, pdl_enabled(), ws, splits, rows,
                    n_cols, q_out, kv, rope, pos_ptr, n_heads);
}

// ------------------------------------------------------------------------------------------------ embedding sum
// grid (S, B): hidden[b*out_stride_b + out_off + s] = sum_q emb[q][ids[b,q,col0+s]], adds rounded to bf16 in q order;
// the row is written twice when dup_off > 0 (cond and uncond copies, model.py:141 / :196-199).
__global__ void embed_sum_kernel(const int32_t* __restrict__ ids, int ids_stride_b, int ids_stride_q, const int* col_ptr,
                                 int col_add, const bf16* __restrict__ emb, bf16* __restrict__ out, int n_q, int vocab,
                                 int d, int out_stride_b, int out_off, int dup_off) {
  pdl_launch_dependents();
  pdl_wait_prior();
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
  return launch_pdl(embed_sum_kernel, grid, dim3(256), 0, st, pdl_enabled(), ids, ids_stride_b, ids_stride_q, col_ptr,
                    col_add, emb, out, n_q, vocab, d, out_stride_b, out_off, dup_off);
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

extern "C" int zb200_embed_sum_at(const int32_t* ids, int ids_stride_b, int ids_stride_q, const int* col_ptr, int col_add,
                                  const void* emb, void* out, int batch, int n_q, int vocab, int d, int dup_rows_offset,
                                  void* stream) {
  ZB_REQUIRE(ids && col_ptr && emb && out && n_q > 0 && n_q <= ZB200_MAX_CODEBOOKS && d % 8 == 0 && batch > 0,
             "embed_sum_at: bad arguments");
  ZB_CHECK_CUDA(launch_embed_sum(ids, ids_stride_b, ids_stride_q, col_ptr, col_add, 1, (const bf16*)emb, (bf16*)out, batch, n_q,
                                 vocab, d, 1, 0, dup_rows_offset, (cudaStream_t)stream));
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

extern "C" int zb200_gemm_big_bf16(const void* a, int m, int lda, const void* w_packed, int n, int k, void* out, int ldo,
                                   void* stream) {
  ZB_REQUIRE(m > 0 && n > 0 && k > 0 && k % 64 == 0 && lda % 8 == 0, "gemm_big: bad shape m=%d n=%d k=%d lda=%d", m, n, k, lda);
  ZB_REQUIRE((reinterpret_cast<uintptr_t>(a) & 15) == 0, "gemm_big: A must be 16-byte aligned");
  int dev = 0, sms = 0;
  ZB_CHECK_CUDA(cudaGetDevice(&dev));
  ZB_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  EpiStoreBf16<64> epi{(bf16*)out, ldo, n};
  ZB_CHECK_CUDA((launch_gemm_big<1>(a, (size_t)lda * 2, m, k, w_packed, ceil_div(n, 64), epi, sms, (cudaStream_t)stream)));
  return ZB200_OK;
}

/* debug: device buffer [64][160][12] uint64 receiving the in-kernel globaltimer timeline of the next
 * zb200_gemm_skinny_bf16 launches (libraries built with ZB200_TRACE=1); returns launches traced so far and resets. */
extern "C" int zb200_debug_trace(unsigned long long* device_buf) {
  const int n = g_trace_next;
  g_trace_buf = device_buf;
  g_trace_next = 0;
#ifdef SK_TRACE
  return n;
#else
  (void)n;
  set_error("debug_trace: library built without ZB200_TRACE=1");
  return ZB200_ERR_STATE;
#endif
}

extern "C" int zb200_gemm_skinny_bf16(const void* a, int m, int lda, const void* w_packed, int n, int k, void* out,
                                      int ldo, int grid, float* ws, int64_t ws_floats, int* tickets, int n_tickets,
                                      void* stream) {
  ZB_REQUIRE(m > 0 && m <= GEMM_BM && n > 0 && k > 0 && k % 64 == 0 && lda % 8 == 0,
             "gemm_skinny: bad shape m=%d n=%d k=%d lda=%d (m <= 128)", m, n, k, lda);
  const int n_tiles = ceil_div(n, SK_NT);
  bool prefer_k1 = false;
  if (grid <= 0) {  // plan the grid the way the engine does for its ticketed GEMMs (aligned to the tile groups)
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const SkPlan p = plan_ticketed(sms, m, n_tiles, k / 64);
    grid = p.grid, prefer_k1 = p.prefer_k1;
  }
  ZB_REQUIRE(grid >= 1 && grid <= 1024, "gemm_skinny: grid=%d out of range", grid);
  ZB_REQUIRE(ws && tickets && ws_floats >= (int64_t)grid * 2 * skinny_tg(skinny_mode(m, n_tiles, k / 64, prefer_k1)) * SK_PARTIAL_F4 * 4 &&
                 n_tickets >= n_tiles,
             "gemm_skinny: workspace too small (needs grid*2*tg*32 KB and n_tiles tickets)");
  SkinnyCore g{};
  g.Wp = w_packed, g.n_tiles = n_tiles, g.k_blocks = k / 64, g.m_valid = m, g.partial_ws = ws, g.tickets = tickets;
  g.n_splits = 1;
  g.trace = g_trace_buf, g.trace_id = g_trace_buf ? g_trace_next++ : -1;
  ARowMajor al{(const uint8_t*)a, (size_t)lda * 2};
  EpiStoreBf16<SK_NT> epi{(bf16*)out, ldo, n};
  ZB_CHECK_CUDA(launch_gemm_skinny(g, al, epi, grid, (cudaStream_t)stream, prefer_k1));
  return ZB200_OK;
}

/* apply_heads + the classifier-free-guidance mix of _compute_logits (zonos/model.py:100-116) as one weight-streaming GEMM:
 * xn bf16 [2*batch][d] (rows [0,B) cond, [B,2B) uncond) x packed heads (row_mode 2) -> logits fp32 [batch][n_q][vocab],
 * uncond + (cond - uncond) * cfg_scale, pad slots (>= 1025) = -inf.  The engine runs the same kernel inside its step. */
extern "C" int zb200_heads_cfg(const void* xn, int batch, int d, const void* w_heads_packed, int n_q, int vocab,
                               int head_rows_padded, float cfg_scale, float* logits, int grid, float* ws, int64_t ws_floats,
                               int* tickets, int n_tickets, void* stream) {
  ZB_REQUIRE(xn && w_heads_packed && logits && batch > 0 && 2 * batch <= GEMM_BM && d > 0 && d % 64 == 0,
             "heads_cfg: bad shape batch=%d d=%d (2*batch <= 128)", batch, d);
  ZB_REQUIRE(n_q > 0 && vocab > 1025 && vocab <= head_rows_padded && head_rows_padded % SK_NT == 0, "heads_cfg: bad head layout");
  const int tiles_per_head = head_rows_padded / SK_NT, n_tiles = n_q * tiles_per_head;
  ZB_REQUIRE(grid >= 1 && grid <= 1024, "heads_cfg: grid=%d out of range", grid);
  ZB_REQUIRE(ws && tickets && ws_floats >= (int64_t)grid * 2 * skinny_tg(skinny_mode(2 * batch, n_tiles, d / 64)) * SK_PARTIAL_F4 * 4 &&
                 n_tickets >= n_tiles,
             "heads_cfg: workspace too small (needs grid*2*tg*32 KB and n_tiles tickets)");
  SkinnyCore g{};
  g.Wp = w_heads_packed, g.n_tiles = n_tiles, g.k_blocks = d / 64, g.m_valid = 2 * batch, g.partial_ws = ws, g.tickets = tickets;
  g.n_splits = 1;
  g.trace = nullptr, g.trace_id = -1;
  ARowCfgPairs al{(const uint8_t*)xn, (size_t)d * 2, batch};
  EpiHeadsCfg<SK_NT> epi{logits, batch, n_q, vocab, tiles_per_head, cfg_scale};
  ZB_CHECK_CUDA(launch_gemm_skinny(g, al, epi, grid, (cudaStream_t)stream));
  return ZB200_OK;
}

/* x <- bf16(x + bf16(a * w^T)); xn <- LayerNorm(x) * ln_w + ln_b: the out_proj / fc2 + residual + next-norm group of the
 * decode step as the engine runs it (deferred split-K GEMM -> resid_ln_kernel).  ws: >= splits*m*n floats. */
extern "C" int zb200_gemm_resid_ln(const void* a, int m, int lda, const void* w_packed, int n, int k, int splits, float* ws,
                                   int64_t ws_floats, void* x, const void* ln_w, const void* ln_b, void* xn, float eps,
                                   void* stream) {
  ZB_REQUIRE(m > 0 && m <= GEMM_BM && n > 0 && n % 256 == 0 && n <= 4096 && k > 0 && k % 64 == 0 && lda % 8 == 0,
             "gemm_resid_ln: bad shape m=%d n=%d k=%d lda=%d", m, n, k, lda);
  const int n_tiles = n / SK_NT, mode = skinny_mode(m, n_tiles, k / 64);
  const int n_groups = n_tiles / skinny_tg(mode), kbu = (k / 64) / skinny_kg(mode);
  ZB_REQUIRE(splits >= 1 && splits <= kbu, "gemm_resid_ln: splits=%d exceeds the %d units of a group", splits, kbu);
  ZB_REQUIRE(ws && ws_floats >= (int64_t)splits * m * n, "gemm_resid_ln: workspace too small");
  SkinnyCore g{};
  g.Wp = w_packed, g.n_tiles = n_tiles, g.k_blocks = k / 64, g.m_valid = m, g.partial_ws = nullptr, g.tickets = nullptr;
  g.n_splits = splits, g.trace = g_trace_buf, g.trace_id = g_trace_buf ? g_trace_next++ : -1;
  ARowMajor al{(const uint8_t*)a, (size_t)lda * 2};
  EpiPartial epi{ws, m, n};
  ZB_CHECK_CUDA(launch_gemm_skinny(g, al, epi, n_groups * splits, (cudaStream_t)stream));
  const int rln_id = g_trace_buf ? g_trace_next++ : -1;
  ZB_CHECK_CUDA(launch_resid_ln((bf16*)x, ws, splits, m, (const bf16*)ln_w, (const bf16*)ln_b, (bf16*)xn, n, eps,
                                (cudaStream_t)stream, g_trace_buf, rln_id));
  return ZB200_OK;
}

/* Hybrid blocks (mamba_ssm Block with fused_add_norm): hidden = a * w^T (bf16), residual <- hidden + residual (fp32),
 * out = norm(residual): the mixer's out_proj (or the MLP's fc2) together with the NEXT block's add + norm, as a deferred
 * split-K weight-streaming GEMM whose partial sums are reduced by the add + norm kernel.  ws: fp32 partials, the split count
 * is planned from its capacity (>= m*n floats). */
extern "C" int zb200_gemm_add_norm(const void* a, int m, int lda, const void* w_packed, int n, int k, float* ws, int64_t ws_floats,
                                   float* residual, int has_residual, const void* norm_w, const void* norm_b, void* out, float eps,
                                   int rms, void* stream) {
  ZB_REQUIRE(a && w_packed && ws && residual && norm_w && out, "gemm_add_norm: null argument");
  ZB_REQUIRE(m > 0 && m <= GEMM_BM && n > 0 && n % 128 == 0 && n <= 8192 && k > 0 && k % 64 == 0 && lda % 8 == 0,
             "gemm_add_norm: bad shape m=%d n=%d k=%d lda=%d (m <= 128, n %% 128 == 0)", m, n, k, lda);
  ZB_REQUIRE(ws_floats >= (int64_t)m * n, "gemm_add_norm: workspace too small (needs >= m*n floats)");
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int n_tiles = n / SK_NT;
  const SkPlan p = plan_deferred(sms, m, n_tiles, k / 64, ws_floats, n);
  SkinnyCore g{};
  g.Wp = w_packed, g.n_tiles = n_tiles, g.k_blocks = k / 64, g.m_valid = m, g.partial_ws = nullptr, g.tickets = nullptr;
  g.n_splits = p.splits, g.trace = g_trace_buf, g.trace_id = g_trace_buf ? g_trace_next++ : -1;
  ARowMajor al{(const uint8_t*)a, (size_t)lda * 2};
  EpiPartial epi{ws, m, n};
  ZB_CHECK_CUDA(launch_gemm_skinny(g, al, epi, p.grid, (cudaStream_t)stream));
  ZB_CHECK_CUDA(launch_add_norm(nullptr, residual, has_residual, (const bf16*)norm_w, (const bf16*)norm_b, (bf16*)out, m, n, eps, rms,
                                ws, p.splits, (cudaStream_t)stream));
  return ZB200_OK;
}

extern "C" int zb200_gemm_swap_bf16(const void* a, int m, int lda, const void* w_packed, int n, int k, void* out, int ldo,
                                    int grid, float* ws, int64_t ws_floats, int* tickets, int n_tickets, void* stream) {
  ZB_REQUIRE(m > 0 && m <= 8 && n > 0 && k > 0 && k % 128 == 0 && lda % 8 == 0,
             "gemm_swap: bad shape m=%d n=%d k=%d lda=%d (m <= 8, k %% 128 == 0)", m, n, k, lda);
  const int n_tiles = ceil_div(n, SK_NT);
  ZB_REQUIRE(grid >= 1 && grid <= 1024, "gemm_swap: grid=%d out of range", grid);
  ZB_REQUIRE(ws && tickets && ws_floats >= (int64_t)grid * 2048 && n_tickets >= (n_tiles + 1) / 2,
             "gemm_swap: workspace too small (needs grid*8 KB and ceil(n/128) tickets)");
  SwapCore g{};
  g.Wp = w_packed, g.n_tiles = n_tiles, g.k_blocks = k / 64, g.m_valid = m, g.partial_ws = ws, g.tickets = tickets;
  g.trace = g_trace_buf, g.trace_id = g_trace_buf ? g_trace_next++ : -1;
  ARowMajor al{(const uint8_t*)a, (size_t)lda * 2};
  EpiTStoreBf16 epi{(bf16*)out, ldo, n};
  ZB_CHECK_CUDA(launch_gemm_swap(g, al, epi, grid, (cudaStream_t)stream));
  return ZB200_OK;
}
