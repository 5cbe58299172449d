// Prefix conditioner on the device: PrefixConditioner.forward (zonos/conditioning.py:373-389) with
// the conditioners of the Zonos checkpoints -- phoneme / integer embedding rows (:295-312, :340-349), Fourier features
// (:318-337), a linearly projected pass-through vector (:352-358 + the base class's `project`, :26-27) and learned
// `uncond_vector`s (:38-40, :46-48) -- concatenated along the sequence and LayerNorm'ed, as ONE launch.
//
// grid (prefix position, utterance), 256 threads; thread t owns 8 consecutive columns (+ 2048 per round).  Every value
// is rounded to bf16 where the eager reference rounds (each conditioner's output, then the norm's output); the norm
// statistics are fp32 over the rounded row, like nn.LayerNorm on a bf16 tensor.
#include "../../include/zonos_b200.h"
#include "common.cuh"

namespace zb {

constexpr int PC_THREADS = 256;
constexpr int PC_MAX_SEGS = ZB200_MAX_COND_SEGMENTS;
struct PcSegs {
  zb200_cond_segment s[PC_MAX_SEGS];
  int n;
};

__device__ __forceinline__ float pc_block_sum(float v, float* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int k = 0; k < PC_THREADS / 32; ++k) t += red[k];
  return t;
}

__global__ void __launch_bounds__(PC_THREADS) prefix_cond_kernel(const __grid_constant__ PcSegs segs, int d, const bf16* __restrict__ ln_w,
                                                                 const bf16* __restrict__ ln_b, float eps, bf16* __restrict__ out,
                                                                 int l_total) {
  __shared__ float red[PC_THREADS / 32];
  __shared__ float s_x[64];  // Fourier: bf16(2 pi x_hat) per input; linear: the input vector is read from global
  const int pos = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  int j = 0;
  while (j + 1 < segs.n && pos >= segs.s[j + 1].start) ++j;
  const zb200_cond_segment& sg = segs.s[j];
  const int p = pos - sg.start;  // position inside the segment
  const bf16* table = reinterpret_cast<const bf16*>(sg.table);
  if (sg.kind == ZB200_COND_FOURIER) {
    const float* x = reinterpret_cast<const float*>(sg.x) + (size_t)b * sg.batch_stride + (size_t)p * sg.in_dim;
    if (tid < sg.in_dim) {
      const float xh = bf16_round((x[tid] - sg.min_val) / (sg.max_val - sg.min_val));
      s_x[tid] = bf16_round(6.283185307179586f * xh);
    }
  }
  __syncthreads();

  constexpr int MAXV = 2;  // d <= 4096
  float v[MAXV][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c0 = (i * PC_THREADS + tid) * 8;
    if (c0 >= d) continue;
    if (sg.kind == ZB200_COND_ROWS || sg.kind == ZB200_COND_VECTOR) {
      size_t row = 0;
      if (sg.kind == ZB200_COND_ROWS) {
        const int32_t* ids = reinterpret_cast<const int32_t*>(sg.x) + (size_t)b * sg.batch_stride;
        row = (size_t)(ids[p] + sg.id_offset);
      }
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(table + row * d + c0));
      const float2 a0 = unpack_bf16x2(u.x), a1 = unpack_bf16x2(u.y), a2 = unpack_bf16x2(u.z), a3 = unpack_bf16x2(u.w);
      v[i][0] = a0.x, v[i][1] = a0.y, v[i][2] = a1.x, v[i][3] = a1.y, v[i][4] = a2.x, v[i][5] = a2.y, v[i][6] = a3.x, v[i][7] = a3.y;
    } else if (sg.kind == ZB200_COND_FOURIER) {
      const int half = d / 2;
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int c = c0 + e, f = c < half ? c : c - half;
        float acc = 0.f;
        for (int k = 0; k < sg.in_dim; ++k) acc = fmaf(s_x[k], __bfloat162float(table[(size_t)f * sg.in_dim + k]), acc);
        acc = bf16_round(acc);                                    // the matmul's bf16 output
        v[i][e] = bf16_round(c < half ? cosf(acc) : sinf(acc));  // torch.cat([f.cos(), f.sin()]) on a bf16 tensor
      }
    } else {  // ZB200_COND_LINEAR: bf16(x W^T + bias), fp32 accumulate
      const bf16* x = reinterpret_cast<const bf16*>(sg.x) + (size_t)b * sg.batch_stride + (size_t)p * sg.in_dim;
      const bf16* bias = reinterpret_cast<const bf16*>(sg.bias);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int c = c0 + e;
        const bf16* w = table + (size_t)c * sg.in_dim;
        float acc = 0.f;
        for (int k = 0; k < sg.in_dim; k += 2) {
          const float2 ww = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(w + k));
          const float2 xx = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(x + k));
          acc = fmaf(xx.x, ww.x, acc);
          acc = fmaf(xx.y, ww.y, acc);
        }
        v[i][e] = bf16_round(acc + (bias ? __bfloat162float(bias[c]) : 0.f));
      }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) sum += v[i][e];
  }
  const float mean = pc_block_sum(sum, red) / (float)d;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i)
    if ((i * PC_THREADS + tid) * 8 < d)
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float t = v[i][e] - mean;
        sq += t * t;
      }
  const float rstd = 1.0f / sqrtf(pc_block_sum(sq, red) / (float)d + eps);
  bf16* orow = out + ((size_t)b * l_total + pos) * d;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c0 = (i * PC_THREADS + tid) * 8;
    if (c0 >= d) continue;
    const uint4 wu = __ldg(reinterpret_cast<const uint4*>(ln_w + c0)), bu = __ldg(reinterpret_cast<const uint4*>(ln_b + c0));
    const uint32_t wv[4] = {wu.x, wu.y, wu.z, wu.w}, bv[4] = {bu.x, bu.y, bu.z, bu.w};
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 ww = unpack_bf16x2(wv[e]), bb = unpack_bf16x2(bv[e]);
      o[e] = pack_bf16x2((v[i][2 * e] - mean) * rstd * ww.x + bb.x, (v[i][2 * e + 1] - mean) * rstd * ww.y + bb.y);
    }
    *reinterpret_cast<uint4*>(orow + c0) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

}  // namespace zb

using namespace zb;

extern "C" int zb200_prefix_conditioner(const zb200_cond_segment* segs, int n_segs, int batch, int l_total, int d, const void* ln_w,
                                        const void* ln_b, float eps, void* out, void* stream) {
  ZB_REQUIRE(segs && n_segs > 0 && n_segs <= PC_MAX_SEGS, "prefix_conditioner: 1..%d segments", PC_MAX_SEGS);
  ZB_REQUIRE(batch > 0 && l_total > 0 && d > 0 && d % 256 == 0 && d <= 4096 && ln_w && ln_b && out,
             "prefix_conditioner: bad shape (d a multiple of 256, <= 4096)");
  PcSegs p{};
  p.n = n_segs;
  int next = 0;
  for (int j = 0; j < n_segs; ++j) {
    const zb200_cond_segment& s = segs[j];
    ZB_REQUIRE(s.start == next && s.len > 0, "prefix_conditioner: segment %d must start at %d (segments tile the prefix in order)", j, next);
    ZB_REQUIRE(s.kind >= ZB200_COND_ROWS && s.kind <= ZB200_COND_LINEAR && s.table, "prefix_conditioner: segment %d: bad kind / null table", j);
    ZB_REQUIRE(s.kind == ZB200_COND_VECTOR || s.x, "prefix_conditioner: segment %d needs its input", j);
    ZB_REQUIRE(s.kind != ZB200_COND_VECTOR || s.len == 1, "prefix_conditioner: an uncond vector covers one position");
    ZB_REQUIRE(s.kind != ZB200_COND_FOURIER || (s.in_dim >= 1 && s.in_dim <= 64 && s.max_val != s.min_val),
               "prefix_conditioner: Fourier input_dim 1..64, max_val != min_val");
    ZB_REQUIRE(s.kind != ZB200_COND_LINEAR || (s.in_dim >= 2 && s.in_dim % 2 == 0), "prefix_conditioner: linear cond_dim must be even");
    next += s.len;
    p.s[j] = s;
  }
  ZB_REQUIRE(next == l_total, "prefix_conditioner: segments cover %d positions, l_total = %d", next, l_total);
  prefix_cond_kernel<<<dim3(l_total, batch), PC_THREADS, 0, (cudaStream_t)stream>>>(p, d, (const bf16*)ln_w, (const bf16*)ln_b, eps,
                                                                                   (bf16*)out, l_total);
  ZB_CHECK_LAUNCH("prefix_cond_kernel");
  return ZB200_OK;
}
