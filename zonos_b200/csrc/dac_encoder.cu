// DAC 44.1 kHz ENCODER + residual vector quantiser: DACAutoencoder.encode (zonos/autoencoder.py:27-28)
// -> HF DacModel.encode (modeling_dac.py:581-605) -> DacEncoder (:442-472), DacEncoderBlock (:210-232), DacResidualUnit
// (:173-207), Snake1d (:85-99), DacResidualVectorQuantizer.forward (:300-343), DacVectorQuantize (:100-170).
//
// The reference runs this path in fp32 (no autocast around encode), once per request, on the audio prefix (the CLIs'
// default is 100 ms of silence = 9 frames), and its result is INTEGER codes picked by an argmax: what matters is fp32
// arithmetic that reproduces those picks, not tensor-core throughput -- fp16 / tf32 operands would flip nearest-neighbour
// decisions.  So: plain fp32 FMA kernels, channels-first tensors [B][C][T] like torch.
//   conv1d_f32_kernel  direct convolution, 64 output steps x 64 output channels per CTA, 4 x 4 outputs per thread, input
//                      window and weight slab of 4 input channels staged in shared memory; stride / dilation / zero padding;
//                      optional residual add (the residual unit's `hidden + output`)
//   snake_f32_kernel   x + (alpha + 1e-9)^-1 sin^2(alpha x)
//   rvq_encode_kernel  one CTA per time step, residual vector (1024) in shared memory; per codebook: in_proj (8 x 1024),
//                      F.normalize, the reference's distance expression, first-index argmax, out_proj of the picked code,
//                      residual update
#include "../../include/zonos_b200.h"
#include "common.cuh"

namespace zb {

constexpr int CE_TT = 64, CE_TC = 64, CE_CI = 4, CE_THREADS = 256, CE_MAXK = 16;
constexpr int CE_MAX_SPAN = (CE_TT - 1) * 8 + (CE_MAXK - 1) * 9 + 1;  // stride <= 8, dilation <= 9

__global__ void __launch_bounds__(CE_THREADS) conv1d_f32_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                                const float* __restrict__ bias, const float* __restrict__ resid,
                                                                float* __restrict__ out, int C_in, int C_out, int T_in, int T_out,
                                                                int K, int stride, int dil, int pad) {
  __shared__ float xs[CE_CI][CE_MAX_SPAN];
  __shared__ float ws[CE_CI][CE_MAXK][CE_TC];
  const int t0 = blockIdx.x * CE_TT, c0 = blockIdx.y * CE_TC, b = blockIdx.z;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int span = (CE_TT - 1) * stride + (K - 1) * dil + 1;
  const int in0 = t0 * stride - pad;  // input index of the window's first element
  const float* xb = x + (size_t)b * C_in * T_in;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int ci0 = 0; ci0 < C_in; ci0 += CE_CI) {
    for (int idx = tid; idx < CE_CI * span; idx += CE_THREADS) {
      const int ci = idx / span, p = idx - ci * span, g = in0 + p;
      xs[ci][p] = (ci0 + ci < C_in && g >= 0 && g < T_in) ? xb[(size_t)(ci0 + ci) * T_in + g] : 0.f;
    }
    for (int idx = tid; idx < CE_CI * K * CE_TC; idx += CE_THREADS) {  // w is [C_in][K][C_out]
      const int co = idx % CE_TC, r = idx / CE_TC, k = r % K, ci = r / K;
      ws[ci][k][co] = (ci0 + ci < C_in && c0 + co < C_out) ? w[((size_t)(ci0 + ci) * K + k) * C_out + c0 + co] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int ci = 0; ci < CE_CI; ++ci)
      for (int k = 0; k < K; ++k) {
        float xv[4], wv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) xv[i] = xs[ci][(tx + 16 * i) * stride + k * dil];
#pragma unroll
        for (int j = 0; j < 4; ++j) wv[j] = ws[ci][k][ty + 16 * j];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(xv[i], wv[j], acc[i][j]);
      }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int co = c0 + ty + 16 * j;
    if (co >= C_out) continue;
    const float bv = bias ? bias[co] : 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int t = t0 + tx + 16 * i;
      if (t >= T_out) continue;
      const size_t o = ((size_t)b * C_out + co) * T_out + t;
      float v = acc[i][j] + bv;
      if (resid) v = resid[o] + v;
      out[o] = v;
    }
  }
}

__global__ void snake_f32_kernel(const float* __restrict__ x, const float* __restrict__ alpha, float* __restrict__ out, int C, int T,
                                 long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)((i / T) % C);
    const float a = alpha[c], v = x[i], s = sinf(a * v);
    out[i] = v + (1.0f / (a + 1e-9f)) * (s * s);  // modeling_dac.py:94-99
  }
}

constexpr int RQ_THREADS = 256, RQ_MAXD = 16;
__global__ void __launch_bounds__(RQ_THREADS) rvq_encode_kernel(const float* __restrict__ latent, const float* __restrict__ in_w,
                                                                const float* __restrict__ in_b, const float* __restrict__ codebook,
                                                                const float* __restrict__ cb_norm, const float* __restrict__ out_w,
                                                                const float* __restrict__ out_b, long long* __restrict__ codes, int Q,
                                                                int H, int D, int N, int T) {
  extern __shared__ float sm[];
  float* res = sm;           // [H] residual of this time step
  float* e = sm + H;         // [D] projected latent, then normalised
  float* red = e + RQ_MAXD;  // [RQ_THREADS] reduction scratch
  int* redi = reinterpret_cast<int*>(red + RQ_THREADS);
  __shared__ int s_pick;
  const int t = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  for (int h = tid; h < H; h += RQ_THREADS) res[h] = latent[((size_t)b * H + h) * T + t];
  __syncthreads();
  for (int q = 0; q < Q; ++q) {
    // in_proj: e[d] = b[d] + sum_h W[d][h] res[h]     (Conv1d k = 1, modeling_dac.py:116,145)
    const float* wq = in_w + (size_t)q * D * H;
    for (int d = 0; d < D; ++d) {
      float p = 0.f;
      for (int h = tid; h < H; h += RQ_THREADS) p = fmaf(wq[(size_t)d * H + h], res[h], p);
      red[tid] = p;
      __syncthreads();
      for (int s = RQ_THREADS / 2; s > 0; s >>= 1) {
        if (tid < s) red[tid] += red[tid + s];
        __syncthreads();
      }
      if (tid == 0) e[d] = red[0] + in_b[q * D + d];
      __syncthreads();
    }
    // F.normalize (eps 1e-12), then dist = -(|e|^2 - 2 e.c) + |c|^2 over the normalised codebook (modeling_dac.py:154-166)
    if (tid == 0) {
      float n2 = 0.f;
      for (int d = 0; d < D; ++d) n2 += e[d] * e[d];
      const float inv = 1.0f / fmaxf(sqrtf(n2), 1e-12f);
      float l2 = 0.f;
      for (int d = 0; d < D; ++d) e[d] *= inv, l2 += e[d] * e[d];
      e[RQ_MAXD - 1] = l2;  // D <= RQ_MAXD - 1
    }
    __syncthreads();
    const float l2 = e[RQ_MAXD - 1];
    float best = -INFINITY;
    int besti = 0x7fffffff;
    for (int n = tid; n < N; n += RQ_THREADS) {
      const float* c = cb_norm + ((size_t)q * N + n) * (D + 1);  // normalised code, then its squared norm
      float dot = 0.f;
      for (int d = 0; d < D; ++d) dot = fmaf(e[d], c[d], dot);
      const float dist = -(l2 - 2.0f * dot) + c[D];
      if (dist > best) best = dist, besti = n;  // n ascending per thread: first index kept on ties
    }
    red[tid] = best, redi[tid] = besti;
    __syncthreads();
    for (int s = RQ_THREADS / 2; s > 0; s >>= 1) {
      if (tid < s) {
        const float ov = red[tid + s];
        const int oi = redi[tid + s];
        if (ov > red[tid] || (ov == red[tid] && oi < redi[tid])) red[tid] = ov, redi[tid] = oi;
      }
      __syncthreads();
    }
    if (tid == 0) {
      s_pick = redi[0];
      codes[((size_t)b * Q + q) * T + t] = redi[0];
    }
    __syncthreads();
    // out_proj of the picked (un-normalised) code and residual update (modeling_dac.py:150-151, :330)
    const float* cq = codebook + ((size_t)q * N + s_pick) * D;
    const float* ow = out_w + (size_t)q * H * D;
    for (int h = tid; h < H; h += RQ_THREADS) {
      float v = out_b[q * H + h];
      for (int d = 0; d < D; ++d) v = fmaf(ow[(size_t)h * D + d], cq[d], v);
      res[h] -= v;
    }
    __syncthreads();
  }
}

}  // namespace zb

using namespace zb;

extern "C" int zb200_conv1d_f32(const float* x, const float* w_ckn, const float* bias, const float* resid, float* out, int batch,
                                int c_in, int c_out, int t_in, int t_out, int k, int stride, int dilation, int pad, void* stream) {
  ZB_REQUIRE(x && w_ckn && out && batch > 0 && c_in > 0 && c_out > 0 && t_in > 0 && t_out > 0, "conv1d_f32: bad arguments");
  ZB_REQUIRE(k >= 1 && k <= CE_MAXK && stride >= 1 && stride <= 8 && dilation >= 1 && dilation <= 9 && pad >= 0,
             "conv1d_f32: kernel <= 16, stride <= 8, dilation <= 9");
  const int t_fit = (t_in + 2 * pad - dilation * (k - 1) - 1) / stride + 1;  // nn.Conv1d's output length
  ZB_REQUIRE(t_out <= t_fit, "conv1d_f32: t_out = %d exceeds the %d outputs the input of %d samples gives", t_out, t_fit, t_in);
  dim3 grid((t_out + CE_TT - 1) / CE_TT, (c_out + CE_TC - 1) / CE_TC, batch);
  conv1d_f32_kernel<<<grid, CE_THREADS, 0, (cudaStream_t)stream>>>(x, w_ckn, bias, resid, out, c_in, c_out, t_in, t_out, k, stride,
                                                                   dilation, pad);
  ZB_CHECK_LAUNCH("conv1d_f32_kernel");
  return ZB200_OK;
}

extern "C" int zb200_snake_f32(const float* x, const float* alpha, float* out, int batch, int channels, int t, void* stream) {
  ZB_REQUIRE(x && alpha && out && batch > 0 && channels > 0 && t > 0, "snake_f32: bad arguments");
  const long long n = (long long)batch * channels * t;
  const int blocks = (int)((n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16);
  snake_f32_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, alpha, out, channels, t, n);
  ZB_CHECK_LAUNCH("snake_f32_kernel");
  return ZB200_OK;
}

extern "C" int zb200_rvq_encode(const float* latent, const float* in_w, const float* in_b, const float* codebook,
                                const float* codebook_norm, const float* out_w, const float* out_b, int64_t* codes, int batch,
                                int n_q, int hidden, int dim, int n_codes, int t, void* stream) {
  ZB_REQUIRE(latent && in_w && in_b && codebook && codebook_norm && out_w && out_b && codes, "rvq_encode: null buffer");
  ZB_REQUIRE(batch > 0 && n_q > 0 && hidden > 0 && dim > 0 && dim < RQ_MAXD && n_codes > 0 && t > 0, "rvq_encode: bad shape (codebook_dim < 16)");
  const size_t smem = (size_t)(hidden + RQ_MAXD + 2 * RQ_THREADS) * sizeof(float);
  ZB_REQUIRE(smem <= 48 * 1024, "rvq_encode: hidden size %d too large", hidden);
  rvq_encode_kernel<<<dim3(t, batch), RQ_THREADS, smem, (cudaStream_t)stream>>>(latent, in_w, in_b, codebook, codebook_norm, out_w, out_b,
                                                                                 reinterpret_cast<long long*>(codes), n_q, hidden, dim,
                                                                                 n_codes, t);
  ZB_CHECK_LAUNCH("rvq_encode_kernel");
  return ZB200_OK;
}
