// Speaker encoder: SpeakerEmbeddingLDA.forward (zonos/speaker_cloning.py:387-411) ->
// SpeakerEmbedding.forward (:375-384: mono mix + torchaudio Resample to 16 kHz) -> ResNet293_based.forward (:216-223:
// logFbankCal :12-34, ResNet293 of SimAMBasicBlock :63-95,152-195, ASP :37-60, bottleneck Linear) -> LDA Linear.
//
// The reference runs this model in fp32, once per new speaker; the kernels here are fp32 too (CUDA-core FMA, no tensor-core
// rounding), channels-first tensors like torch:
//   resample_f32_kernel    polyphase sinc FIR of torchaudio's Resample (kernel table built on the host from its formula)
//   logfbank_kernel        one CTA per frame: reflect-padded, Hann-windowed 512-point DFT by direct summation (fp64
//                          accumulators, twiddles from a 512-entry table), power, 80 mel bands, log(. + 1e-6)
//   sub_time_mean_kernel   the features minus their mean over time
//   conv2d_f32_kernel      3x3 (pad 1) / 1x1 convolution with stride 1 or 2 as an implicit GEMM on the CUDA cores: 128 output
//                          positions of one output row x 64 output channels per CTA, 8 x 4 outputs per thread, input rows and
//                          the weight slab of 8 input channels staged in shared memory; eval-mode BatchNorm folded into a
//                          per-channel scale / shift in the epilogue, optional ReLU
//   simam_residual_kernel  SimAM (:90-95) on a whole (sample, channel) plane + shortcut add + ReLU
//   relu_affine_kernel     ReLU followed by a per-channel affine (ASP's Conv1d -> ReLU -> BatchNorm1d)
//   asp_pool_kernel        softmax over time of the attention logits, weighted mean and standard deviation (:54-57)
//   linear_f32_kernel      nn.Linear on one vector per sample (bottleneck, LDA)
// The 1x1 Conv1d layers of the pooling run on zb200_conv1d_f32 (dac_encoder.cu).
#include "../../include/zonos_b200.h"
#include "common.cuh"

namespace zb {

// ------------------------------------------------------------------------------------------------ resample
// out[b][j * new_f + q] = sum_k kern[q][k] * x[b][j * orig_f + k - width]   (zero outside [0, n_in))
__global__ void __launch_bounds__(256) resample_f32_kernel(const float* __restrict__ x, const float* __restrict__ kern,
                                                           float* __restrict__ out, int n_in, int n_out, int orig_f, int new_f,
                                                           int width, int taps) {
  const int o = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
  if (o >= n_out) return;
  const int j = o / new_f, q = o - j * new_f;
  const float* kq = kern + (size_t)q * taps;
  const float* xb = x + (size_t)b * n_in;
  const int i0 = j * orig_f - width;
  float acc = 0.f;
  for (int k = 0; k < taps; ++k) {
    const int i = i0 + k;
    if (i >= 0 && i < n_in) acc = fmaf(__ldg(kq + k), xb[i], acc);
  }
  out[(size_t)b * n_out + o] = acc;
}

// ------------------------------------------------------------------------------------------------ log mel filterbank
constexpr int FB_NFFT = 512, FB_WIN = 400, FB_HOP = 160, FB_BINS = 257, FB_MELS = 80, FB_THREADS = 288;
__global__ void __launch_bounds__(FB_THREADS) logfbank_kernel(const float* __restrict__ wav, int n, const float* __restrict__ window,
                                                              const float* __restrict__ mel_fb, float* __restrict__ out, int n_frames) {
  __shared__ double tw_c[FB_NFFT], tw_s[FB_NFFT];
  __shared__ float frame[FB_WIN];
  __shared__ float power[FB_BINS];
  const int t = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  const float* w = wav + (size_t)b * n;
  constexpr int OFF = (FB_NFFT - FB_WIN) / 2;  // torch.stft centres the window inside the n_fft frame
  for (int i = tid; i < FB_NFFT; i += FB_THREADS) sincospi(2.0 * (double)i / FB_NFFT, &tw_s[i], &tw_c[i]);
  for (int i = tid; i < FB_WIN; i += FB_THREADS) {
    int s = t * FB_HOP + OFF + i - FB_NFFT / 2;  // centre = True: the signal is reflect-padded by n_fft / 2 on both sides
    if (s < 0) s = -s;
    if (s >= n) s = 2 * (n - 1) - s;
    frame[i] = w[s] * window[i];
  }
  __syncthreads();
  if (tid < FB_BINS) {
    double re = 0.0, im = 0.0;
    for (int i = 0; i < FB_WIN; ++i) {
      const int ph = (tid * (i + OFF)) & (FB_NFFT - 1);
      const double v = (double)frame[i];
      re += v * tw_c[ph];
      im -= v * tw_s[ph];
    }
    power[tid] = (float)(re * re + im * im);
  }
  __syncthreads();
  if (tid < FB_MELS) {
    float acc = 0.f;
    for (int f = 0; f < FB_BINS; ++f) acc = fmaf(power[f], __ldg(mel_fb + f * FB_MELS + tid), acc);
    out[((size_t)b * FB_MELS + tid) * n_frames + t] = logf(acc + 1e-6f);
  }
}

__global__ void __launch_bounds__(256) sub_time_mean_kernel(float* __restrict__ x, int t) {
  __shared__ float red[8];
  __shared__ float mean;
  float* row = x + (size_t)blockIdx.x * t;
  float s = 0.f;
  for (int i = threadIdx.x; i < t; i += 256) s += row[i];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f;
    for (int k = 0; k < 8; ++k) a += red[k];
    mean = a / (float)t;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < t; i += 256) row[i] -= mean;
}

// ------------------------------------------------------------------------------------------------ conv2d + folded BatchNorm
// CTA = 16 PT consecutive output positions of one output row x 64 output channels, 256 threads: thread (tx, ty) owns the PT
// CONSECUTIVE positions PT tx .. PT tx + PT - 1 and the 4 channels 4 ty .. 4 ty + 3.  Per input channel and filter row it
// reads its (PT - 1) S + K input values once (shared-memory index p is stored at p + p / 32, which keeps the 16 position lanes
// of a warp on different banks) and one float4 of weights per tap: 13 shared-memory reads for 96 FMAs at PT = 8, stride 1.
// The slab of the NEXT 8 input channels is fetched into registers while the current one is multiplied (one exposed global
// round trip per launch instead of one per slab).  PT = 8 / 4 / 2 (128 / 64 / 32 positions) is chosen per launch so that
// the grid has several CTAs per SM: one clip is a small problem (251 x 20 positions x 256 channels in the deep stage).
constexpr int C2_TC = 64, C2_CI = 8, C2_THREADS = 256;
__device__ __forceinline__ int c2_pad(int p) { return p + (p >> 5); }
template <int K, int S, int PT>
__global__ void __launch_bounds__(C2_THREADS, 2) conv2d_f32_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                                const float* __restrict__ scale, const float* __restrict__ shift,
                                                                float* __restrict__ out, int C_in, int H, int W, int C_out, int Ho,
                                                                int Wo, int relu) {
  constexpr int TW = 16 * PT, SPAN = (TW - 1) * S + K, XS_W = SPAN + SPAN / 32 + 1, KK = K * K, NX = (PT - 1) * S + K;
  constexpr int XN = C2_CI * K * SPAN, NXL = (XN + C2_THREADS - 1) / C2_THREADS;
  constexpr int WN = C2_CI * KK * C2_TC, NWL = (WN + C2_THREADS - 1) / C2_THREADS;
  __shared__ float xs[C2_CI * K][XS_W];
  __shared__ __align__(16) float ws[C2_CI * KK][C2_TC];
  const int n_ct = (C_out + C2_TC - 1) / C2_TC;
  const int wo0 = blockIdx.x * TW, ho = blockIdx.y, b = blockIdx.z / n_ct, c0 = (blockIdx.z % n_ct) * C2_TC;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  constexpr int pad = K / 2;
  const int gx0 = wo0 * S - pad, gy0 = ho * S - pad;
  const float* xb = x + (size_t)b * C_in * H * W;
  float acc[PT][4];
#pragma unroll
  for (int i = 0; i < PT; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  float xreg[NXL], wreg[NWL];
  auto fetch = [&](int ci0) {
#pragma unroll
    for (int k = 0; k < NXL; ++k) {
      const int idx = tid + k * C2_THREADS, row = idx / SPAN, p = idx - row * SPAN, ci = row / K, r = row - ci * K;
      const int gy = gy0 + r, gx = gx0 + p;
      xreg[k] = (idx < XN && ci0 + ci < C_in && gy >= 0 && gy < H && gx >= 0 && gx < W) ? xb[((size_t)(ci0 + ci) * H + gy) * W + gx] : 0.f;
    }
#pragma unroll
    for (int k = 0; k < NWL; ++k) {  // w is [C_in][K*K][C_out]: row = ci * KK + tap
      const int idx = tid + k * C2_THREADS, co = idx & (C2_TC - 1), row = idx / C2_TC;
      wreg[k] = (idx < WN && ci0 * KK + row < C_in * KK && c0 + co < C_out) ? __ldg(w + ((size_t)ci0 * KK + row) * C_out + c0 + co) : 0.f;
    }
  };
  fetch(0);
  for (int ci0 = 0; ci0 < C_in; ci0 += C2_CI) {
#pragma unroll
    for (int k = 0; k < NXL; ++k) {
      const int idx = tid + k * C2_THREADS, row = idx / SPAN, p = idx - row * SPAN;
      if (idx < XN) xs[row][c2_pad(p)] = xreg[k];
    }
#pragma unroll
    for (int k = 0; k < NWL; ++k) {
      const int idx = tid + k * C2_THREADS;
      if (idx < WN) ws[idx / C2_TC][idx & (C2_TC - 1)] = wreg[k];
    }
    __syncthreads();
    if (ci0 + C2_CI < C_in) fetch(ci0 + C2_CI);
#pragma unroll 2
    for (int ci = 0; ci < C2_CI; ++ci)
#pragma unroll
      for (int kh = 0; kh < K; ++kh) {
        float xr[NX];
#pragma unroll
        for (int e = 0; e < NX; ++e) xr[e] = xs[ci * K + kh][c2_pad(tx * PT * S + e)];
#pragma unroll
        for (int kw = 0; kw < K; ++kw) {
          const float4 wv = *reinterpret_cast<const float4*>(&ws[ci * KK + kh * K + kw][ty * 4]);
#pragma unroll
          for (int i = 0; i < PT; ++i) {
            const float xv = xr[i * S + kw];
            acc[i][0] = fmaf(xv, wv.x, acc[i][0]);
            acc[i][1] = fmaf(xv, wv.y, acc[i][1]);
            acc[i][2] = fmaf(xv, wv.z, acc[i][2]);
            acc[i][3] = fmaf(xv, wv.w, acc[i][3]);
          }
        }
      }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int co = c0 + ty * 4 + j;
    if (co >= C_out) continue;
    const float sc = scale ? scale[co] : 1.f, sh = shift ? shift[co] : 0.f;
    float* o = out + (((size_t)b * C_out + co) * Ho + ho) * Wo;
#pragma unroll
    for (int i = 0; i < PT; ++i) {
      const int wo = wo0 + tx * PT + i;
      if (wo >= Wo) continue;
      float v = fmaf(acc[i][j], sc, sh);
      if (relu) v = fmaxf(v, 0.f);
      o[wo] = v;
    }
  }
}

template <int K, int S>
static void launch_conv2d(const float* x, const float* w, const float* scale, const float* shift, float* out, int batch, int c_in,
                          int h, int wd, int c_out, int ho, int wo, int relu, cudaStream_t st) {
  const int n_ct = ceil_div(c_out, C2_TC);
  const long long rows = (long long)ho * batch * n_ct;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const long long want = 4LL * sms;  // CTAs for an even spread over the SMs
  const int pt = ceil_div(wo, 128) * rows >= want ? 8 : ceil_div(wo, 64) * rows >= want ? 4 : 2;
  const dim3 grid((unsigned)ceil_div(wo, 16 * pt), (unsigned)ho, (unsigned)(batch * n_ct));
  if (pt == 8) conv2d_f32_kernel<K, S, 8><<<grid, C2_THREADS, 0, st>>>(x, w, scale, shift, out, c_in, h, wd, c_out, ho, wo, relu);
  else if (pt == 4) conv2d_f32_kernel<K, S, 4><<<grid, C2_THREADS, 0, st>>>(x, w, scale, shift, out, c_in, h, wd, c_out, ho, wo, relu);
  else conv2d_f32_kernel<K, S, 2><<<grid, C2_THREADS, 0, st>>>(x, w, scale, shift, out, c_in, h, wd, c_out, ho, wo, relu);
}

// ------------------------------------------------------------------------------------------------ SimAM + shortcut + ReLU
__device__ __forceinline__ float block_sum_256(float v, float* red, float* bc) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) a += red[k];
    *bc = a;
  }
  __syncthreads();
  return *bc;
}

// one CTA per (sample, channel) plane of hw values:  n = hw - 1, d = (x - mean)^2, v = sum(d) / n,
//   out = relu(x * sigmoid(d / (4 (v + lambda)) + 0.5) + shortcut)            (speaker_cloning.py:85-95)
__global__ void __launch_bounds__(256) simam_residual_kernel(const float* __restrict__ x, const float* __restrict__ shortcut,
                                                             float* __restrict__ out, int hw, float lambda_p) {
  __shared__ float red[8];
  __shared__ float bc;
  const float* xp = x + (size_t)blockIdx.x * hw;
  const float* sp = shortcut + (size_t)blockIdx.x * hw;
  float* op = out + (size_t)blockIdx.x * hw;
  float s = 0.f;
  for (int i = threadIdx.x; i < hw; i += 256) s += xp[i];
  const float mean = block_sum_256(s, red, &bc) / (float)hw;
  float ss = 0.f;
  for (int i = threadIdx.x; i < hw; i += 256) {
    const float t = xp[i] - mean;
    ss = fmaf(t, t, ss);
  }
  const float v = block_sum_256(ss, red, &bc) / (float)(hw - 1);
  const float inv = 1.0f / (4.0f * (v + lambda_p));
  for (int i = threadIdx.x; i < hw; i += 256) {
    const float xv = xp[i], t = xv - mean;
    const float e = fmaf(t * t, inv, 0.5f);
    const float y = xv * (1.0f / (1.0f + expf(-e))) + sp[i];
    op[i] = fmaxf(y, 0.f);
  }
}

// y[b][c][t] = relu(x) * scale[c] + shift[c]
__global__ void relu_affine_kernel(const float* __restrict__ x, const float* __restrict__ scale, const float* __restrict__ shift,
                                   float* __restrict__ out, int C, int T, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)((i / T) % C);
    out[i] = fmaf(fmaxf(x[i], 0.f), scale[c], shift[c]);
  }
}

// ------------------------------------------------------------------------------------------------ attentive statistics pooling
// grid (channel, sample), 128 threads: w = softmax_t(logits), mu = sum_t x w, sg = sqrt(max(sum_t x^2 w - mu^2, 1e-5));
// out[b][c] = mu, out[b][C + c] = sg                                            (speaker_cloning.py:54-57)
__global__ void __launch_bounds__(128) asp_pool_kernel(const float* __restrict__ x, const float* __restrict__ logits,
                                                       float* __restrict__ out, int C, int T) {
  __shared__ float red[4];
  __shared__ float bc;
  const int c = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  const float* xr = x + ((size_t)b * C + c) * T;
  const float* lr = logits + ((size_t)b * C + c) * T;
  float m = -INFINITY;
  for (int t = tid; t < T; t += 128) m = fmaxf(m, lr[t]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((tid & 31) == 0) red[tid >> 5] = m;
  __syncthreads();
  m = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  float se = 0.f, s1 = 0.f, s2 = 0.f;
  for (int t = tid; t < T; t += 128) {
    const float e = expf(lr[t] - m), xv = xr[t];
    se += e;
    s1 = fmaf(xv, e, s1);
    s2 = fmaf(xv * xv, e, s2);
  }
  se = block_sum_256(se, red, &bc);
  s1 = block_sum_256(s1, red, &bc);
  s2 = block_sum_256(s2, red, &bc);
  if (tid == 0) {
    const float mu = s1 / se;
    out[(size_t)b * 2 * C + c] = mu;
    out[(size_t)b * 2 * C + C + c] = sqrtf(fmaxf(s2 / se - mu * mu, 1e-5f));
  }
}

// ------------------------------------------------------------------------------------------------ Linear on one vector
// out[b][co] = bias[co] + sum_ci x[b][ci] * w[ci][co]   (w = the nn.Linear weight transposed).  grid (ceil(C_out / 32), batch),
// 256 threads = 32 output columns x 8 slices of the reduction; coalesced weight rows, the slices summed through shared memory.
__global__ void __launch_bounds__(256) linear_f32_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                         const float* __restrict__ bias, float* __restrict__ out, int C_in, int C_out) {
  __shared__ float part[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5, co = blockIdx.x * 32 + tx, b = blockIdx.y;
  const float* xb = x + (size_t)b * C_in;
  float acc = 0.f;
  if (co < C_out) {
#pragma unroll 8
    for (int ci = ty; ci < C_in; ci += 8) acc = fmaf(xb[ci], __ldg(w + (size_t)ci * C_out + co), acc);
  }
  part[ty][tx] = acc;
  __syncthreads();
  if (ty == 0 && co < C_out) {
    float s = bias ? bias[co] : 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += part[k][tx];
    out[(size_t)b * C_out + co] = s;
  }
}

}  // namespace zb

using namespace zb;

extern "C" int zb200_linear_f32(const float* x, const float* w_t, const float* bias, float* out, int batch, int c_in, int c_out,
                                void* stream) {
  ZB_REQUIRE(x && w_t && out && batch > 0 && batch <= 65535 && c_in > 0 && c_out > 0, "linear_f32: bad arguments");
  linear_f32_kernel<<<dim3((unsigned)ceil_div(c_out, 32), (unsigned)batch), 256, 0, (cudaStream_t)stream>>>(x, w_t, bias, out, c_in, c_out);
  ZB_CHECK_LAUNCH("linear_f32_kernel");
  return ZB200_OK;
}

extern "C" int zb200_resample_f32(const float* x, const float* kernel, float* out, int batch, int n_in, int n_out, int orig_freq,
                                  int new_freq, int width, void* stream) {
  ZB_REQUIRE(x && kernel && out && batch > 0 && n_in > 0 && n_out > 0 && orig_freq > 0 && new_freq > 0 && width >= 0,
             "resample_f32: bad arguments");
  ZB_REQUIRE((long long)n_out <= ((long long)n_in * new_freq + orig_freq - 1) / orig_freq,
             "resample_f32: n_out=%d exceeds ceil(n_in * new / orig)", n_out);
  resample_f32_kernel<<<dim3((unsigned)ceil_div(n_out, 256), (unsigned)batch), 256, 0, (cudaStream_t)stream>>>(
      x, kernel, out, n_in, n_out, orig_freq, new_freq, width, 2 * width + orig_freq);
  ZB_CHECK_LAUNCH("resample_f32_kernel");
  return ZB200_OK;
}

extern "C" int zb200_logfbank_f32(const float* wav16k, const float* window, const float* mel_fb, float* out, int batch, int n,
                                  int n_frames, void* stream) {
  ZB_REQUIRE(wav16k && window && mel_fb && out && batch > 0, "logfbank_f32: bad arguments");
  ZB_REQUIRE(n > FB_NFFT / 2 && n_frames == 1 + n / FB_HOP, "logfbank_f32: n=%d needs more than %d samples and n_frames = 1 + n / %d", n,
             FB_NFFT / 2, FB_HOP);
  logfbank_kernel<<<dim3((unsigned)n_frames, (unsigned)batch), FB_THREADS, 0, (cudaStream_t)stream>>>(wav16k, n, window, mel_fb, out,
                                                                                                       n_frames);
  ZB_CHECK_LAUNCH("logfbank_kernel");
  sub_time_mean_kernel<<<(unsigned)(batch * FB_MELS), 256, 0, (cudaStream_t)stream>>>(out, n_frames);
  ZB_CHECK_LAUNCH("sub_time_mean_kernel");
  return ZB200_OK;
}

extern "C" int zb200_conv2d_f32(const float* x, const float* w_ckn, const float* scale, const float* shift, float* out, int batch,
                                int c_in, int h, int w, int c_out, int k, int stride, int relu, void* stream) {
  ZB_REQUIRE(x && w_ckn && out && batch > 0 && c_in > 0 && c_out > 0 && h > 0 && w > 0, "conv2d_f32: bad arguments");
  ZB_REQUIRE((k == 1 || k == 3) && (stride == 1 || stride == 2), "conv2d_f32: k=%d stride=%d (3x3 pad 1 or 1x1, stride 1 or 2)", k, stride);
  const int pad = k / 2, ho = (h + 2 * pad - k) / stride + 1, wo = (w + 2 * pad - k) / stride + 1;
  ZB_REQUIRE(ho <= 65535 && (long long)batch * ceil_div(c_out, C2_TC) <= 65535, "conv2d_f32: grid too large");
  cudaStream_t st = (cudaStream_t)stream;
  if (k == 3 && stride == 1) launch_conv2d<3, 1>(x, w_ckn, scale, shift, out, batch, c_in, h, w, c_out, ho, wo, relu, st);
  else if (k == 3) launch_conv2d<3, 2>(x, w_ckn, scale, shift, out, batch, c_in, h, w, c_out, ho, wo, relu, st);
  else if (stride == 1) launch_conv2d<1, 1>(x, w_ckn, scale, shift, out, batch, c_in, h, w, c_out, ho, wo, relu, st);
  else launch_conv2d<1, 2>(x, w_ckn, scale, shift, out, batch, c_in, h, w, c_out, ho, wo, relu, st);
  ZB_CHECK_LAUNCH("conv2d_f32_kernel");
  return ZB200_OK;
}

extern "C" int zb200_simam_residual_f32(const float* x, const float* shortcut, float* out, int planes, int hw, float lambda_p,
                                        void* stream) {
  ZB_REQUIRE(x && shortcut && out && planes > 0 && hw > 1, "simam_residual_f32: bad arguments");
  simam_residual_kernel<<<(unsigned)planes, 256, 0, (cudaStream_t)stream>>>(x, shortcut, out, hw, lambda_p);
  ZB_CHECK_LAUNCH("simam_residual_kernel");
  return ZB200_OK;
}

extern "C" int zb200_relu_affine_f32(const float* x, const float* scale, const float* shift, float* out, int batch, int channels,
                                     int t, void* stream) {
  ZB_REQUIRE(x && scale && shift && out && batch > 0 && channels > 0 && t > 0, "relu_affine_f32: bad arguments");
  const long long n = (long long)batch * channels * t;
  const int grid = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
  relu_affine_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(x, scale, shift, out, channels, t, n);
  ZB_CHECK_LAUNCH("relu_affine_kernel");
  return ZB200_OK;
}

extern "C" int zb200_asp_pool_f32(const float* x, const float* logits, float* out, int batch, int channels, int t, void* stream) {
  ZB_REQUIRE(x && logits && out && batch > 0 && channels > 0 && t > 0 && batch <= 65535, "asp_pool_f32: bad arguments");
  asp_pool_kernel<<<dim3((unsigned)channels, (unsigned)batch), 128, 0, (cudaStream_t)stream>>>(x, logits, out, channels, t);
  ZB_CHECK_LAUNCH("asp_pool_kernel");
  return ZB200_OK;
}
