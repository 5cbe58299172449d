// Delay-pattern apply / revert on int64 code matrices (bit-exact index shuffles).
// Reference: zonos/codebook_pattern.py:5-7 (pad by n_q with mask, roll row k by k+1) and :10-12.
#include "../../include/zonos_b200.h"
#include "common.cuh"

namespace zb {

// out[b,k,t] = codes[b,k,t-(k+1)] if 0 <= t-(k+1) < T else mask;   t in [0, T+Q)
__global__ void delay_apply_kernel(const int64_t* __restrict__ codes, int64_t* __restrict__ out, int rows, int Q, int T,
                                   int64_t mask) {
  const int Tp = T + Q;
  const long long total = (long long)rows * Tp;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int row = (int)(i / Tp), t = (int)(i % Tp);
    const int s = t - (row % Q + 1);
    out[i] = (s >= 0 && s < T) ? codes[(long long)row * T + s] : mask;
  }
}

// out[b,k,t] = delayed[b,k,t+k+1];   t in [0, Tp-Q)
__global__ void delay_revert_kernel(const int64_t* __restrict__ delayed, int64_t* __restrict__ out, int rows, int Q,
                                    int Tp) {
  const int T = Tp - Q;
  const long long total = (long long)rows * T;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int row = (int)(i / T), t = (int)(i % T);
    out[i] = delayed[(long long)row * Tp + t + row % Q + 1];
  }
}

}  // namespace zb

extern "C" int zb200_delay_apply(const int64_t* codes, int64_t* delayed, int batch, int n_q, int t, int64_t mask_token,
                                 void* stream) {
  ZB_REQUIRE(batch >= 0 && n_q > 0 && t >= 0, "delay_apply: bad shape B=%d Q=%d T=%d", batch, n_q, t);
  const long long total = (long long)batch * n_q * (t + n_q);
  if (total == 0) return ZB200_OK;
  const int threads = 256;
  const int blocks = (int)((total + threads - 1) / threads < 148 * 8 ? (total + threads - 1) / threads : 148 * 8);
  zb::delay_apply_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(codes, delayed, batch * n_q, n_q, t, mask_token);
  ZB_CHECK_LAUNCH("delay_apply_kernel");
  return ZB200_OK;
}

extern "C" int zb200_delay_revert(const int64_t* delayed, int64_t* codes, int batch, int n_q, int t_delayed, void* stream) {
  ZB_REQUIRE(batch >= 0 && n_q > 0 && t_delayed >= n_q, "delay_revert: bad shape B=%d Q=%d Tp=%d", batch, n_q, t_delayed);
  const long long total = (long long)batch * n_q * (t_delayed - n_q);
  if (total == 0) return ZB200_OK;
  const int threads = 256;
  const int blocks = (int)((total + threads - 1) / threads < 148 * 8 ? (total + threads - 1) / threads : 148 * 8);
  zb::delay_revert_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(delayed, codes, batch * n_q, n_q, t_delayed);
  ZB_CHECK_LAUNCH("delay_revert_kernel");
  return ZB200_OK;
}
