// Experiment: cycles per tcgen05.mma (M=128, K=16, fp16, SWIZZLE_128B K-major operands from shared memory) as a
// function of N and of the row shift of the A descriptor (the DAC convolution taps).  One CTA, one issuing thread.
//   nvcc -gencode arch=compute_100a,code=sm_100a -I zonos_b200/csrc -o scripts/_bin/exp_umma_rate scripts/exp_umma_rate.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "common.cuh"

namespace zb {
void set_error(const char*, ...) {}
int cuda_fail(cudaError_t, const char*) { return -1; }
}  // namespace zb
using namespace zb;

constexpr int A_ROWS = 192;

// mode 0: every MMA reads the same A k-slice; mode 1: cycles through the 4 k-slices and 7 "taps" (shift*tap rows)
__global__ void __launch_bounds__(128) rate_kernel(long long* out, int N, int shift, int n_mma, int mode, int b_rows_shift) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sa = smem;
  uint8_t* sb = smem + A_ROWS * 128;
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < (A_ROWS + 256 + 64) * 128 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(&slot, 256);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  if (tid == 0) {
    const uint32_t idesc = umma_idesc(128, N, 0);
    const long long t0 = clock64();
    for (int i = 0; i < n_mma; ++i) {
      const int k = mode ? (i & 3) : 0;
      const int tap = mode ? ((i >> 2) % 7) : 1;
      const uint64_t adesc = umma_desc_sw128(smem_u32(sa) + tap * shift * 128) + (uint64_t)(k * 2);
      const uint64_t bdesc = umma_desc_sw128(smem_u32(sb) + b_rows_shift * 128) + (uint64_t)(k * 2);
      umma_f16(tmem, adesc, bdesc, idesc, i != 0);
    }
    const long long t1 = clock64();
    umma_commit(&bar);
    mbar_wait(&bar, 0);
    const long long t2 = clock64();
    out[0] = t1 - t0, out[1] = t2 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  const int smem = (A_ROWS + 256 + 64) * 128 + 1024;
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int n_mma = 2048;
  const int Ns[] = {32, 64, 96, 128, 192, 256};
  const int shifts[] = {0, 1, 3, 8, 9};
  for (int mode = 0; mode < 2; ++mode)
    for (int N : Ns)
      for (int shift : shifts) {
        if (mode == 1 && shift * 6 + 128 > A_ROWS) continue;
        long long h[2];
        for (int rep = 0; rep < 2; ++rep) {
          rate_kernel<<<1, 128, smem>>>(d, N, shift, n_mma, mode, 0);
          if (cudaDeviceSynchronize() != cudaSuccess) {
            printf("CUDA error\n");
            return 1;
          }
        }
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("mode %d N %3d shift %d: issue %.1f clk/MMA, complete %.1f clk/MMA\n", mode, N, shift, (double)h[0] / n_mma,
               (double)h[1] / n_mma);
      }
  // B operand with a row shift (control)
  for (int bs : {0, 3}) {
    long long h[2];
    rate_kernel<<<1, 128, smem>>>(d, 96, 0, n_mma, 0, bs);
    cudaDeviceSynchronize();
    cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("B row shift %d, N 96: complete %.1f clk/MMA\n", bs, (double)h[1] / n_mma);
  }
  return 0;
}
