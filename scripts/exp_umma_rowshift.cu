// Experiment: may the A-operand UMMA descriptor (SWIZZLE_128B, K-major) start at an arbitrary 128-byte row of a
// larger swizzled tile?  The DAC convolutions want this: tap i of a dilated conv is the SAME activation tile
// shifted by i*dilation rows, so one halo tile in shared memory could feed all taps.
//   nvcc -gencode arch=compute_100a,code=sm_100a -I zonos_b200/csrc -o scripts/_bin/exp_umma_rowshift scripts/exp_umma_rowshift.cu
// Prints, per (row shift, variant), the number of mismatching outputs.  variant 0: descriptor base-offset field 0,
// variant 1: base offset = (start address >> 7) & 7.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "common.cuh"

namespace zb {
void set_error(const char*, ...) {}
int cuda_fail(cudaError_t, const char*) { return -1; }
}  // namespace zb
using namespace zb;

constexpr int A_ROWS = 192, N = 64;

__global__ void __launch_bounds__(128) shift_kernel(const __half* A, const __half* B, float* D, int r0, int variant) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sa = smem;                  // A_ROWS x 128 B
  uint8_t* sb = smem + A_ROWS * 128;   // N x 128 B   (A_ROWS*128 is a multiple of 1024)
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int idx = tid; idx < A_ROWS * 8; idx += 128) {
    const int r = idx >> 3, c = idx & 7;
    *reinterpret_cast<uint4*>(sa + r * 128 + ((c ^ (r & 7)) << 4)) = *reinterpret_cast<const uint4*>(A + r * 64 + c * 8);
  }
  for (int idx = tid; idx < N * 8; idx += 128) {
    const int r = idx >> 3, c = idx & 7;
    *reinterpret_cast<uint4*>(sb + r * 128 + ((c ^ (r & 7)) << 4)) = *reinterpret_cast<const uint4*>(B + r * 64 + c * 8);
  }
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(&slot, 64);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  if (tid == 0) {
    const uint32_t idesc = umma_idesc(128, N, 0);
    const uint32_t a_addr = smem_u32(sa) + r0 * 128;
    uint64_t adesc = umma_desc_sw128(a_addr);
    if (variant == 1) adesc |= (uint64_t)((a_addr >> 7) & 7) << 49;
    const uint64_t bdesc = umma_desc_sw128(smem_u32(sb));
    for (int k = 0; k < 4; ++k) umma_f16(tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, k != 0);
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  float acc[64];
  const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
  tmem_ld32(taddr, acc);
  tmem_ld32(taddr + 32, acc + 32);
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int n = 0; n < N; ++n) D[tid * N + n] = acc[n];
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 64);
}

int main() {
  std::vector<__half> hA(A_ROWS * 64), hB(N * 64);
  std::vector<float> fA(A_ROWS * 64), fB(N * 64);
  srand(1);
  for (size_t i = 0; i < hA.size(); ++i) fA[i] = (float)(rand() % 7 - 3), hA[i] = __float2half(fA[i]);
  for (size_t i = 0; i < hB.size(); ++i) fB[i] = (float)(rand() % 5 - 2), hB[i] = __float2half(fB[i]);
  __half *dA, *dB;
  float* dD;
  cudaMalloc(&dA, hA.size() * 2), cudaMalloc(&dB, hB.size() * 2), cudaMalloc(&dD, 128 * N * 4);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  const int smem = (A_ROWS + N) * 128 + 1024;
  cudaFuncSetAttribute(shift_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  std::vector<float> hD(128 * N);
  const int shifts[] = {0, 8, 1, 2, 3, 5, 7, 9, 18, 27, 54, 63};
  for (int variant = 0; variant < 2; ++variant)
    for (int r0 : shifts) {
      cudaMemset(dD, 0, 128 * N * 4);
      shift_kernel<<<1, 128, smem>>>(dA, dB, dD, r0, variant);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) {
        printf("variant %d shift %d: CUDA error %s\n", variant, r0, cudaGetErrorString(e));
        return 1;
      }
      cudaMemcpy(hD.data(), dD, 128 * N * 4, cudaMemcpyDeviceToHost);
      int bad = 0;
      for (int i = 0; i < 128; ++i)
        for (int n = 0; n < N; ++n) {
          float ref = 0.f;
          for (int k = 0; k < 64; ++k) ref += fA[(r0 + i) * 64 + k] * fB[n * 64 + k];
          bad += (ref != hD[i * N + n]);
        }
      printf("variant %d shift %2d: %d mismatches of %d\n", variant, r0, bad, 128 * N);
    }
  return 0;
}
