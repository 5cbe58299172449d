// Experiment: HBM -> shared-memory streaming rate of 1-D TMA bulk copies (cp.async.bulk, UBLKCP) as a function of the
// grid size, the copy size and the bytes kept in flight per CTA.  No consumer work: the ceiling of every
// weight-streaming kernel of the decode step.
//   nvcc -gencode arch=compute_100a,code=sm_100a -I zonos_b200/csrc -o scripts/_bin/exp_stream scripts/exp_stream.cu
#include <cstdio>
#include <cstdlib>

#include "common.cuh"

namespace zb {
void set_error(const char*, ...) {}
int cuda_fail(cudaError_t, const char*) { return -1; }
}  // namespace zb
using namespace zb;

// Each CTA streams `per_cta` contiguous bytes starting at base + cta*per_cta (streams == 1) or two interleaved
// streams `stream_gap` bytes apart (streams == 2, like two adjacent n_tiles of the packed weights).
__global__ void __launch_bounds__(64) stream_kernel(const uint8_t* base, long long per_cta, int copy_bytes, int depth, int streams,
                                                    long long stream_gap, int evict_first, int issuers, int same_warp) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t full_all[64];
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < 64; ++s) mbar_init(&full_all[s], 1);
    fence_barrier_init();
  }
  __syncthreads();
  // issuer q: lane q of warp 0 (same_warp) or lane 0 of warp q; each owns depth/issuers ring slots and 1/issuers of the range
  const int q = same_warp ? tid : (tid >> 5);
  const bool is_issuer = same_warp ? (tid < issuers) : ((tid & 31) == 0 && q < issuers);
  if (is_issuer) {
    const int my_depth = depth / issuers;
    uint64_t* full = full_all + q * my_depth;
    uint8_t* ring = smem + (size_t)q * my_depth * copy_bytes;
    const uint64_t pol = l2_policy_evict_first();
    const long long my_bytes = per_cta * streams / issuers / copy_bytes * copy_bytes;
    const uint8_t* src = base + (long long)blockIdx.x * per_cta * streams + q * my_bytes;
    const long long n = my_bytes / copy_bytes;
    long long issued = 0, done = 0;
    while (done < n) {
      while (issued < n && issued - done < my_depth) {
        const int s = (int)(issued % my_depth);
        mbar_arrive_expect_tx(&full[s], copy_bytes);
        const uint8_t* p = src + issued * copy_bytes;
        if (evict_first)
          bulk_g2s_hint(ring + (size_t)s * copy_bytes, p, copy_bytes, &full[s], pol);
        else
          bulk_g2s(ring + (size_t)s * copy_bytes, p, copy_bytes, &full[s]);
        ++issued;
      }
      mbar_wait(&full[done % my_depth], (uint32_t)((done / my_depth) & 1));
      ++done;
    }
  }
}

// fc1-like: CTA c owns 512 KB = two adjacent 256 KB tile rows, read as 32 units of (8 KB from row 0, 8 KB from row 1);
// `skew`: CTA c starts at unit (c * skew) % 32 and wraps (breaks the lockstep of the address pattern across CTAs).
__global__ void __launch_bounds__(64) fc1_like_kernel(const uint8_t* base, int lanes, int skew, int depth_units) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t full[16];
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < 16; ++s) mbar_init(&full[s], 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (tid < lanes) {
    const uint64_t pol = l2_policy_evict_first();
    const uint8_t* src = base + (size_t)blockIdx.x * 524288;
    const int start = (blockIdx.x * skew) % 32;
    // lane l owns ring slots l, l+lanes, ... and units l, l+lanes, ...
    for (int i = tid; i < 32; i += lanes) {
      const int s = i % depth_units;
      if (i >= depth_units) mbar_wait(&full[s], (uint32_t)(((i / depth_units) - 1) & 1));  // previous use landed (no consumer)
      mbar_arrive_expect_tx(&full[s], 16384);
      const int ku = (start + i) % 32;
      bulk_g2s_hint(smem + s * 16384, src + (size_t)ku * 8192, 8192, &full[s], pol);
      bulk_g2s_hint(smem + s * 16384 + 8192, src + 262144 + (size_t)ku * 8192, 8192, &full[s], pol);
    }
    for (int i = 32 - depth_units + tid; i < 32; i += lanes)
      if (i >= 0 && (i % lanes) == tid % lanes) {}
  }
  __syncthreads();
  if (tid == 0)
    for (int s = 0; s < depth_units; ++s) {
      const int uses = (32 - s + depth_units - 1) / depth_units;
      mbar_wait(&full[s], (uint32_t)((uses - 1) & 1));
    }
}

int main() {
  const long long total = 4096ll << 20;  // 4 GB buffer, > L2
  uint8_t* d;
  if (cudaMalloc(&d, total) != cudaSuccess) return 1;
  cudaMemset(d, 1, total);
  cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024 + 1024);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  {
    cudaFuncSetAttribute(fc1_like_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024 + 1024);
    for (int lanes : {1, 4})
      for (int skew : {0, 1, 5, 13})
        for (int depth : {8}) {
          const int g = 128;
          float best = 1e9f;
          for (int rep = 0; rep < 6; ++rep) {
            cudaEventRecord(e0);
            fc1_like_kernel<<<g, 64, depth * 16384 + 1024>>>(d + (size_t)rep * 67108864 * 2, lanes, skew, depth);
            cudaEventRecord(e1);
            if (cudaDeviceSynchronize() != cudaSuccess) {
              printf("CUDA error %s\n", cudaGetErrorString(cudaGetLastError()));
              return 1;
            }
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (rep > 0 && ms < best) best = ms;
          }
          printf("fc1-like grid 128 (64 MB, 2 x 8 KB copies per unit) lanes %d skew %2d depth %d: %.2f us, %.0f GB/s\n", lanes, skew, depth,
                 best * 1e3, 67108864.0 / best / 1e6);
        }
  }
  const int grids[] = {128};
  const int copies[] = {8192, 16384, 32768};
  const int inflight_kb[] = {128};
  const int streams = 1;
  for (int mode = 0; mode < 4; ++mode)  // issuers: 1, 2 lanes, 4 lanes, 2 warps (wait: 64 threads = 2 warps), evict-normal
    for (int g : grids)
      for (int cb : copies)
        for (int fk : inflight_kb) {
          const int issuers = mode == 1 ? 2 : mode == 2 ? 4 : mode == 3 ? 2 : 1;
          const int same_warp = mode != 3;
          const int evict_first = mode != 4;
          const int ctas_per_sm = g > 148 ? 2 : 1;
          if (fk * ctas_per_sm > 200) continue;
          const int depth = fk * 1024 / cb;
          if (depth > 64 || depth < issuers) continue;
          const long long bytes = 1024ll << 20;  // 1 GB per launch
          long long per_cta = bytes / g / streams / cb * cb;
          const long long gap = per_cta;  // second stream right after the first
          const int smem = fk * 1024 + 1024;
          for (int rep = 0; rep < 2; ++rep) {
            cudaEventRecord(e0);
            // rotate over the buffer so that nothing is L2-resident
            stream_kernel<<<g, 64, smem>>>(d + (rep * bytes) % (total - bytes), per_cta, cb, depth, streams, gap, evict_first, issuers, same_warp);
            cudaEventRecord(e1);
            if (cudaDeviceSynchronize() != cudaSuccess) {
              printf("CUDA error %s\n", cudaGetErrorString(cudaGetLastError()));
              return 1;
            }
          }
          float ms;
          cudaEventElapsedTime(&ms, e0, e1);
          const double moved = (double)per_cta * streams * g;
          printf("mode %d (issuers %d %s%s) grid %3d copy %2d KB in-flight/CTA %3d KB: %6.0f GB/s (%.1f GB/s per SM)\n", mode, issuers,
                 same_warp ? "lanes" : "warps", evict_first ? "" : ", no hint", g, cb / 1024, fk,
                 moved / ms / 1e6, moved / ms / 1e6 / (g > 148 ? 148 : g));
        }
  return 0;
}
