// Paged-KV GQA attention.
//
// attn_decode_kernel: one query position per sequence (the decode step).  Replaces F.scaled_dot_product_attention over
// the dense cache view (zonos/backbone/_torch.py:131-138) with a split-KV flash-decoding kernel over the paged pool:
//   grid (kv_head, sequence, split);  warp 4 = producer: one elected lane streams whole K and V pages (16 KB each,
//   contiguous in HBM, already chunk-swizzled by the in_proj epilogue) into a shared-memory ring with 1-D TMA bulk
//   copies + mbarriers;  warps 0-3 = consumers: each owns 16 of the 64 keys of a page, S = Q K^T and O += P V with
//   mma.sync.m16n8k16 (the 4 query heads of the GQA group occupy rows 0-3 of the M=16 tile: the op is bandwidth
//   bound, tensor throughput is irrelevant), online softmax in fp32.  Splits are merged by the last CTA to arrive.
//
// attn_prefill_kernel: causal flash attention for the prefill: 64 query positions of one q head per CTA, mma.sync.
#include "../../include/zonos_b200.h"
#include "attention.cuh"
#include "common.cuh"

namespace zb {

constexpr int PAGE = ZB200_PAGE;           // 64 positions
constexpr int HD = ZB200_HEAD_DIM;         // 128
constexpr int PAGE_BYTES = PAGE * HD * 2;  // 16 KB per (K or V, head, page)
constexpr int ATT_CONSUMER_WARPS = 4;
constexpr int ATT_THREADS = (ATT_CONSUMER_WARPS + 1) * 32;

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float* c, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                               uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <int STAGES>
struct AttSmem {
  static constexpr int RING = STAGES * 2 * PAGE_BYTES;
  static constexpr int BAR = RING;
  static constexpr int TOTAL = RING + 2 * STAGES * 8;
  static constexpr int DYN = TOTAL + 1024;
};

template <int STAGES>
__global__ void __launch_bounds__(ATT_THREADS) attn_decode_kernel(AttnDecodeArgs p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + AttSmem<STAGES>::BAR);
  uint64_t* empty_bar = full_bar + STAGES;
  __shared__ float s_o[ATT_CONSUMER_WARPS][4][HD];  // per-warp partial outputs of the 4 heads
  __shared__ float s_m[ATT_CONSUMER_WARPS][4], s_l[ATT_CONSUMER_WARPS][4];
  __shared__ int s_is_last;

  pdl_launch_dependents();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kvh = blockIdx.x, seq = blockIdx.y, split = blockIdx.z, nsplit = gridDim.z;
  const int group = p.n_heads / p.n_kv_heads;  // 4 query heads share one kv head
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], ATT_CONSUMER_WARPS);
    }
    fence_barrier_init();
  }
  // The position was written by the previous frame (complete before this graph launch started), and every KV page but
  // the one holding the current position is older than this frame: the producer streams those before the dependency on
  // the in_proj kernel is resolved; q, the current K/V row and all outputs wait for it.
  const int n_keys = (p.pos_ptr ? *p.pos_ptr : 0) + p.pos_add + 1;
  const int n_pages = (n_keys + PAGE - 1) / PAGE;
  const int pps = (n_pages + nsplit - 1) / nsplit;
  const int pg0 = min(n_pages, split * pps), pg1 = min(n_pages, pg0 + pps);

  __syncthreads();

  if (warp == ATT_CONSUMER_WARPS) {
    // ---------------------------------------------------------------------------------- producer
    if (lane == 0) {
      const uint64_t pol = l2_policy_evict_first();
      const int* bt = p.kv.block_table + seq * p.kv.pages_per_seq;
      const int cur_page = (n_keys - 1) / PAGE;
      bool waited = false;
      for (int i = 0; i < pg1 - pg0; ++i) {
        const int s = i % STAGES;
        if (!waited && pg0 + i >= cur_page) {
          pdl_wait_prior();
          waited = true;
        }
        mbar_wait(&empty_bar[s], ((i / STAGES) & 1) ^ 1);
        const int page = bt[pg0 + i];
        const size_t kbase = ((((size_t)p.kv.layer * p.kv.num_pages + page) * 2 + 0) * p.kv.n_kv_heads + kvh) * (size_t)(PAGE * HD);
        const size_t vbase = kbase + (size_t)p.kv.n_kv_heads * PAGE * HD;
        mbar_arrive_expect_tx(&full_bar[s], 2 * PAGE_BYTES);
        bulk_g2s_hint(smem + s * 2 * PAGE_BYTES, p.kv.pool + kbase, PAGE_BYTES, &full_bar[s], pol);
        bulk_g2s_hint(smem + s * 2 * PAGE_BYTES + PAGE_BYTES, p.kv.pool + vbase, PAGE_BYTES, &full_bar[s], pol);
      }
    }
  } else {
    // ---------------------------------------------------------------------------------- consumers
    pdl_wait_prior();
    // Q fragments (A operand, rows 0..group-1 are heads, the rest zero), kept in registers for the whole kernel
    const int qrow = lane >> 2;
    uint32_t qa0[8], qa2[8];
    {
      const bf16* qp = p.q + (size_t)seq * p.n_heads * HD + (size_t)(kvh * group + qrow) * HD + (lane & 3) * 2;
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        qa0[ks] = qa2[ks] = 0u;
        if (qrow < group) {
          qa0[ks] = *reinterpret_cast<const uint32_t*>(qp + ks * 16);
          qa2[ks] = *reinterpret_cast<const uint32_t*>(qp + ks * 16 + 8);
        }
      }
    }
    float o[16][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;  // for row qrow (valid on lanes 0..15)
    const float sc = p.scale * 1.4426950408889634f;

    const int key_in_page = warp * 16;
    // ldmatrix row addresses: matrix index = lane >> 3, row within matrix = lane & 7
    const int mat = lane >> 3, mrow = lane & 7;
    for (int i = 0; i < pg1 - pg0; ++i) {
      const int s = i % STAGES;
      mbar_wait(&full_bar[s], (i / STAGES) & 1);
      const uint32_t kbase = smem_u32(smem + s * 2 * PAGE_BYTES);
      const uint32_t vbase = kbase + PAGE_BYTES;

      // ---- S[16 x 16 keys] = Q K^T
      float sacc[2][4];
#pragma unroll
      for (int j = 0; j < 2; ++j) sacc[j][0] = sacc[j][1] = sacc[j][2] = sacc[j][3] = 0.f;
      {
        const int key = key_in_page + (mat >> 1) * 8 + mrow;  // matrices 0,1: keys 0-7; 2,3: keys 8-15
        const uint32_t rowaddr = kbase + key * 256;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const int chunk = ks * 2 + (mat & 1);
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4(rowaddr + ((chunk ^ (key & 7)) << 4), b0, b1, b2, b3);
          mma_bf16_16816(sacc[0], qa0[ks], 0u, qa2[ks], 0u, b0, b1);
          mma_bf16_16816(sacc[1], qa0[ks], 0u, qa2[ks], 0u, b2, b3);
        }
      }
      // ---- online softmax on rows 0..3 (c0,c1 of lanes 0..15)
      const int key0 = (pg0 + i) * PAGE + key_in_page + (lane & 3) * 2;
      float sv[4];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        sv[2 * j] = (key0 + j * 8 < n_keys) ? sacc[j][0] * sc : -INFINITY;
        sv[2 * j + 1] = (key0 + j * 8 + 1 < n_keys) ? sacc[j][1] * sc : -INFINITY;
      }
      float mx = fmaxf(fmaxf(sv[0], sv[1]), fmaxf(sv[2], sv[3]));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float m_new = fmaxf(m_run, mx);
      const float m_safe = (m_new == -INFINITY) ? 0.f : m_new;
      const float corr = exp2f(m_run - m_safe);
      float pv[4], ps = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        pv[j] = exp2f(sv[j] - m_safe);
        ps += pv[j];
      }
      ps += __shfl_xor_sync(0xffffffffu, ps, 1);
      ps += __shfl_xor_sync(0xffffffffu, ps, 2);
      l_run = l_run * corr + ps;
      m_run = m_new;
      const uint32_t pa0 = (qrow < group) ? pack_bf16x2(pv[0], pv[1]) : 0u;  // keys (lane&3)*2 + {0,1}
      const uint32_t pa2 = (qrow < group) ? pack_bf16x2(pv[2], pv[3]) : 0u;  // keys 8 + ...
      // ---- O[16 x 128] = O*corr + P V
      {
        const int key = key_in_page + (mat & 1) * 8 + mrow;  // matrices 0,2: keys 0-7; 1,3: keys 8-15
        const uint32_t rowaddr = vbase + key * 256;
#pragma unroll
        for (int nt = 0; nt < 16; nt += 2) {
          const int chunk = nt + (mat >> 1);  // 8 dims per chunk
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4_trans(rowaddr + ((chunk ^ (key & 7)) << 4), b0, b1, b2, b3);
          o[nt][0] *= corr, o[nt][1] *= corr, o[nt + 1][0] *= corr, o[nt + 1][1] *= corr;
          mma_bf16_16816(o[nt], pa0, 0u, pa2, 0u, b0, b1);
          mma_bf16_16816(o[nt + 1], pa0, 0u, pa2, 0u, b2, b3);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[s]);
    }
    // ---- publish this warp's partial (rows 0..group-1)
    if (qrow < group) {
#pragma unroll
      for (int nt = 0; nt < 16; ++nt) {
        s_o[warp][qrow][nt * 8 + (lane & 3) * 2] = o[nt][0];
        s_o[warp][qrow][nt * 8 + (lane & 3) * 2 + 1] = o[nt][1];
      }
      if ((lane & 3) == 0) s_m[warp][qrow] = m_run, s_l[warp][qrow] = l_run;
    }
  }
  __syncthreads();

  // -------------------------------------------------------------------------------------- merge the 4 warps
  // thread t < 128: head row r = t / 32, dims d = (t % 32) * 4 .. +3
  float acc[4] = {0.f, 0.f, 0.f, 0.f}, m_all = -INFINITY, l_all = 0.f;
  const int r = tid >> 5, d0 = (tid & 31) * 4;
  if (tid < 128 && r < group) {
#pragma unroll
    for (int w = 0; w < ATT_CONSUMER_WARPS; ++w) m_all = fmaxf(m_all, s_m[w][r]);
    const float m_safe = (m_all == -INFINITY) ? 0.f : m_all;
#pragma unroll
    for (int w = 0; w < ATT_CONSUMER_WARPS; ++w) {
      const float f = exp2f(s_m[w][r] - m_safe);
      l_all += s_l[w][r] * f;
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[j] += s_o[w][r][d0 + j] * f;
    }
  }
  bf16* orow = p.out + (size_t)seq * p.n_heads * HD + (size_t)(kvh * group + r) * HD + d0;
  if (nsplit == 1) {
    if (tid < 128 && r < group) {
      const float inv = 1.0f / l_all;
      *reinterpret_cast<uint2*>(orow) = make_uint2(pack_bf16x2(acc[0] * inv, acc[1] * inv), pack_bf16x2(acc[2] * inv, acc[3] * inv));
    }
    return;
  }
  // -------------------------------------------------------------------------------------- merge the splits
  const int item = seq * gridDim.x + kvh;
  float* ws = p.partial_ws + (size_t)item * nsplit * (4 * (HD + 2));
  if (tid < 128 && r < group) {
    float* mine = ws + (size_t)split * (4 * (HD + 2));
    *reinterpret_cast<float4*>(mine + r * HD + d0) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    if ((tid & 31) == 0) mine[4 * HD + r] = m_all, mine[4 * HD + 4 + r] = l_all;
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const int old = atomicAdd(&p.tickets[item], 1);
    s_is_last = (old == nsplit - 1);
    if (s_is_last) p.tickets[item] = 0;
  }
  __syncthreads();
  if (!s_is_last) return;
  __threadfence();
  if (tid < 128 && r < group) {
    float mm = -INFINITY;
    for (int sp = 0; sp < nsplit; ++sp) mm = fmaxf(mm, __ldcg(ws + (size_t)sp * (4 * (HD + 2)) + 4 * HD + r));
    const float m_safe = (mm == -INFINITY) ? 0.f : mm;
    float a4[4] = {0.f, 0.f, 0.f, 0.f}, ll = 0.f;
    for (int sp = 0; sp < nsplit; ++sp) {
      const float* part = ws + (size_t)sp * (4 * (HD + 2));
      const float f = exp2f(__ldcg(part + 4 * HD + r) - m_safe);
      ll += __ldcg(part + 4 * HD + 4 + r) * f;
      const float4 v = __ldcg(reinterpret_cast<const float4*>(part + r * HD + d0));
      a4[0] += v.x * f, a4[1] += v.y * f, a4[2] += v.z * f, a4[3] += v.w * f;
    }
    const float inv = 1.0f / ll;
    *reinterpret_cast<uint2*>(orow) = make_uint2(pack_bf16x2(a4[0] * inv, a4[1] * inv), pack_bf16x2(a4[2] * inv, a4[3] * inv));
  }
}

cudaError_t launch_attn_decode(const AttnDecodeArgs& a, int n_seq, int nsplit, cudaStream_t st) {
  constexpr int STAGES = 3;
  auto kern = attn_decode_kernel<STAGES>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, AttSmem<STAGES>::DYN);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  dim3 grid(a.n_kv_heads, n_seq, nsplit);
  return launch_pdl(kern, grid, dim3(ATT_THREADS), (size_t)AttSmem<STAGES>::DYN, st, pdl_enabled(), a);
}

// ------------------------------------------------------------------------------------------------ prefill
// Causal flash attention over the paged pool for S > 1 query positions (keys 0..pos0+s visible to query s, no padding
// mask: Appendix B.7).  Grid (query tile of 64 positions, q head, sequence); 4 warps, each 16 query rows x 128 dims
// (Q fragments in registers).  One thread streams the K and V pages of the kv head (16 KB each, already
// chunk-swizzled) into a 2-stage shared-memory ring with 1-D TMA bulk copies; S = Q K^T and O += P V run on
// mma.sync.m16n8k16 with the online softmax in fp32 (P rounded to bf16 for P V like the fused SDPA kernels).
// The op is 10 GFLOP per layer at the C3 prefill (0.7 % of the layer's flops): mma.sync is ample, what matters is
// that it is not the scalar one-warp-per-query loop it replaces (2.07 ms -> see DESIGN.md).
constexpr int PF_QTILE = 64;
constexpr int PF_THREADS = 128;
constexpr int PF_STAGES = 2;
constexpr int PF_DYN_SMEM = PF_STAGES * 2 * PAGE_BYTES + 1024;

__global__ void __launch_bounds__(PF_THREADS) attn_prefill_kernel(AttnPrefillArgs p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t full_bar[PF_STAGES];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q0 = blockIdx.x * PF_QTILE, head = blockIdx.y, seq = blockIdx.z;
  const int group = p.n_heads / p.kv.n_kv_heads, kvh = head / group;
  const int q_end = min(q0 + PF_QTILE, p.seq_len);
  const int n_pages = (p.pos0 + q_end + PAGE - 1) / PAGE;  // pages holding keys visible to the tile's last query
  if (tid == 0) {
    for (int s = 0; s < PF_STAGES; ++s) mbar_init(&full_bar[s], 1);
    fence_barrier_init();
  }
  __syncthreads();
  const int* bt = p.kv.block_table + seq * p.kv.pages_per_seq;
  auto issue = [&](int i) {
    const int s = i % PF_STAGES;
    const size_t kbase = ((((size_t)p.kv.layer * p.kv.num_pages + bt[i]) * 2 + 0) * p.kv.n_kv_heads + kvh) * (size_t)(PAGE * HD);
    const size_t vbase = kbase + (size_t)p.kv.n_kv_heads * PAGE * HD;
    mbar_arrive_expect_tx(&full_bar[s], 2 * PAGE_BYTES);
    bulk_g2s(smem + s * 2 * PAGE_BYTES, p.kv.pool + kbase, PAGE_BYTES, &full_bar[s]);
    bulk_g2s(smem + s * 2 * PAGE_BYTES + PAGE_BYTES, p.kv.pool + vbase, PAGE_BYTES, &full_bar[s]);
  };
  if (tid == 0) issue(0);

  // Q fragments (A operand of m16n8k16): rows r0 = lane/4 and r0 + 8 of the warp's 16 query positions
  const int r0 = q0 + warp * 16 + (lane >> 2), r1 = r0 + 8;
  const size_t row_stride = (size_t)p.n_heads * HD;
  uint32_t qa[8][4];
  {
    const bf16* qp0 = p.q + ((size_t)seq * p.seq_len + r0) * row_stride + (size_t)head * HD + (lane & 3) * 2;
    const bf16* qp1 = qp0 + 8 * row_stride;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      qa[ks][0] = qa[ks][1] = qa[ks][2] = qa[ks][3] = 0u;
      if (r0 < p.seq_len) {
        qa[ks][0] = *reinterpret_cast<const uint32_t*>(qp0 + ks * 16);
        qa[ks][2] = *reinterpret_cast<const uint32_t*>(qp0 + ks * 16 + 8);
      }
      if (r1 < p.seq_len) {
        qa[ks][1] = *reinterpret_cast<const uint32_t*>(qp1 + ks * 16);
        qa[ks][3] = *reinterpret_cast<const uint32_t*>(qp1 + ks * 16 + 8);
      }
    }
  }
  float o[16][4];
#pragma unroll
  for (int i = 0; i < 16; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  const float sc = p.scale * 1.4426950408889634f;
  const int lim0 = p.pos0 + r0 + 1, lim1 = p.pos0 + r1 + 1;  // keys visible to the two rows
  const int warp_last_key = p.pos0 + q0 + warp * 16 + 15;     // last key any row of this warp may see
  const int mat = lane >> 3, mrow = lane & 7;

  for (int i = 0; i < n_pages; ++i) {
    if (tid == 0 && i + 1 < n_pages) issue(i + 1);  // its stage was released by the barrier ending iteration i - 1
    mbar_wait(&full_bar[i % PF_STAGES], (i / PF_STAGES) & 1);
    const int page_key0 = i * PAGE;
    if (page_key0 <= warp_last_key) {
      const uint32_t kbase = smem_u32(smem + (i % PF_STAGES) * 2 * PAGE_BYTES);
      const uint32_t vbase = kbase + PAGE_BYTES;
      // ---- S[16 x 64 keys] = Q K^T
      float sacc[8][4];
#pragma unroll
      for (int j = 0; j < 8; ++j) sacc[j][0] = sacc[j][1] = sacc[j][2] = sacc[j][3] = 0.f;
#pragma unroll
      for (int kb = 0; kb < 4; ++kb) {
        const int key = kb * 16 + (mat >> 1) * 8 + mrow;  // matrices 0,1: keys 0-7; 2,3: keys 8-15
        const uint32_t rowaddr = kbase + key * 256;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const int chunk = ks * 2 + (mat & 1);
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4(rowaddr + ((chunk ^ (key & 7)) << 4), b0, b1, b2, b3);
          mma_bf16_16816(sacc[2 * kb], qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], b0, b1);
          mma_bf16_16816(sacc[2 * kb + 1], qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], b2, b3);
        }
      }
      // ---- causal mask + online softmax (row r0: c0,c1; row r1: c2,c3; a row lives in the 4 lanes of a quad)
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        const int kc = page_key0 + nt * 8 + (lane & 3) * 2;
        sacc[nt][0] = (kc < lim0) ? sacc[nt][0] * sc : -INFINITY;
        sacc[nt][1] = (kc + 1 < lim0) ? sacc[nt][1] * sc : -INFINITY;
        sacc[nt][2] = (kc < lim1) ? sacc[nt][2] * sc : -INFINITY;
        sacc[nt][3] = (kc + 1 < lim1) ? sacc[nt][3] * sc : -INFINITY;
        mx0 = fmaxf(mx0, fmaxf(sacc[nt][0], sacc[nt][1]));
        mx1 = fmaxf(mx1, fmaxf(sacc[nt][2], sacc[nt][3]));
      }
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
      const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
      const float ms0 = (mn0 == -INFINITY) ? 0.f : mn0, ms1 = (mn1 == -INFINITY) ? 0.f : mn1;
      const float corr0 = exp2f(m0 - ms0), corr1 = exp2f(m1 - ms1);
      float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        sacc[nt][0] = exp2f(sacc[nt][0] - ms0), sacc[nt][1] = exp2f(sacc[nt][1] - ms0);
        sacc[nt][2] = exp2f(sacc[nt][2] - ms1), sacc[nt][3] = exp2f(sacc[nt][3] - ms1);
        ps0 += sacc[nt][0] + sacc[nt][1];
        ps1 += sacc[nt][2] + sacc[nt][3];
      }
      ps0 += __shfl_xor_sync(0xffffffffu, ps0, 1);
      ps0 += __shfl_xor_sync(0xffffffffu, ps0, 2);
      ps1 += __shfl_xor_sync(0xffffffffu, ps1, 1);
      ps1 += __shfl_xor_sync(0xffffffffu, ps1, 2);
      l0 = l0 * corr0 + ps0, l1 = l1 * corr1 + ps1;
      m0 = mn0, m1 = mn1;
#pragma unroll
      for (int nt = 0; nt < 16; ++nt) o[nt][0] *= corr0, o[nt][1] *= corr0, o[nt][2] *= corr1, o[nt][3] *= corr1;
      // ---- O[16 x 128] += P V
#pragma unroll
      for (int kb = 0; kb < 4; ++kb) {
        const uint32_t pa0 = pack_bf16x2(sacc[2 * kb][0], sacc[2 * kb][1]), pa1 = pack_bf16x2(sacc[2 * kb][2], sacc[2 * kb][3]);
        const uint32_t pa2 = pack_bf16x2(sacc[2 * kb + 1][0], sacc[2 * kb + 1][1]);
        const uint32_t pa3 = pack_bf16x2(sacc[2 * kb + 1][2], sacc[2 * kb + 1][3]);
        const int key = kb * 16 + (mat & 1) * 8 + mrow;  // matrices 0,2: keys 0-7; 1,3: keys 8-15
        const uint32_t rowaddr = vbase + key * 256;
#pragma unroll
        for (int nt = 0; nt < 16; nt += 2) {
          const int chunk = nt + (mat >> 1);  // 8 dims per chunk
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4_trans(rowaddr + ((chunk ^ (key & 7)) << 4), b0, b1, b2, b3);
          mma_bf16_16816(o[nt], pa0, pa1, pa2, pa3, b0, b1);
          mma_bf16_16816(o[nt + 1], pa0, pa1, pa2, pa3, b2, b3);
        }
      }
    }
    __syncthreads();  // every warp is done with this stage before it is refilled
  }
  const float inv0 = 1.0f / l0, inv1 = 1.0f / l1;
  bf16* op0 = p.out + ((size_t)seq * p.seq_len + r0) * row_stride + (size_t)head * HD + (lane & 3) * 2;
  bf16* op1 = op0 + 8 * row_stride;
#pragma unroll
  for (int nt = 0; nt < 16; ++nt) {
    if (r0 < p.seq_len) *reinterpret_cast<uint32_t*>(op0 + nt * 8) = pack_bf16x2(o[nt][0] * inv0, o[nt][1] * inv0);
    if (r1 < p.seq_len) *reinterpret_cast<uint32_t*>(op1 + nt * 8) = pack_bf16x2(o[nt][2] * inv1, o[nt][3] * inv1);
  }
}

cudaError_t launch_attn_prefill(const AttnPrefillArgs& a, int n_seq, cudaStream_t st) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(attn_prefill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PF_DYN_SMEM);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  dim3 grid(ceil_div(a.seq_len, PF_QTILE), a.n_heads, n_seq);
  attn_prefill_kernel<<<grid, PF_THREADS, PF_DYN_SMEM, st>>>(a);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ test helper
// Writes rows of K and V [n_seq, seq_len, n_kv, 128] bf16 into the paged pool at positions pos0.. (test hook only;
// the product path writes the cache from the in_proj epilogue).
__global__ void kv_write_kernel(const bf16* __restrict__ k, const bf16* __restrict__ v, KvLayout kv, int seq_len, int pos0) {
  const int s = blockIdx.x, seq = blockIdx.y;
  const int head = threadIdx.x / 16, chunk = threadIdx.x % 16;
  const int pos = pos0 + s;
  const size_t src = (((size_t)seq * seq_len + s) * kv.n_kv_heads + head) * HD + chunk * 8;
  const int pc = (chunk ^ (pos & 7)) * 8;
  *reinterpret_cast<uint4*>(kv.row(seq, pos, 0, head) + pc) = *reinterpret_cast<const uint4*>(k + src);
  *reinterpret_cast<uint4*>(kv.row(seq, pos, 1, head) + pc) = *reinterpret_cast<const uint4*>(v + src);
}

// Decode form with the position in device memory (graph replay): one token per sequence at position *pos_ptr.
__global__ void kv_write_at_kernel(const bf16* __restrict__ k, const bf16* __restrict__ v, KvLayout kv, const int* __restrict__ pos_ptr) {
  const int seq = blockIdx.x;
  const int head = threadIdx.x / 16, chunk = threadIdx.x % 16;
  const int pos = *pos_ptr;
  if (pos < 0 || pos >= kv.pages_per_seq * ZB200_PAGE) return;
  const size_t src = ((size_t)seq * kv.n_kv_heads + head) * HD + chunk * 8;
  const int pc = (chunk ^ (pos & 7)) * 8;
  *reinterpret_cast<uint4*>(kv.row(seq, pos, 0, head) + pc) = *reinterpret_cast<const uint4*>(k + src);
  *reinterpret_cast<uint4*>(kv.row(seq, pos, 1, head) + pc) = *reinterpret_cast<const uint4*>(v + src);
}

}  // namespace zb

using namespace zb;

static KvLayout make_kv(void* pool, const int* block_table, int pages_per_seq, int num_pages, int n_kv_heads, int layer) {
  KvLayout kv;
  kv.pool = (bf16*)pool, kv.block_table = block_table, kv.pages_per_seq = pages_per_seq, kv.num_pages = num_pages;
  kv.n_kv_heads = n_kv_heads, kv.layer = layer;
  return kv;
}

extern "C" int zb200_kv_write(const void* k, const void* v, void* pool, const int* block_table, int pages_per_seq,
                              int num_pages, int n_kv_heads, int layer, int n_seq, int seq_len, int pos0, void* stream) {
  ZB_REQUIRE(n_seq > 0 && seq_len > 0 && n_kv_heads > 0 && n_kv_heads <= 64, "kv_write: bad shape");
  ZB_REQUIRE(pos0 + seq_len <= pages_per_seq * ZB200_PAGE, "kv_write: positions exceed the block table");
  dim3 grid(seq_len, n_seq);
  kv_write_kernel<<<grid, n_kv_heads * 16, 0, (cudaStream_t)stream>>>(
      (const bf16*)k, (const bf16*)v, make_kv(pool, block_table, pages_per_seq, num_pages, n_kv_heads, layer), seq_len, pos0);
  ZB_CHECK_LAUNCH("kv_write_kernel");
  return ZB200_OK;
}

extern "C" int zb200_attn_decode(const void* q, const void* pool, const int* block_table, int pages_per_seq, int num_pages,
                                 int n_heads, int n_kv_heads, int layer, int n_seq, int n_keys, int nsplit, float* ws,
                                 int* tickets, void* out, void* stream) {
  ZB_REQUIRE(n_seq > 0 && n_keys > 0 && nsplit >= 1 && n_heads % n_kv_heads == 0 && n_heads / n_kv_heads <= 4,
             "attn_decode: bad shape (GQA group must be <= 4)");
  ZB_REQUIRE(n_keys <= pages_per_seq * ZB200_PAGE, "attn_decode: n_keys exceeds the block table");
  ZB_REQUIRE(nsplit == 1 || (ws && tickets), "attn_decode: split-KV needs workspace and tickets");
  AttnDecodeArgs a;
  a.q = (const bf16*)q, a.out = (bf16*)out;
  a.kv = make_kv((void*)pool, block_table, pages_per_seq, num_pages, n_kv_heads, layer);
  a.pos_ptr = nullptr, a.pos_add = n_keys - 1, a.n_heads = n_heads, a.n_kv_heads = n_kv_heads;
  a.scale = 1.0f / sqrtf((float)ZB200_HEAD_DIM);
  a.partial_ws = ws, a.tickets = tickets;
  ZB_CHECK_CUDA(launch_attn_decode(a, n_seq, nsplit, (cudaStream_t)stream));
  return ZB200_OK;
}

extern "C" int zb200_kv_write_at(const void* k, const void* v, void* pool, const int* block_table, int pages_per_seq,
                                 int num_pages, int n_kv_heads, int layer, int n_seq, const int* pos_ptr, void* stream) {
  ZB_REQUIRE(k && v && pool && pos_ptr && n_seq > 0 && n_kv_heads > 0 && n_kv_heads <= 64, "kv_write_at: bad arguments");
  kv_write_at_kernel<<<n_seq, n_kv_heads * 16, 0, (cudaStream_t)stream>>>(
      (const bf16*)k, (const bf16*)v, make_kv(pool, block_table, pages_per_seq, num_pages, n_kv_heads, layer), pos_ptr);
  ZB_CHECK_LAUNCH("kv_write_at_kernel");
  return ZB200_OK;
}

extern "C" int zb200_attn_decode_at(const void* q, const void* pool, const int* block_table, int pages_per_seq, int num_pages,
                                    int n_heads, int n_kv_heads, int layer, int n_seq, const int* pos_ptr, int nsplit, float* ws,
                                    int* tickets, void* out, void* stream) {
  ZB_REQUIRE(n_seq > 0 && pos_ptr && nsplit >= 1 && n_heads % n_kv_heads == 0 && n_heads / n_kv_heads <= 4,
             "attn_decode_at: bad shape (GQA group must be <= 4)");
  ZB_REQUIRE(nsplit == 1 || (ws && tickets), "attn_decode_at: split-KV needs workspace and tickets");
  AttnDecodeArgs a;
  a.q = (const bf16*)q, a.out = (bf16*)out;
  a.kv = make_kv((void*)pool, block_table, pages_per_seq, num_pages, n_kv_heads, layer);
  a.pos_ptr = pos_ptr, a.pos_add = 0, a.n_heads = n_heads, a.n_kv_heads = n_kv_heads;
  a.scale = 1.0f / sqrtf((float)ZB200_HEAD_DIM);
  a.partial_ws = ws, a.tickets = tickets;
  ZB_CHECK_CUDA(launch_attn_decode(a, n_seq, nsplit, (cudaStream_t)stream));
  return ZB200_OK;
}

extern "C" int zb200_attn_prefill(const void* q, const void* pool, const int* block_table, int pages_per_seq,
                                  int num_pages, int n_heads, int n_kv_heads, int layer, int n_seq, int seq_len, int pos0,
                                  void* out, void* stream) {
  ZB_REQUIRE(n_seq > 0 && seq_len > 0 && n_heads % n_kv_heads == 0, "attn_prefill: bad shape");
  ZB_REQUIRE(pos0 + seq_len <= pages_per_seq * ZB200_PAGE, "attn_prefill: positions exceed the block table");
  AttnPrefillArgs a;
  a.q = (const bf16*)q, a.out = (bf16*)out;
  a.kv = make_kv((void*)pool, block_table, pages_per_seq, num_pages, n_kv_heads, layer);
  a.seq_len = seq_len, a.pos0 = pos0, a.n_heads = n_heads, a.scale = 1.0f / sqrtf((float)ZB200_HEAD_DIM);
  ZB_CHECK_CUDA(launch_attn_prefill(a, n_seq, (cudaStream_t)stream));
  return ZB200_OK;
}

// ---------------------------------------------------------------------------------------------- allocation-free variant
// zb200_attn_decode_noalloc: the split-KV decode attention with its scratch (split partials + merge tickets) in a
// file-scope device array instead of caller-allocated buffers, for callers that cannot allocate (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Every GPU that loads the library holds the array (8.14 MiB) until exit.
//
// In HOST code the array's name is a host-side stand-in, not a device address (a kernel handed it faults or, where the
// GPU can reach host memory, silently uses the wrong memory): the device address is resolved once per device, with that
// device current, by cudaGetSymbolAddress below, and that pointer is what the kernel receives.
namespace zb {
constexpr int NOALLOC_MAX_UNITS = ZB200_ATTN_NOALLOC_MAX_UNITS;  // n_seq * n_kv_heads * nsplit
constexpr int NOALLOC_UNIT_FLOATS = 520;                         // 4 heads x (128 + 2) floats per split partial
constexpr int NOALLOC_MAX_DEVICES = 64;
struct AttnNoallocScratch {
  float partial[(size_t)NOALLOC_MAX_UNITS * NOALLOC_UNIT_FLOATS];
  int tickets[NOALLOC_MAX_UNITS];
};
static __device__ __align__(256) AttnNoallocScratch g_attn_noalloc_scratch;

}  // namespace zb

#include <mutex>

namespace {
std::mutex g_noalloc_mu;
zb::AttnNoallocScratch* g_noalloc_ptr[zb::NOALLOC_MAX_DEVICES] = {};  // device address per device ordinal

// Device address of the scratch array on the CURRENT device (resolved on first use on that device).
int noalloc_scratch(zb::AttnNoallocScratch** out) {
  int dev = -1;
  ZB_CHECK_CUDA(cudaGetDevice(&dev));
  ZB_REQUIRE(dev >= 0 && dev < zb::NOALLOC_MAX_DEVICES, "attn_decode_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    zb::AttnNoallocScratch* p = nullptr;
    ZB_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, zb::g_attn_noalloc_scratch));
    ZB_REQUIRE(p != nullptr, "attn_decode_noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return ZB200_OK;
}
}  // namespace

extern "C" int zb200_attn_decode_noalloc(const void* q, const void* pool, const int* block_table, int pages_per_seq,
                                         int num_pages, int n_heads, int n_kv_heads, int layer, int n_seq, int n_keys,
                                         int nsplit, void* out, void* stream) {
  using namespace zb;
  ZB_REQUIRE(q && pool && block_table && out, "attn_decode_noalloc: null buffer");
  ZB_REQUIRE(n_seq > 0 && n_keys > 0 && nsplit >= 1 && n_kv_heads > 0 && n_heads % n_kv_heads == 0 && n_heads / n_kv_heads <= 4,
             "attn_decode_noalloc: bad shape (GQA group must be <= 4)");
  ZB_REQUIRE(n_keys <= pages_per_seq * ZB200_PAGE, "attn_decode_noalloc: n_keys exceeds the block table");
  const long long units = (long long)n_seq * n_kv_heads * nsplit;
  ZB_REQUIRE(units <= NOALLOC_MAX_UNITS,
             "attn_decode_noalloc: n_seq*n_kv_heads*nsplit = %lld exceeds the built-in scratch (ZB200_ATTN_NOALLOC_MAX_UNITS = %d); "
             "use zb200_attn_decode with caller-allocated scratch",
             units, NOALLOC_MAX_UNITS);
  AttnNoallocScratch* s = nullptr;
  const int rc = noalloc_scratch(&s);
  if (rc != ZB200_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // The array is shared by every call and keeps the last call's contents.  The partials are fully overwritten before
  // they are read; the merge tickets must start at zero (they reset themselves, but a call that was aborted half-way
  // would leave them dirty), so they are cleared on the caller's stream -- a memset, not an allocation.
  if (nsplit > 1) ZB_CHECK_CUDA(cudaMemsetAsync(s->tickets, 0, sizeof(int) * (size_t)n_seq * n_kv_heads, st));
  AttnDecodeArgs a;
  a.q = (const bf16*)q, a.out = (bf16*)out;
  a.kv = make_kv((void*)pool, block_table, pages_per_seq, num_pages, n_kv_heads, layer);
  a.pos_ptr = nullptr, a.pos_add = n_keys - 1, a.n_heads = n_heads, a.n_kv_heads = n_kv_heads;
  a.scale = 1.0f / sqrtf((float)ZB200_HEAD_DIM);
  a.partial_ws = s->partial, a.tickets = s->tickets;
  ZB_CHECK_CUDA(launch_attn_decode(a, n_seq, nsplit, st));
  return ZB200_OK;
}
