// Fused per-row sampler + the device-side EOS state machine of Zonos.generate.
//
// One CTA (1024 threads, 2 logits each) per (utterance, codebook) row does, in shared memory / registers:
//   logit bias + EOS masks (zonos/model.py:321-324,353,358-361,387) -> windowed repetition penalty
//   (zonos/sampling.py:131-169) -> temperature softmax -> unified sampler (:54-75) -> top-p (:96-111, bitonic sort +
//   scan) -> top-k (:77-93) -> min-p (:114-128) -> exponential-race draw argmax(p/q) (:11-33) or argmax (:326).
// The last CTA of a pass to finish (ticket) runs the per-utterance bookkeeping that the reference does on the host
// with >= 3+B device syncs per step (model.py:345-424): first-EOS detection, grace period, whole-batch re-draw flag,
// remaining-steps clamp, diagonal EOS fill, frame write, loop-termination flag.  A step is therefore two launches
// (pass 1, pass 2) and no host round trip.
#include "../../include/zonos_b200.h"
#include "common.cuh"
#include "epilogues.cuh"  // StepStateIdx
#include "sampler.cuh"

namespace zb {

constexpr int SMP_THREADS = 1024;
constexpr int SMP_MAXV = 2048;
constexpr int EOS_ID = 1024, MASK_ID = 1025, UNKNOWN_ID = -1, GRACE = 6;

// ------------------------------------------------------------------------------------------------ block reductions
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = red[threadIdx.x & 31];
  return warp_sum(t);  // 32 warps -> every lane holds one partial
}
__device__ __forceinline__ float block_max(float v, float* red) {
  v = warp_max(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = red[threadIdx.x & 31];
  return warp_max(t);
}
// argmax with first-index tie break (torch.argmax semantics); returns the index to every thread
__device__ __forceinline__ int block_argmax(float v0, int i0, float v1, int i1, float* red, int* redi) {
  float bv = v0;
  int bi = i0;
  if (v1 > bv || (v1 == bv && i1 < bi)) bv = v1, bi = i1;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > bv || (ov == bv && oi < bi)) bv = ov, bi = oi;
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = bv, redi[threadIdx.x >> 5] = bi;
  __syncthreads();
  bv = red[threadIdx.x & 31], bi = redi[threadIdx.x & 31];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > bv || (ov == bv && oi < bi)) bv = ov, bi = oi;
  }
  return bi;
}

// ------------------------------------------------------------------------------------------------ Philox4x32-10
__device__ __forceinline__ uint32_t philox_first(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0, c1 = lo1, c2 = n2, c3 = lo0;
    k0 += 0x9E3779B9u, k1 += 0xBB67AE85u;
  }
  return c0;
}
// Exp(1) variate for (seed, draw, global row, vocab index)
__device__ __forceinline__ float exp1_noise(unsigned long long seed, uint32_t draw, uint32_t grow, uint32_t v) {
  const uint32_t x = philox_first(v, draw, grow, 0x5A6F6E6Fu, (uint32_t)seed, (uint32_t)(seed >> 32));
  // 24 random bits: (k + 0.5) * 2^-24 is exact in fp32 and strictly inside (0, 1), so q = -log(u) > 0 like
  // torch.exponential_ (which never returns 0).  ((float)x + 0.5f) * 2^-32 rounds to 1.0f for the top 128 values of x.
  const float u = ((float)(x >> 8) + 0.5f) * 5.9604644775390625e-08f;
  return -logf(u);
}

// ------------------------------------------------------------------------------------------------ the row sampler
// returns the sampled index (valid in every thread)
__device__ int sample_row(const SamplerArgs& a, int b, int q, int pass, int offset, int draw, float* sval, int* sidx,
                          float* red, int* redi) {
  const int V = a.vocab, t = threadIdx.x;
  const int v0 = t, v1 = t + SMP_THREADS;
  const float* lrow = a.logits + ((size_t)b * a.n_q + q) * V;
  float x0 = (v0 < V) ? lrow[v0] : -INFINITY, x1 = (v1 < V) ? lrow[v1] : -INFINITY;

  if (a.loop_mode) {
    // logit_bias (model.py:321-324): EOS forbidden on codebooks 1..8, -ln(1024) on codebook 0
    const bool masked = (q > 0) || (pass == 1 ? a.eos_mode[b] && a.grace[b] > 0 : a.mask2[b] != 0);
    if (v0 == EOS_ID) x0 = masked ? -INFINITY : x0 - 6.931471805599453f;
    if (v1 == EOS_ID) x1 = masked ? -INFINITY : x1 - 6.931471805599453f;
    // repetition penalty over the last `window` columns of the delayed codes (sampling.py:131-169)
    const bool in_eos = (pass == 1) ? a.eos_mode[b] != 0 : a.eos_old[b] != 0;
    const float rp = in_eos ? 1.0f : a.rep_penalty;
    {  // factor 1 is exact, so rows with rp == 1 need no special case
      const int w = min(a.rep_window, offset);
      const int32_t* hist = a.delayed + ((size_t)b * a.n_q + q) * a.t_delayed + (offset - w);
      float f0 = 1.f, f1 = 1.f;
      for (int j = 0; j < w; ++j) {
        const int tok = min(hist[j], V - 1);
        if (tok == v0) f0 *= rp;
        if (tok == v1) f1 *= rp;
      }
      x0 = (x0 <= 0.f) ? x0 * f0 : x0 / f0;
      x1 = (x1 <= 0.f) ? x1 * f1 : x1 / f1;
    }
  }
  if (a.temperature <= 0.f) return block_argmax(x0, v0, x1, v1 < V ? v1 : SMP_MAXV, red, redi);

  // ---- softmax(logits / T)
  x0 = x0 / a.temperature, x1 = x1 / a.temperature;
  float mx = block_max(fmaxf(x0, x1), red);
  float p0 = (v0 < V) ? expf(x0 - mx) : 0.f, p1 = (v1 < V) ? expf(x1 - mx) : 0.f;
  float sum = block_sum(p0 + p1, red);
  p0 /= sum, p1 /= sum;

  // ---- unified sampler (sampling.py:54-75)
  if (a.linear > 0.f) {
    const float lp0 = logf(fmaxf(p0, 1e-20f)), lp1 = logf(fmaxf(p1, 1e-20f));
    const float ent = -block_sum((v0 < V ? p0 * lp0 : 0.f) + (v1 < V ? p1 * lp1 : 0.f), red);
    const float lin = a.linear + ent * a.conf;
    const float r0 = (v0 < V) ? lp0 * lin - lp0 * lp0 * a.quad : -INFINITY;
    const float r1 = (v1 < V) ? lp1 * lin - lp1 * lp1 * a.quad : -INFINITY;
    mx = block_max(fmaxf(r0, r1), red);
    p0 = (v0 < V) ? expf(r0 - mx) : 0.f, p1 = (v1 < V) ? expf(r1 - mx) : 0.f;
    sum = block_sum(p0 + p1, red);
    p0 /= sum, p1 /= sum;
  }

  // ---- top-p / top-k need the descending order: bitonic sort of (p, index) in shared memory
  if (a.top_p > 0.f || a.top_k > 0) {
    __syncthreads();
    sval[v0] = p0, sidx[v0] = v0;
    sval[v1] = (v1 < V) ? p1 : -1.f, sidx[v1] = v1;  // padding sorts last
    __syncthreads();
    for (int k = 2; k <= SMP_MAXV; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));  // lower index of the pair
        const int l = i | j;
        const bool desc = (i & k) == 0;
        const float a0 = sval[i], a1 = sval[l];
        const int i0 = sidx[i], i1 = sidx[l];
        // "greater" = larger p, ties by smaller index
        const bool first_greater = (a0 > a1) || (a0 == a1 && i0 < i1);
        if (first_greater != desc) sval[i] = a1, sval[l] = a0, sidx[i] = i1, sidx[l] = i0;
        __syncthreads();
      }
    }
    // my two sorted slots are 2t, 2t+1 (consecutive -> simple block scan)
    float s0 = fmaxf(sval[2 * t], 0.f), s1 = fmaxf(sval[2 * t + 1], 0.f);
    if (a.top_p > 0.f) {
      // inclusive scan of the sorted probabilities
      const float pair = s0 + s1;
      float inc = pair;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float n = __shfl_up_sync(0xffffffffu, inc, o);
        if ((t & 31) >= o) inc += n;
      }
      __syncthreads();
      if ((t & 31) == 31) red[t >> 5] = inc;
      __syncthreads();
      float wsum = (t & 31) < 32 ? red[t & 31] : 0.f, wprefix = wsum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float n = __shfl_up_sync(0xffffffffu, wprefix, o);
        if ((t & 31) >= o) wprefix += n;
      }
      const float warp_excl = __shfl_sync(0xffffffffu, wprefix - wsum, t >> 5);
      const float before0 = warp_excl + inc - pair;  // mass strictly before slot 2t
      const float before1 = before0 + s0;
      if (before0 > a.top_p) s0 = 0.f;
      if (before1 > a.top_p) s1 = 0.f;
      const float tot = block_sum(s0 + s1, red);
      s0 /= tot, s1 /= tot;
    }
    if (a.top_k > 0) {
      __syncthreads();
      sval[2 * t] = s0, sval[2 * t + 1] = s1;
      __syncthreads();
      const float kth = sval[min(a.top_k, V) - 1];
      if (s0 < kth) s0 = 0.f;
      if (s1 < kth) s1 = 0.f;
      const float tot = block_sum(s0 + s1, red);
      s0 /= tot, s1 /= tot;
    }
    // scatter back to vocabulary order
    const int j0 = sidx[2 * t], j1 = sidx[2 * t + 1];
    __syncthreads();
    sval[j0] = s0, sval[j1] = s1;
    __syncthreads();
    p0 = sval[v0], p1 = (v1 < V) ? sval[v1] : 0.f;
  }

  // ---- min-p (sampling.py:114-128)
  if (a.min_p > 0.f) {
    const float top = block_max(fmaxf(p0, p1), red);
    if (p0 < a.min_p * top) p0 = 0.f;
    if (p1 < a.min_p * top) p1 = 0.f;
    sum = block_sum(p0 + p1, red);
    p0 /= sum, p1 /= sum;
  }

  // ---- exponential race: argmax(p / q), q ~ Exp(1)
  float q0, q1;
  if (a.noise) {
    const float* nrow = a.noise + (((size_t)draw * a.batch + b) * a.n_q + q) * V;
    q0 = (v0 < V) ? nrow[v0] : 1.f, q1 = (v1 < V) ? nrow[v1] : 1.f;
  } else {
    const uint32_t grow = (uint32_t)((a.row_base + b) * a.n_q + q);
    q0 = exp1_noise(a.seed, (uint32_t)draw, grow, v0), q1 = exp1_noise(a.seed, (uint32_t)draw, grow, v1);
  }
  // p / q (sampling.py:26-28).  A zero-probability entry (masked EOS, the pad slot, anything a filter removed) scores
  // exactly 0 like in the reference -- also when injected noise is 0 there (0 / 0 would be NaN) -- and a NaN never wins.
  float r0 = (v0 < V) ? (p0 > 0.f ? p0 / q0 : 0.f) : -INFINITY, r1 = (v1 < V) ? (p1 > 0.f ? p1 / q1 : 0.f) : -INFINITY;
  if (!(r0 == r0)) r0 = -INFINITY;
  if (!(r1 == r1)) r1 = -INFINITY;
  return block_argmax(r0, v0, r1, v1 < V ? v1 : SMP_MAXV, red, redi);
}

// ------------------------------------------------------------------------------------------------ bookkeeping
// after pass 1: detect first EOS per utterance, start the grace period, request the whole-batch re-draw
__device__ void update_after_pass1(const SamplerArgs& a) {
  __shared__ int s_any;
  if (threadIdx.x == 0) s_any = 0;
  __syncthreads();
  for (int b = threadIdx.x; b < a.batch; b += blockDim.x) {
    const bool mode = a.eos_mode[b] != 0;
    const bool active = mode && a.grace[b] > 0;  // model.py:360
    if (active) a.grace[b] -= 1;                 // :362
    const bool hit = a.tok1[b * a.n_q] == EOS_ID;
    const bool fresh = hit && !mode;             // :379
    a.mask2[b] = (active || fresh) ? 1 : 0;
    a.eos_old[b] = mode ? 1 : 0;
    if (fresh) {
      a.eos_mode[b] = 1;  // :383-384
      a.grace[b] = GRACE;
      atomicOr(&s_any, 1);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    a.state[ST_ANY_FRESH] = s_any;
    if (a.temperature > 0.f) a.state[ST_DRAW] += 1;
  }
}

// after pass 2: remaining-steps clamp, diagonal fill, frame write, counters, termination flag (model.py:398-424)
__device__ void finalize_step(const SamplerArgs& a, int offset) {
  __shared__ int s_max;
  if (threadIdx.x == 0) s_max = INT_MIN;
  __syncthreads();
  const int any = a.state[ST_ANY_FRESH];
  const int32_t* tok = any ? a.tok2 : a.tok1;
  const int Q = a.n_q;
  for (int b = threadIdx.x; b < a.batch; b += blockDim.x) {
    int rem = a.remaining[b];
    const bool hit = tok[b * Q] == EOS_ID;
    if (hit) rem = min(rem, Q), a.stopping[b] = 1;
    const bool stop = a.stopping[b] != 0;
    const int diag = min(Q - rem, Q - 1);
    for (int q = 0; q < Q; ++q) {
      int tk = tok[b * Q + q];
      if (stop) tk = (q < diag) ? MASK_ID : (q == diag ? EOS_ID : tk);
      // the reference's last iteration addresses column T+Q, an empty slice (model.py:417-418): nothing is written
      if (offset < a.t_delayed) {
        int32_t* cell = a.delayed + ((size_t)b * Q + q) * a.t_delayed + offset;
        if (*cell == UNKNOWN_ID) *cell = tk;
      }
    }
    rem -= 1;
    a.remaining[b] = rem;
    atomicMax(&s_max, rem);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (any && a.temperature > 0.f) a.state[ST_DRAW] += 1;
    a.state[ST_OFFSET] = offset + 1;
    a.state[ST_POS] += 1;
    a.state[ST_STEP] += 1;
    a.state[ST_ANY_FRESH] = 0;
    if (s_max <= 0) {
      a.state[ST_DONE] = 1;
      a.state[ST_FINAL_OFFSET] = offset;
    }
    if (a.host_flags) {  // pinned mirror the host polls between graph launches
      a.host_flags[0] = a.state[ST_DONE];
      a.host_flags[1] = a.state[ST_STEP];
    }
  }
}

// first frame after the prefill (model.py:304-314): plain write, no state machine
__device__ void finalize_first(const SamplerArgs& a, int offset) {
  const int Q = a.n_q;
  for (int i = threadIdx.x; i < a.batch * Q; i += blockDim.x) {
    int32_t* cell = a.delayed + (size_t)i * a.t_delayed + offset;
    if (*cell == UNKNOWN_ID) *cell = a.tok1[i];
  }
  if (threadIdx.x == 0 && a.temperature > 0.f) a.state[ST_DRAW] += 1;
}

__global__ void __launch_bounds__(SMP_THREADS) sampler_kernel(SamplerArgs a, int pass) {
  __shared__ float sval[SMP_MAXV];
  __shared__ int sidx[SMP_MAXV];
  __shared__ float red[32];
  __shared__ int redi[32];
  __shared__ int s_last;
  pdl_launch_dependents();
  pdl_wait_prior();
  const int q = blockIdx.x, b = blockIdx.y;
  if (a.loop_mode && a.state[ST_DONE]) return;  // steps launched past the end of the loop are no-ops
  const int offset = a.loop_mode ? a.state[ST_OFFSET] : a.first_offset;
  const bool need = (pass == 1) || a.state[ST_ANY_FRESH] != 0;
  if (need) {
    const int draw = a.state[ST_DRAW];
    const int tk = sample_row(a, b, q, pass, offset, draw, sval, sidx, red, redi);
    if (threadIdx.x == 0) (pass == 1 ? a.tok1 : a.tok2)[b * a.n_q + q] = tk;
  }
  // ticket: the last CTA of this pass runs the per-utterance bookkeeping
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int idx = pass == 1 ? ST_TICKET_A : ST_TICKET_B;
    const int old = atomicAdd(&a.state[idx], 1);
    s_last = (old == (int)(gridDim.x * gridDim.y) - 1);
    if (s_last) a.state[idx] = 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (!a.loop_mode) {
    finalize_first(a, offset);
  } else if (pass == 1) {
    update_after_pass1(a);
  } else {
    finalize_step(a, offset);
  }
}

cudaError_t launch_sampler(const SamplerArgs& a, int pass, cudaStream_t st) {
  dim3 grid(a.n_q, a.batch);
  return launch_pdl(sampler_kernel, grid, dim3(SMP_THREADS), 0, st, pdl_enabled(), a, pass);
}

}  // namespace zb

using namespace zb;

extern "C" int zb200_sample(const zb200_sampler_args* u, int pass, void* stream) {
  ZB_REQUIRE(u && u->batch > 0 && u->n_q > 0 && u->n_q <= ZB200_MAX_CODEBOOKS, "sample: bad batch/codebooks");
  ZB_REQUIRE(u->vocab > 1025 && u->vocab <= SMP_MAXV, "sample: vocab %d must be in (1025, 2048]", u->vocab);
  ZB_REQUIRE(pass == 1 || pass == 2, "sample: pass must be 1 or 2");
  ZB_REQUIRE(u->logits && u->state && u->tok1 && u->tok2 && u->delayed, "sample: null buffer");
  ZB_REQUIRE(!u->loop_mode || (u->remaining && u->stopping && u->eos_mode && u->grace && u->mask2 && u->eos_old),
             "sample: loop mode needs the per-utterance state arrays");
  SamplerArgs a;
  a.logits = u->logits, a.noise = u->noise, a.delayed = u->delayed, a.tok1 = u->tok1, a.tok2 = u->tok2;
  a.state = u->state, a.remaining = u->remaining, a.grace = u->grace;
  a.stopping = u->stopping, a.eos_mode = u->eos_mode, a.mask2 = u->mask2, a.eos_old = u->eos_old;
  a.host_flags = u->host_flags;
  a.noise = u->noise;
  a.batch = u->batch, a.n_q = u->n_q, a.vocab = u->vocab, a.t_delayed = u->t_delayed;
  a.loop_mode = u->loop_mode, a.first_offset = u->first_offset, a.row_base = u->row_base;
  a.seed = u->seed, a.temperature = u->temperature, a.top_p = u->top_p, a.top_k = u->top_k, a.min_p = u->min_p;
  a.linear = u->linear, a.conf = u->conf, a.quad = u->quad, a.rep_penalty = u->rep_penalty, a.rep_window = u->rep_window;
  ZB_CHECK_CUDA(launch_sampler(a, pass, (cudaStream_t)stream));
  return ZB200_OK;
}
