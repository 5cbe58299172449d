// DACAutoencoder.codes_to_wavs post-processing on a ragged batch of decoded waveforms (zonos/autoencoder.py:188-245):
// BS.1770 integrated loudness as pyloudnorm 0.1.1 computes it (normalize_loudness, autoencoder.py:172-186), the
// energy trim (trim_silence, autoencoder.py:49-90) and the two fades (autoencoder.py:233-241).
//
// The only sequential part is the K-weighting IIR (two cascaded biquads).  It is split into chunks of `chunk` samples:
//   1. kw_chunk<state>: every chunk filtered from a zero state -> its end state (one thread per chunk, fp64)
//   2. kw_chunk_scan:   s_in[c+1] = M^chunk s_in[c] + s_end[c]            (one warp per utterance, 32 chunks per
//                       5-step warp scan)
//   3. kw_chunk<energy>: every chunk filtered again from its true start state; the K-weighted sample is rounded to
//                       fp32 (pyloudnorm filters in place in the float32 input array), squared, and summed per
//                       SEGMENT -- the pieces between consecutive gating-block edges that fall inside the chunk
//   4. kw_segment_sum:  the pieces of every segment added up in chunk order (deterministic, no atomics)
//   5. loudness_trim:   block energies (a block is a run of segments), gating, LUFS, gain, then the frame energies of
//                       the scaled signal and the trim bounds
//   6. apply:           out = ((x * gain) * fade_in) * fade_out over the kept range
// All of it is HBM-bound byte streaming: the signal is read three times and written once (fp32), 16 B per sample
// against the 12 B a single-pass filter would need.
#include "../../include/zonos_b200.h"
#include "common.cuh"

#include <math.h>

namespace zb {
namespace {

struct KwCoef {
  double b0a, b1a, b2a, a1a, a2a;
  double b0b, b1b, b2b, a1b, a2b;
};

// One sample through both transposed direct-form-II biquads (scipy.signal.lfilter's structure).  pyloudnorm writes the
// first stage's output back into the float32 array before the second stage runs, hence the rounding in between.
__device__ __forceinline__ double kw_step(const KwCoef& c, double x, double (&s)[4]) {
  const double ya = fma(c.b0a, x, s[0]);
  s[0] = fma(-c.a1a, ya, fma(c.b1a, x, s[1]));
  s[1] = fma(-c.a2a, ya, c.b2a * x);
  const double xb = (double)(float)ya;
  const double yb = fma(c.b0b, xb, s[2]);
  s[2] = fma(-c.a1b, yb, fma(c.b1b, xb, s[3]));
  s[3] = fma(-c.a2b, yb, c.b2b * xb);
  return yb;
}

constexpr int KW_THREADS = 128;

constexpr int KW_MAX_PIECES = 8;  // segments one chunk may touch (host checks the tables against it)

// ENERGY == false: end state of the chunk from a zero start state.  ENERGY == true: the chunk from its true start state,
// sum of squares per segment piece -> part[b][c][0..], index of the chunk's first segment -> seg0[b][c].
template <bool ENERGY>
__global__ void __launch_bounds__(KW_THREADS)
kw_chunk_kernel(const float* __restrict__ wav, int64_t pitch, const int32_t* __restrict__ n_samples,
                const int32_t* __restrict__ blk_sel, int chunk, int chunks_pitch, KwCoef coef,
                double* __restrict__ s_end, const double* __restrict__ s_in, const int32_t* __restrict__ edges0,
                const int32_t* __restrict__ edges1, int n_edges0, int n_edges1, double* __restrict__ part,
                int32_t* __restrict__ seg0) {
  const int b = blockIdx.y;
  const int c = blockIdx.x * KW_THREADS + threadIdx.x;
  const int n = n_samples[b];
  const int sel = blk_sel[b];
  if (sel < 0) return;
  const int64_t first = (int64_t)c * chunk;
  if (first >= n) return;
  const bool last = first + chunk >= n;
  if (!ENERGY && last) return;  // nobody consumes the end state of the last chunk
  const int len = last ? (int)(n - first) : chunk;
  double s[4] = {0.0, 0.0, 0.0, 0.0};
  const float* x = wav + (int64_t)b * pitch + first;
  if (!ENERGY) {
    int i = 0;
    for (; i + 4 <= len; i += 4) {
      const float4 v = *reinterpret_cast<const float4*>(x + i);
      kw_step(coef, (double)v.x, s);
      kw_step(coef, (double)v.y, s);
      kw_step(coef, (double)v.z, s);
      kw_step(coef, (double)v.w, s);
    }
    for (; i < len; ++i) kw_step(coef, (double)x[i], s);
    double* se = s_end + ((int64_t)b * chunks_pitch + c) * 4;
#pragma unroll
    for (int k = 0; k < 4; ++k) se[k] = s[k];
    return;
  }
  const double* si = s_in + ((int64_t)b * chunks_pitch + c) * 4;
#pragma unroll
  for (int i = 0; i < 4; ++i) s[i] = si[i];
  const int32_t* E = sel ? edges1 : edges0;
  const int ne = sel ? n_edges1 : n_edges0;
  int seg = 0;  // largest index with E[seg] <= first (E[0] == 0)
  for (int lo = 0, hi = ne; hi - lo > 1;) {
    const int mid = (lo + hi) >> 1;
    if ((int64_t)E[mid] <= first) lo = mid; else hi = mid;
    seg = lo;
  }
  seg0[(int64_t)b * chunks_pitch + c] = seg;
  double* pc = part + ((int64_t)b * chunks_pitch + c) * KW_MAX_PIECES;
  int piece = 0;
  int64_t next_edge = seg + 1 < ne ? (int64_t)E[seg + 1] : INT64_MAX;
  double acc = 0.0;
  auto sample = [&](float xv, int64_t idx) {
    while (idx >= next_edge) {
      if (piece < KW_MAX_PIECES) pc[piece] = acc;
      acc = 0.0;
      ++piece;
      ++seg;
      next_edge = seg + 1 < ne ? (int64_t)E[seg + 1] : INT64_MAX;
    }
    const float yf = (float)kw_step(coef, (double)xv, s);
    acc += (double)__fmul_rn(yf, yf);  // np.square of the float32 array, summed in fp64
  };
  int i = 0;
  for (; i + 4 <= len; i += 4) {
    const float4 v = *reinterpret_cast<const float4*>(x + i);
    sample(v.x, first + i);
    sample(v.y, first + i + 1);
    sample(v.z, first + i + 2);
    sample(v.w, first + i + 3);
  }
  for (; i < len; ++i) sample(x[i], first + i);
  if (piece < KW_MAX_PIECES) pc[piece] = acc;
}

// seg_sum[b][s] = sum of the pieces of segment s = [E[s], E[s+1]) (clipped to the utterance), chunk by chunk.
__global__ void __launch_bounds__(128)
kw_segment_sum_kernel(const int32_t* __restrict__ n_samples, const int32_t* __restrict__ blk_sel, int chunk,
                      int chunks_pitch, const int32_t* __restrict__ edges0, const int32_t* __restrict__ edges1,
                      int n_edges0, int n_edges1, const double* __restrict__ part, const int32_t* __restrict__ seg0,
                      int seg_pitch, double* __restrict__ seg_sum) {
  const int b = blockIdx.y, sg = blockIdx.x * 128 + threadIdx.x;
  const int sel = blk_sel[b];
  if (sel < 0) return;
  const int32_t* E = sel ? edges1 : edges0;
  const int ne = sel ? n_edges1 : n_edges0;
  if (sg + 1 >= ne) return;
  const int n = n_samples[b];
  const int lo = min(E[sg], n), hi = min(E[sg + 1], n);
  double t = 0.0;
  if (lo < hi) {
    for (int c = lo / chunk; c <= (hi - 1) / chunk; ++c) {
      const int p = sg - seg0[(int64_t)b * chunks_pitch + c];
      if (p >= 0 && p < KW_MAX_PIECES) t += part[((int64_t)b * chunks_pitch + c) * KW_MAX_PIECES + p];
    }
  }
  seg_sum[(int64_t)b * seg_pitch + sg] = t;
}

__device__ __forceinline__ double shfl_f64(double v, int src) {
  int lo = __double2loint(v), hi = __double2hiint(v);
  lo = __shfl_sync(0xffffffffu, lo, src);
  hi = __shfl_sync(0xffffffffu, hi, src);
  return __hiloint2double(hi, lo);
}

struct KwPow {
  double m[6][16];  // M^1, M^2, M^4, M^8, M^16, M^32 with M the transition over one chunk (row-major 4x4)
};

__device__ __forceinline__ void kw_matvec_acc(const double* __restrict__ m, const double (&v)[4], double (&acc)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    double t = acc[i];
#pragma unroll
    for (int j = 0; j < 4; ++j) t = fma(m[i * 4 + j], v[j], t);
    acc[i] = t;
  }
}

// One warp per utterance: s_in[c + 1] = M s_in[c] + s_end[c].  32 chunks at a time, one per lane, as an inclusive
// Hillis-Steele scan of the affine recurrence (step k adds M^(2^k) times the value 2^k lanes below); the state entering
// the group rides in lane 0's element, the last lane's result enters the next group.
__global__ void __launch_bounds__(32)
kw_chunk_scan_kernel(const int32_t* __restrict__ n_samples, const int32_t* __restrict__ blk_sel, int chunk,
                     int chunks_pitch, KwPow P, const double* __restrict__ s_end, double* __restrict__ s_in) {
  const int b = blockIdx.x, lane = threadIdx.x;
  if (blk_sel[b] < 0) return;
  const int n = n_samples[b];
  const int n_chunks = (n + chunk - 1) / chunk;
  double sb[4] = {0.0, 0.0, 0.0, 0.0};  // state entering chunk `base`
  auto load_end = [&](int c, double (&v)[4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = 0.0;
    if (c < n_chunks - 1) {
      const double2* p = reinterpret_cast<const double2*>(s_end + ((int64_t)b * chunks_pitch + c) * 4);
      const double2 lo = p[0], hi = p[1];
      v[0] = lo.x; v[1] = lo.y; v[2] = hi.x; v[3] = hi.y;
    }
  };
  double nxt[4];
  load_end(lane, nxt);
  for (int base = 0; base < n_chunks; base += 32) {
    const int c = base + lane;
    double e[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) e[i] = nxt[i];
    load_end(c + 32, nxt);  // the next group's end states travel while this group is scanned
    if (lane == 0) {
      double* q = s_in + ((int64_t)b * chunks_pitch + base) * 4;
#pragma unroll
      for (int i = 0; i < 4; ++i) q[i] = sb[i];
      kw_matvec_acc(P.m[0], sb, e);
    }
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      const int d = 1 << k;
      double t[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) t[i] = shfl_f64(e[i], lane >= d ? lane - d : lane);
      if (lane >= d) kw_matvec_acc(P.m[k], t, e);
    }
    // e = state after chunk c = state entering chunk c + 1
    if (lane < 31 && c + 1 < n_chunks) {
      double* q = s_in + ((int64_t)b * chunks_pitch + c + 1) * 4;
#pragma unroll
      for (int i = 0; i < 4; ++i) q[i] = e[i];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) sb[i] = shfl_f64(e[i], 31);
  }
}

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += shfl_f64(v, (threadIdx.x & 31) ^ o);
  return v;
}

constexpr int LT_THREADS = 1024;
constexpr int TRIM_FRAME = 512;      // autoencoder.py:53
constexpr int TRIM_MAX_FRAMES = 16;  // autoencoder.py:67

__global__ void __launch_bounds__(LT_THREADS)
loudness_trim_kernel(const float* __restrict__ wav, int64_t pitch, const int32_t* __restrict__ n_samples,
                     const int32_t* __restrict__ blk_sel, const int32_t* __restrict__ n_blocks, int max_blocks,
                     const int32_t* __restrict__ e_lo0, const int32_t* __restrict__ e_hi0,
                     const int32_t* __restrict__ e_lo1, const int32_t* __restrict__ e_hi1, double inv0, double inv1,
                     const double* __restrict__ seg_sum, int seg_pitch, double* __restrict__ z, double target_lufs,
                     float threshold, int32_t* __restrict__ bounds,
                     double* __restrict__ lufs_out, float* __restrict__ gain_out) {
  const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __shared__ float gain_s;
  __shared__ int over[2 * TRIM_MAX_FRAMES];
  const int n = n_samples[b];
  if (warp == 0) {
    double lufs = nan("");
    float gain = 1.0f;
    if (blk_sel[b] >= 0) {
      const int nb = n_blocks[b];
      double* zb_ = z + (int64_t)b * max_blocks;
      {
        // z_j = (1 / (T_g * rate)) * sum(y[lo_j : hi_j]^2)  (pyloudnorm meter.py, eq. 1): block j is the run of
        // segments e_lo[j] .. e_hi[j] - 1
        const int sel = blk_sel[b];
        const int32_t* e_lo = sel ? e_lo1 : e_lo0;
        const int32_t* e_hi = sel ? e_hi1 : e_hi0;
        const double inv = sel ? inv1 : inv0;
        const double* ss = seg_sum + (int64_t)b * seg_pitch;
        for (int j = lane; j < nb; j += 32) {
          double t = 0.0;
          for (int sg = e_lo[j]; sg < e_hi[j]; ++sg) t += ss[sg];
          zb_[j] = inv * t;  // read back below by the same lane
        }
      }
      // absolute gate (Gamma_a = -70 LKFS), then the mean of the blocks that pass it
      double sum1 = 0.0, cnt1 = 0.0;
      for (int j = lane; j < nb; j += 32) {
        const double l = -0.691 + 10.0 * log10(zb_[j]);
        if (l >= -70.0) { sum1 += zb_[j]; cnt1 += 1.0; }
      }
      sum1 = warp_sum_f64(sum1);
      cnt1 = warp_sum_f64(cnt1);
      const double gamma_r = -0.691 + 10.0 * log10(sum1 / cnt1) - 10.0;  // 0/0 -> NaN -> every comparison false
      double sum2 = 0.0, cnt2 = 0.0;
      for (int j = lane; j < nb; j += 32) {
        const double l = -0.691 + 10.0 * log10(zb_[j]);
        if (l > gamma_r && l > -70.0) { sum2 += zb_[j]; cnt2 += 1.0; }
      }
      sum2 = warp_sum_f64(sum2);
      cnt2 = warp_sum_f64(cnt2);
      const double zavg = cnt2 > 0.0 ? sum2 / cnt2 : 0.0;  // np.nan_to_num(mean([]))
      lufs = -0.691 + 10.0 * log10(zavg);
      gain = (float)pow(10.0, (target_lufs - lufs) / 20.0);
    }
    if (lane == 0) {
      gain_s = gain;
      lufs_out[b] = lufs;
      gain_out[b] = gain;
    }
  }
  __syncthreads();
  const float gain = gain_s;
  const int nf = min((n / TRIM_FRAME) / 4, TRIM_MAX_FRAMES);
  {
    // warps 0..15: leading frame i; warps 16..31: trailing frame i (its slice is empty for i == 0: wav[:, -512:-0])
    const bool tail = warp >= TRIM_MAX_FRAMES;
    const int i = tail ? warp - TRIM_MAX_FRAMES : warp;
    int flag = 0;
    if (i < nf && !(tail && i == 0)) {
      const int first = tail ? n - (i + 1) * TRIM_FRAME : i * TRIM_FRAME;
      const float* p = wav + (int64_t)b * pitch + first;
      float acc = 0.0f;
#pragma unroll
      for (int k = 0; k < TRIM_FRAME / 32; ++k) {
        const float v = __fmul_rn(p[k * 32 + lane], gain);  // no fma contraction: torch squares, then sums
        acc += __fmul_rn(v, v);
      }
      acc = warp_sum(acc);
      flag = (acc / (float)TRIM_FRAME) > threshold;
    }
    if (lane == 0) over[warp] = flag;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int start = 0, end = n;
    for (int i = 0; i < nf; ++i)
      if (over[i]) { start = i * TRIM_FRAME; break; }
    for (int i = 1; i < nf; ++i)
      if (over[TRIM_MAX_FRAMES + i]) { end = n - (i + 1) * TRIM_FRAME; break; }  // cuts the detected frame too
    if (end < start) end = start;
    bounds[2 * b] = start;
    bounds[2 * b + 1] = end;
  }
}

constexpr int AP_THREADS = 256;
constexpr int AP_VEC_PER_THREAD = 4;   // float4 each
constexpr int FADE_BLOCK = 512;      // autoencoder.py:234
constexpr int FADE_MAX_BLOCKS = 20;  // autoencoder.py:239

// out[b][i] = ((x[b][start + i] * gain) * fade_in[i]) * fade_out[i - (n_out - fade_len)], zero from n_out to the pitch.
// start is a multiple of 512 and the pitch a multiple of 4, so the float4 path is always aligned.
__global__ void __launch_bounds__(AP_THREADS)
postproc_apply_kernel(const float* __restrict__ wav, int64_t pitch, const int32_t* __restrict__ bounds,
                      const float* __restrict__ gain_in, const float* __restrict__ fade_in,
                      const float* __restrict__ fade_out, int no_fades, float* __restrict__ out) {
  const int b = blockIdx.y;
  const int start = bounds[2 * b], n_out = bounds[2 * b + 1] - start;
  const float gain = gain_in[b];
  const int fo_blocks = no_fades ? 0 : min((n_out / FADE_BLOCK) / 4, FADE_MAX_BLOCKS);
  const int fo_first = n_out - fo_blocks * FADE_BLOCK;
  const float* fo = fade_out + (int64_t)FADE_BLOCK * (fo_blocks * (fo_blocks - 1) / 2);  // tables of 512*k, k = 1..20
  const int fi_end = no_fades ? 0 : FADE_BLOCK;
  const float* x = wav + (int64_t)b * pitch + start;
  float* o = out + (int64_t)b * pitch;
  const int64_t base = (int64_t)blockIdx.x * (AP_THREADS * AP_VEC_PER_THREAD * 4);
#pragma unroll
  for (int k = 0; k < AP_VEC_PER_THREAD; ++k) {
    const int64_t i = base + ((int64_t)k * AP_THREADS + threadIdx.x) * 4;
    if (i >= pitch) break;
    float v[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    if (i + 3 < n_out && start + i + 3 < pitch) {
      const float4 t = *reinterpret_cast<const float4*>(x + i);
      v[0] = t.x * gain; v[1] = t.y * gain; v[2] = t.z * gain; v[3] = t.w * gain;
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (i + e < n_out) v[e] = x[i + e] * gain;
    }
    if (i < fi_end || (fo_blocks > 0 && i + 3 >= fo_first)) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int64_t ie = i + e;
        if (ie < n_out) {
          if (ie < fi_end) v[e] = v[e] * fade_in[ie];
          if (fo_blocks > 0 && ie >= fo_first) v[e] = v[e] * fo[ie - fo_first];
        }
      }
    }
    *reinterpret_cast<float4*>(o + i) = make_float4(v[0], v[1], v[2], v[3]);
  }
}

struct WsLayout {
  int chunks_pitch, seg_pitch;
  int64_t s_end, s_in, part, seg_sum, z, seg0, total;
};
WsLayout ws_layout(int batch, int max_n, int chunk, int max_blocks) {
  WsLayout w;
  w.chunks_pitch = (max_n + chunk - 1) / chunk + 1;
  w.seg_pitch = 2 * max_blocks + 2;  // at most one lower and one upper edge per block
  const int64_t chunks = (int64_t)batch * w.chunks_pitch, d = (int64_t)sizeof(double);
  w.s_end = 0;
  w.s_in = w.s_end + chunks * 4 * d;
  w.part = w.s_in + chunks * 4 * d;
  w.seg_sum = w.part + chunks * KW_MAX_PIECES * d;
  w.z = w.seg_sum + (int64_t)batch * w.seg_pitch * d;
  w.seg0 = w.z + (int64_t)batch * (max_blocks > 0 ? max_blocks : 1) * d;
  w.total = w.seg0 + chunks * (int64_t)sizeof(int32_t);
  return w;
}

}  // namespace
}  // namespace zb

using namespace zb;

extern "C" int64_t zb200_postproc_ws_bytes(int batch, int max_n, int chunk, int max_blocks) {
  if (batch <= 0 || max_n < 0 || chunk <= 0 || max_blocks < 0) return 0;
  return ws_layout(batch, max_n, chunk, max_blocks).total;
}

extern "C" int zb200_codes_postprocess(const zb200_postproc_plan* plan, const float* wav, int64_t pitch,
                                       const int32_t* n_samples, const int32_t* blk_sel, const int32_t* n_blocks,
                                       int batch, int max_n, float* out, int32_t* bounds, double* lufs, float* gain,
                                       void* ws, int64_t ws_bytes, void* stream) {
  ZB_REQUIRE(plan && wav && n_samples && blk_sel && n_blocks && out && bounds && lufs && gain && ws,
             "codes_postprocess: null pointer");
  ZB_REQUIRE(batch > 0 && max_n > 0 && pitch >= max_n, "codes_postprocess: bad batch / max_n / pitch");
  ZB_REQUIRE(pitch % 4 == 0 && (reinterpret_cast<uintptr_t>(wav) & 15) == 0 &&
                 (reinterpret_cast<uintptr_t>(out) & 15) == 0,
             "codes_postprocess: wav / out must be 16-byte aligned with a pitch that is a multiple of 4 samples");
  ZB_REQUIRE(wav != out, "codes_postprocess: out must not alias wav");
  ZB_REQUIRE(plan->chunk >= 4 && plan->chunk % 4 == 0, "codes_postprocess: chunk must be a positive multiple of 4");
  ZB_REQUIRE(plan->max_blocks > 0 && plan->blk_lo[0] && plan->blk_hi[0] && plan->blk_lo[1] && plan->blk_hi[1] &&
                 plan->edges[0] && plan->edges[1],
             "codes_postprocess: gating-block tables missing");
  for (int k = 0; k < 2; ++k)
    ZB_REQUIRE(plan->n_edges[k] >= 2 && plan->n_edges[k] <= 2 * plan->max_blocks + 2,
               "codes_postprocess: n_edges[%d] = %d out of range", k, plan->n_edges[k]);
  ZB_REQUIRE(plan->fade_in && plan->fade_out, "codes_postprocess: fade tables missing");
  const WsLayout w = ws_layout(batch, max_n, plan->chunk, plan->max_blocks);
  ZB_REQUIRE(ws_bytes >= w.total, "codes_postprocess: workspace too small (%lld < %lld bytes)", (long long)ws_bytes,
             (long long)w.total);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  KwCoef c;
  c.b0a = plan->shelf_b[0]; c.b1a = plan->shelf_b[1]; c.b2a = plan->shelf_b[2];
  c.a1a = plan->shelf_a[1]; c.a2a = plan->shelf_a[2];
  c.b0b = plan->hp_b[0]; c.b1b = plan->hp_b[1]; c.b2b = plan->hp_b[2];
  c.a1b = plan->hp_a[1]; c.a2b = plan->hp_a[2];
  KwPow P;
  for (int i = 0; i < 16; ++i) P.m[0][i] = plan->chunk_transition[i];
  for (int k = 1; k < 6; ++k)  // squares: M^(2^k)
    for (int i = 0; i < 4; ++i)
      for (int j = 0; j < 4; ++j) {
        double t = 0.0;
        for (int l = 0; l < 4; ++l) t += P.m[k - 1][i * 4 + l] * P.m[k - 1][l * 4 + j];
        P.m[k][i * 4 + j] = t;
      }
  char* wsb = static_cast<char*>(ws);
  double* s_end = reinterpret_cast<double*>(wsb + w.s_end);
  double* s_in = reinterpret_cast<double*>(wsb + w.s_in);
  double* part = reinterpret_cast<double*>(wsb + w.part);
  double* seg_sum = reinterpret_cast<double*>(wsb + w.seg_sum);
  double* z = reinterpret_cast<double*>(wsb + w.z);
  int32_t* seg0 = reinterpret_cast<int32_t*>(wsb + w.seg0);
  const int n_chunks = ceil_div(max_n, plan->chunk);
  const dim3 kw_grid(ceil_div(n_chunks, KW_THREADS), batch);
  kw_chunk_kernel<false><<<kw_grid, KW_THREADS, 0, st>>>(wav, pitch, n_samples, blk_sel, plan->chunk, w.chunks_pitch,
                                                         c, s_end, nullptr, nullptr, nullptr, 0, 0, nullptr, nullptr);
  ZB_CHECK_LAUNCH("kw_chunk_kernel<state>");
  kw_chunk_scan_kernel<<<batch, 32, 0, st>>>(n_samples, blk_sel, plan->chunk, w.chunks_pitch, P, s_end, s_in);
  ZB_CHECK_LAUNCH("kw_chunk_scan_kernel");
  kw_chunk_kernel<true><<<kw_grid, KW_THREADS, 0, st>>>(wav, pitch, n_samples, blk_sel, plan->chunk, w.chunks_pitch, c,
                                                        nullptr, s_in, plan->edges[0], plan->edges[1],
                                                        plan->n_edges[0], plan->n_edges[1], part, seg0);
  ZB_CHECK_LAUNCH("kw_chunk_kernel<energy>");
  const int max_edges = plan->n_edges[0] > plan->n_edges[1] ? plan->n_edges[0] : plan->n_edges[1];
  kw_segment_sum_kernel<<<dim3(ceil_div(max_edges, 128), batch), 128, 0, st>>>(
      n_samples, blk_sel, plan->chunk, w.chunks_pitch, plan->edges[0], plan->edges[1], plan->n_edges[0],
      plan->n_edges[1], part, seg0, w.seg_pitch, seg_sum);
  ZB_CHECK_LAUNCH("kw_segment_sum_kernel");
  loudness_trim_kernel<<<batch, LT_THREADS, 0, st>>>(
      wav, pitch, n_samples, blk_sel, n_blocks, plan->max_blocks, plan->blk_lo[0], plan->blk_hi[0], plan->blk_lo[1],
      plan->blk_hi[1], plan->inv_norm[0], plan->inv_norm[1], seg_sum, w.seg_pitch, z, plan->target_lufs,
      plan->trim_threshold, bounds, lufs, gain);
  ZB_CHECK_LAUNCH("loudness_trim_kernel");
  const int ap_tile = AP_THREADS * AP_VEC_PER_THREAD * 4;
  const dim3 ap_grid((unsigned)((pitch + ap_tile - 1) / ap_tile), batch);
  postproc_apply_kernel<<<ap_grid, AP_THREADS, 0, st>>>(wav, pitch, bounds, gain, plan->fade_in, plan->fade_out,
                                                        (plan->flags & 1) != 0, out);
  ZB_CHECK_LAUNCH("postproc_apply_kernel");
  return ZB200_OK;
}
