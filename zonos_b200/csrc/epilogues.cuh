// Fused epilogues of the tcgen05 GEMM (gemm_tc.cuh).  Each functor is called by the 128 epilogue threads with
//   (tile_row, row_valid, m_tile, n_tile, acc[NT])   where acc is the fp32 accumulator row of that thread.
// Rounding mirrors the eager reference: every nn.Linear output is rounded to bf16 before the next op
// (SURVEY.md Appendix B.10), so acc is first passed through bf16_round().
#pragma once
#include "common.cuh"
#include "gemm_tc.cuh"
#include "kv.cuh"

namespace zb {

// Device-side step state shared by all kernels of a decode step (one int32 array per session).
enum StepStateIdx {
  ST_OFFSET = 0,        // column of the delayed-code matrix the NEXT frame is written to (model.py `offset` + 1 ...)
  ST_POS = 1,           // tokens already in the KV cache == position of the token being decoded
  ST_DONE = 2,          // 1 once max(remaining_steps) <= 0
  ST_FINAL_OFFSET = 3,  // value of the reference's `offset` when its while-loop exits
  ST_DRAW = 4,          // number of noise draws consumed so far
  ST_ANY_FRESH = 5,     // this step: some row produced its first EOS -> whole batch is re-drawn (model.py:380-393)
  ST_STEP = 6,          // loop iterations executed
  ST_TICKET_A = 7,      // last-CTA tickets of the two sampler passes
  ST_TICKET_B = 8,
  ST_ERROR = 9,         // sticky device-side error flags
  ST_COUNT = 16
};

template <int NT>
struct EpiStoreBf16 {
  bf16* out;
  int ldo;      // elements
  int n_valid;  // columns that exist
  __device__ void operator()(int row, bool valid, int m_tile, int n_tile, float* acc) const {
    if (!valid) return;
    bf16* o = out + (size_t)(m_tile * GEMM_BM + row) * ldo + n_tile * NT;
    const int n0 = n_tile * NT;
    if (n0 + NT <= n_valid && (ldo % 16) == 0 && (NT % 16) == 0 && (reinterpret_cast<uintptr_t>(out) & 31) == 0) {
#pragma unroll
      for (int c = 0; c < NT; c += 16) {  // lane == row: 32 bytes per store instruction (256-bit STG)
        uint4 v, u;
        v.x = pack_bf16x2(acc[c], acc[c + 1]), v.y = pack_bf16x2(acc[c + 2], acc[c + 3]);
        v.z = pack_bf16x2(acc[c + 4], acc[c + 5]), v.w = pack_bf16x2(acc[c + 6], acc[c + 7]);
        u.x = pack_bf16x2(acc[c + 8], acc[c + 9]), u.y = pack_bf16x2(acc[c + 10], acc[c + 11]);
        u.z = pack_bf16x2(acc[c + 12], acc[c + 13]), u.w = pack_bf16x2(acc[c + 14], acc[c + 15]);
        stg256(o + c, v, u);
      }
    } else if (n0 + NT <= n_valid && (ldo % 8) == 0) {
#pragma unroll
      for (int c = 0; c < NT; c += 8) {
        uint4 v;
        v.x = pack_bf16x2(acc[c], acc[c + 1]), v.y = pack_bf16x2(acc[c + 2], acc[c + 3]);
        v.z = pack_bf16x2(acc[c + 4], acc[c + 5]), v.w = pack_bf16x2(acc[c + 6], acc[c + 7]);
        *reinterpret_cast<uint4*>(o + c) = v;
      }
    } else {
      for (int c = 0; c < NT; ++c)
        if (n0 + c < n_valid) o[c] = __float2bfloat16_rn(acc[c]);
    }
  }
};

// Deferred split-K (gemm_skinny.cuh): the raw fp32 partial sums of one K split go to ws[split][row][n]; the consumer
// (resid_ln_kernel, ops.cu) adds the splits in order, rounds, adds the residual and normalises.
struct EpiPartial {
  static constexpr bool kDeferred = true;
  float* ws;  // [splits][rows][ldn]
  int rows, ldn;
  __device__ void partial(int row, bool valid, int n_tile, int split, const float* acc) const {
    if (!valid) return;
    float* o = ws + ((size_t)split * rows + row) * ldn + n_tile * 64;
    const uint32_t* a = reinterpret_cast<const uint32_t*>(acc);
#pragma unroll
    for (int c = 0; c < 64; c += 8)  // ws is 256-byte aligned and ldn a multiple of 64: 32-byte pieces
      stg256(o + c, make_uint4(a[c], a[c + 1], a[c + 2], a[c + 3]), make_uint4(a[c + 4], a[c + 5], a[c + 6], a[c + 7]));
  }
  __device__ void operator()(int, bool, int, int, float*) const {}
  // Warp-collective variant at 128 rows: lane == row would store 16 bytes to 32 different 128-byte lines per
  // instruction (32 LSU wavefronts, ~2 us per 64 KB tile pair with the 4 epilogue warps sharing one LSU).  The warp
  // transposes its 32 x 64 tile through `stage` (8 KB of idle ring space, 16-byte chunks XOR-swizzled by the row) and
  // stores two whole 256-byte row segments per instruction.
  __device__ void partial_warp(float* stage, int lane, int row0, int n_tile, int split, const float* acc) const {
    float4* st = reinterpret_cast<float4*>(stage);
#pragma unroll
    for (int c = 0; c < 16; ++c) st[lane * 16 + (c ^ (lane & 15))] = make_float4(acc[4 * c], acc[4 * c + 1], acc[4 * c + 2], acc[4 * c + 3]);
    __syncwarp();
    const int half = lane >> 4, c = lane & 15;
#pragma unroll
    for (int r = 0; r < 32; r += 2) {
      const int rr = r + half;  // row of the tile this lane moves a chunk of
      const float4 v = st[rr * 16 + (c ^ (rr & 15))];
      if (row0 + rr < rows)
        reinterpret_cast<float4*>(ws + ((size_t)split * rows + row0 + rr) * ldn + n_tile * 64)[c] = v;
    }
    __syncwarp();
  }
};

// x[row, n] = bf16( x[row, n] + bf16(acc) )      (out_proj / fc2 followed by the residual add, _torch.py:100-101)
template <int NT>
struct EpiResidual {
  bf16* x;
  int ldx;
  __device__ void operator()(int row, bool valid, int m_tile, int n_tile, float* acc) const {
    if (!valid) return;
    bf16* o = x + (size_t)(m_tile * GEMM_BM + row) * ldx + n_tile * NT;
    auto add8 = [&](uint4& r, int c) {
      const float2 a = unpack_bf16x2(r.x), b = unpack_bf16x2(r.y), cc = unpack_bf16x2(r.z), d = unpack_bf16x2(r.w);
      r.x = pack_bf16x2(a.x + bf16_round(acc[c]), a.y + bf16_round(acc[c + 1]));
      r.y = pack_bf16x2(b.x + bf16_round(acc[c + 2]), b.y + bf16_round(acc[c + 3]));
      r.z = pack_bf16x2(cc.x + bf16_round(acc[c + 4]), cc.y + bf16_round(acc[c + 5]));
      r.w = pack_bf16x2(d.x + bf16_round(acc[c + 6]), d.y + bf16_round(acc[c + 7]));
    };
    if ((NT % 16) == 0 && (ldx % 16) == 0 && (reinterpret_cast<uintptr_t>(x) & 31) == 0) {
#pragma unroll
      for (int c = 0; c < NT; c += 16) {  // lane == row: 32 bytes per instruction (256-bit LDG / STG)
        uint4 r, q;
        ldg256(o + c, r, q);
        add8(r, c), add8(q, c + 8);
        stg256(o + c, r, q);
      }
    } else {
#pragma unroll
      for (int c = 0; c < NT; c += 8) {
        uint4 r = *reinterpret_cast<const uint4*>(o + c);
        add8(r, c);
        *reinterpret_cast<uint4*>(o + c) = r;
      }
    }
  }
};

// fc1 with the pairing done at pack time (row_mode 1): columns [0,NT/2) = value, [NT/2,NT) = gate of the same
// hidden units.  h = bf16( bf16(value) * bf16(silu(bf16(gate))) )     (_torch.py:151-152)
template <int NT>
struct EpiSwiGLU {
  bf16* h;
  int ldh;
  __device__ void operator()(int row, bool valid, int m_tile, int n_tile, float* acc) const {
    if (!valid) return;
    constexpr int HN = NT / 2;
    bf16* o = h + (size_t)(m_tile * GEMM_BM + row) * ldh + n_tile * HN;
    float r[HN];
#pragma unroll
    for (int c = 0; c < HN; ++c) {
      const float val = bf16_round(acc[c]);
      const float gate = bf16_round(acc[HN + c]);
      const float s = bf16_round(gate / (1.0f + expf(-gate)));
      r[c] = val * s;
    }
    // lane == row: 32 bytes per store instruction and lane (256-bit STG) when the rows allow it
    const bool wide = (HN % 16 == 0) && ((ldh & 15) == 0) && ((reinterpret_cast<uintptr_t>(h) & 31) == 0);
#pragma unroll
    for (int c = 0; c < HN; c += 16) {
      uint4 v[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int cc = c + 8 * k;
        if (cc < HN) {
          v[k].x = pack_bf16x2(r[cc], r[cc + 1]), v[k].y = pack_bf16x2(r[cc + 2], r[cc + 3]);
          v[k].z = pack_bf16x2(r[cc + 4], r[cc + 5]), v[k].w = pack_bf16x2(r[cc + 6], r[cc + 7]);
        }
      }
      if (wide) {
        stg256(o + c, v[0], v[1]);
      } else {
        *reinterpret_cast<uint4*>(o + c) = v[0];
        if (c + 8 < HN) *reinterpret_cast<uint4*>(o + c + 8) = v[1];
      }
    }
  }
};

// in_proj epilogue (NT = 64 or 16 columns of one head): interleaved-pair RoPE in fp32 on q and k (separately rounded
// products, result cast to bf16: _torch.py:18-30), q -> q_out [rows, H*128], k/v -> paged cache (replaces
// _update_kv_cache, _torch.py:33-49).  Row m is (sequence m / seq_len, position pos0 + m % seq_len).
template <int NT>
struct EpiQkvRopeT {
  static_assert(128 % NT == 0 && NT % 8 == 0, "a tile must not straddle heads");
  bf16* q_out;
  KvLayout kv;
  const float* rope;   // fp32 [n_pos][64][2] (cos, sin)
  const int* pos_ptr;  // device: base position (decode) or nullptr
  int pos_add, seq_len, n_heads;
  __device__ void operator()(int row, bool valid, int m_tile, int n_tile, float* acc) const {
    if (!valid) return;
    const int grow = m_tile * GEMM_BM + row;
    const int seq = grow / seq_len;
    const int pos = (pos_ptr ? *pos_ptr : 0) + pos_add + grow % seq_len;
    const int col0 = n_tile * NT;
    const int head_all = col0 >> 7, within = col0 & 127;  // 128 dims per head
    const int q_heads = n_heads, kv_heads = kv.n_kv_heads;
    float r[NT];
    if (head_all < q_heads + kv_heads) {  // q or k: rotate
      const float2* cs = reinterpret_cast<const float2*>(rope) + (size_t)pos * 64 + within / 2;
#pragma unroll
      for (int i = 0; i < NT / 2; ++i) {
        const float2 t = __ldg(cs + i);
        const float a = bf16_round(acc[2 * i]), b = bf16_round(acc[2 * i + 1]);
        r[2 * i] = __fsub_rn(__fmul_rn(a, t.x), __fmul_rn(b, t.y));
        r[2 * i + 1] = __fadd_rn(__fmul_rn(b, t.x), __fmul_rn(a, t.y));
      }
    } else {
#pragma unroll
      for (int i = 0; i < NT; ++i) r[i] = acc[i];
    }
    uint4 v[NT / 8];
#pragma unroll
    for (int c = 0; c < NT / 8; ++c) {
      v[c].x = pack_bf16x2(r[8 * c], r[8 * c + 1]), v[c].y = pack_bf16x2(r[8 * c + 2], r[8 * c + 3]);
      v[c].z = pack_bf16x2(r[8 * c + 4], r[8 * c + 5]), v[c].w = pack_bf16x2(r[8 * c + 6], r[8 * c + 7]);
    }
    if (head_all < q_heads) {
      uint4* o = reinterpret_cast<uint4*>(q_out + (size_t)grow * q_heads * 128 + head_all * 128 + within);
#pragma unroll
      for (int c = 0; c < NT / 8; ++c) o[c] = v[c];
    } else {
      const int is_v = head_all >= q_heads + kv_heads;
      const int head = head_all - q_heads - (is_v ? kv_heads : 0);
      uint4* o = reinterpret_cast<uint4*>(kv.row(seq, pos, is_v, head));
#pragma unroll
      for (int c = 0; c < NT / 8; ++c) o[(within / 8 + c) ^ (pos & 7)] = v[c];
    }
  }
};
typedef EpiQkvRopeT<64> EpiQkvRope;

// 9 output heads + classifier-free guidance + pad mask (model.py:100-116).  A rows are loaded as (cond, uncond)
// pairs (ARowCfgPairs) so the partner logit sits in the neighbouring lane.
// logits[u, head, col] = uncond + (cond - uncond) * cfg  (fp32, unfused);  col >= 1025 -> -inf.
template <int NT>
struct EpiHeadsCfg {
  float* logits;  // [B, n_heads_out, vocab]
  int batch, n_heads_out, vocab, tiles_per_head;
  float cfg_scale;
  __device__ void operator()(int row, bool valid, int m_tile, int n_tile, float* acc) const {
    const int head = n_tile / tiles_per_head;
    const int col0 = (n_tile % tiles_per_head) * NT;
    const int u = m_tile * (GEMM_BM / 2) + (row >> 1);
    float* o = logits + ((size_t)u * n_heads_out + head) * vocab + col0;
    const bool writer = valid && !(row & 1) && u < batch;
#pragma unroll
    for (int c = 0; c < NT; ++c) {
      const float mine = bf16_round(acc[c]);
      const float other = __shfl_xor_sync(0xffffffffu, mine, 1);  // odd lane holds uncond
      if (writer && col0 + c < vocab) {
        const float v = __fadd_rn(other, __fmul_rn(__fsub_rn(mine, other), cfg_scale));
        o[c] = (col0 + c >= 1025) ? -INFINITY : v;
      }
    }
  }
};

}  // namespace zb
