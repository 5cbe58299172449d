// Transposed fused epilogues of the M <= 8 weight-streaming GEMM (gemm_swap.cuh): the thread owns ONE output feature
// (packed tile `tile`, row `r` of it) and acc[b] is batch row b.  Same arithmetic and rounding points as the row-major
// epilogues in epilogues.cuh (every nn.Linear output is rounded to bf16 first: SURVEY.md Appendix B.10).
// All 128 epilogue threads call the functor (tile_valid == false for the padding half of an odd last group).
#pragma once
#include "common.cuh"
#include "epilogues.cuh"
#include "kv.cuh"

namespace zb {

struct EpiTStoreBf16 {  // test hook: out[b, n] = bf16(acc)
  bf16* out;
  int ldo, n_valid;
  __device__ void operator()(int, int tile, bool tile_valid, int r, int rows, const float* acc, float*) const {
    const int n = tile * 64 + r;
    if (!tile_valid || n >= n_valid) return;
#pragma unroll
    for (int b = 0; b < 8; ++b)
      if (b < rows) out[(size_t)b * ldo + n] = __float2bfloat16_rn(acc[b]);
  }
};

// x[b, n] = bf16( x[b, n] + bf16(acc) )      (out_proj / fc2 followed by the residual add, _torch.py:100-101)
struct EpiTResidual {
  bf16* x;
  int ldx;
  __device__ void operator()(int, int tile, bool tile_valid, int r, int rows, const float* acc, float*) const {
    if (!tile_valid) return;
    bf16* o = x + tile * 64 + r;
#pragma unroll
    for (int b = 0; b < 8; ++b)
      if (b < rows) o[(size_t)b * ldx] = __float2bfloat16_rn(__bfloat162float(o[(size_t)b * ldx]) + bf16_round(acc[b]));
  }
};

// fc1 packed with row_mode 1: rows [0,32) of a tile are the values of hidden units tile*32 + r, rows [32,64) their
// gates.  h = bf16( bf16(value) * bf16(silu(bf16(gate))) )   (_torch.py:151-152).  The gate sits 32 lanes away: it is
// handed over through the scratch buffer.
struct EpiTSwiGLU {
  bf16* h;
  int ldh;
  __device__ void operator()(int lane_row, int tile, bool tile_valid, int r, int rows, const float* acc, float* scratch) const {
    if (r >= 32) {
#pragma unroll
      for (int b = 0; b < 8; ++b) {
        const float gate = bf16_round(acc[b]);
        scratch[lane_row * 8 + b] = bf16_round(gate / (1.0f + expf(-gate)));
      }
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (r < 32 && tile_valid) {
      bf16* o = h + tile * 32 + r;
#pragma unroll
      for (int b = 0; b < 8; ++b)
        if (b < rows) o[(size_t)b * ldh] = __float2bfloat16_rn(bf16_round(acc[b]) * scratch[(lane_row + 32) * 8 + b]);
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");  // the scratch is reused by the CTA's next fragment
  }
};

// in_proj: interleaved-pair RoPE in fp32 on q and k (_torch.py:18-30; the partner of a pair is the neighbouring
// lane), q -> q_out [rows, H*128], k/v -> paged cache (replaces _update_kv_cache, _torch.py:33-49).  Decode only:
// batch row b is sequence b at position *pos_ptr.
struct EpiTQkvRope {
  bf16* q_out;
  KvLayout kv;
  const float* rope;   // fp32 [n_pos][64][2] (cos, sin)
  const int* pos_ptr;
  int n_heads;
  __device__ void operator()(int, int tile, bool tile_valid, int r, int rows, const float* acc, float*) const {
    const int n = tile * 64 + r;
    const int head_all = n >> 7, within = n & 127;
    const int q_heads = n_heads, kv_heads = kv.n_kv_heads;
    const bool rotate = head_all < q_heads + kv_heads;
    const int pos = *pos_ptr;
    float2 cs = make_float2(1.f, 0.f);
    if (rotate && tile_valid) cs = __ldg(reinterpret_cast<const float2*>(rope) + (size_t)pos * 64 + (within >> 1));
    float res[8];
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const float mine = bf16_round(acc[b]);
      const float other = __shfl_xor_sync(0xffffffffu, mine, 1);
      // even feature: a*cos - b*sin with (a, b) = (mine, other); odd feature: b*cos + a*sin with (a, b) = (other, mine)
      const float rot = (n & 1) ? __fadd_rn(__fmul_rn(mine, cs.x), __fmul_rn(other, cs.y))
                                : __fsub_rn(__fmul_rn(mine, cs.x), __fmul_rn(other, cs.y));
      res[b] = rotate ? rot : acc[b];
    }
    if (!tile_valid) return;
    if (head_all < q_heads) {
      bf16* o = q_out + head_all * 128 + within;
#pragma unroll
      for (int b = 0; b < 8; ++b)
        if (b < rows) o[(size_t)b * q_heads * 128] = __float2bfloat16_rn(res[b]);
    } else {
      const int is_v = head_all >= q_heads + kv_heads;
      const int head = head_all - q_heads - (is_v ? kv_heads : 0);
      const int elem = ((((within >> 3) ^ (pos & 7))) << 3) + (within & 7);  // 16-byte chunks are XOR-swizzled (kv.cuh)
#pragma unroll
      for (int b = 0; b < 8; ++b)
        if (b < rows) kv.row(b, pos, is_v, head)[elem] = __float2bfloat16_rn(res[b]);
    }
  }
};

// 9 output heads + classifier-free guidance + pad mask (model.py:100-116).  Batch rows [0,B) are conditional, [B,2B)
// unconditional: both logits of an utterance are in this thread.
struct EpiTHeadsCfg {
  float* logits;  // [B, n_heads_out, vocab]
  int batch, n_heads_out, vocab, tiles_per_head;
  float cfg_scale;
  __device__ void operator()(int, int tile, bool tile_valid, int r, int, const float* acc, float*) const {
    if (!tile_valid) return;
    const int head = tile / tiles_per_head;
    const int col = (tile - head * tiles_per_head) * 64 + r;
    if (col >= vocab) return;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (u < batch) {
        float unc_raw = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j == batch + u) unc_raw = acc[j];
        const float cond = bf16_round(acc[u]), unc = bf16_round(unc_raw);
        const float v = __fadd_rn(unc, __fmul_rn(__fsub_rn(cond, unc), cfg_scale));
        logits[((size_t)u * n_heads_out + head) * vocab + col] = (col >= 1025) ? -INFINITY : v;
      }
  }
};

}  // namespace zb
