#pragma once
#include "common.cuh"
#include "kv.cuh"

namespace zb {

struct AttnDecodeArgs {
  const bf16* q;  // [n_seq, n_heads*128], RoPE applied
  bf16* out;      // [n_seq, n_heads*128]
  KvLayout kv;
  const int* pos_ptr;  // device: position of the query token (keys visible = pos + 1), or nullptr
  int pos_add;
  int n_heads, n_kv_heads;
  float scale;
  float* partial_ws;  // [n_seq*n_kv_heads][nsplit][4*(128+2)] floats
  int* tickets;       // [n_seq*n_kv_heads], zero-initialised, self-resetting
};

struct AttnPrefillArgs {
  const bf16* q;  // [n_seq, seq_len, n_heads*128]
  bf16* out;
  KvLayout kv;
  int seq_len, pos0, n_heads;
  float scale;
};

cudaError_t launch_attn_decode(const AttnDecodeArgs& a, int n_seq, int nsplit, cudaStream_t st);
cudaError_t launch_attn_prefill(const AttnPrefillArgs& a, int n_seq, cudaStream_t st);

}  // namespace zb
