// Internal launchers shared between the C-ABI test hooks (ops.cu) and the session engine (engine.cu).
#pragma once
#include "common.cuh"

namespace zb {
cudaError_t launch_layernorm(const bf16* x, const bf16* w, const bf16* b, bf16* out, int rows, int d, float eps,
                             int in_row_mul, int in_row_add, cudaStream_t st);
cudaError_t launch_resid_ln(bf16* x, const float* ws, int splits, int rows, const bf16* w, const bf16* b, bf16* xn, int d,
                            float eps, cudaStream_t st, unsigned long long* trace = nullptr, int trace_id = -1);
struct KvLayout;
cudaError_t launch_qkv_finish(const float* ws, int splits, int rows, int n_cols, bf16* q_out, const KvLayout& kv,
                              const float* rope, const int* pos_ptr, int n_heads, cudaStream_t st);
cudaError_t launch_embed_sum(const int32_t* ids, int ids_stride_b, int ids_stride_q, const int* col_ptr, int col_add,
                             int n_cols, const bf16* emb, bf16* out, int batch, int n_q, int vocab, int d,
                             int out_stride_b, int out_off, int dup_off, cudaStream_t st);
// mamba2.cu: fused residual add + (RMS|Layer)Norm of the hybrid blocks; with `ws` it also reduces a deferred split-K GEMM
cudaError_t launch_add_norm(const bf16* x, float* residual, int has_residual, const bf16* w, const bf16* b, bf16* out, int rows,
                            int d, float eps, int rms, const float* ws, int splits, cudaStream_t st);
}  // namespace zb
