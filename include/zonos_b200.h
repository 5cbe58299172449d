/* zonos_b200 -- C ABI of the B200-native Zonos hot path (libzonos_b200.so).
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch types.  Every device buffer is owned by the
 * caller (the Python shim allocates torch tensors and passes data_ptr()); the library owns only opaque session
 * handles.  Every entry point takes the CUDA stream (cudaStream_t as void*) it must run on, never synchronises
 * unless documented, and returns 0 (ZB200_OK) or a negative status with a message in zb200_last_error().
 *
 * Each entry point names the reference (coezbek/Zonos) interface it replaces.
 */
#ifndef ZONOS_B200_H
#define ZONOS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ZB200_OK 0
#define ZB200_ERR_INVALID (-1) /* bad argument / shape -> ValueError in the shim  */
#define ZB200_ERR_CUDA (-2)    /* CUDA runtime error   -> RuntimeError            */
#define ZB200_ERR_STATE (-3)   /* call order / overflow -> AssertionError (e.g. KV overflow, _torch.py:44-45) */

#define ZB200_ABI_VERSION 1
#define ZB200_PAGE 64        /* KV positions per page */
#define ZB200_HEAD_DIM 128   /* head dim of every Zonos-v0.1 attention layer */
#define ZB200_MAX_CODEBOOKS 9

int zb200_version(void);
const char* zb200_last_error(void);

/* ---- zonos/codebook_pattern.py:5-7 apply_delay_pattern, :10-12 revert_delay_pattern (int64, bit-exact) ---- */
int zb200_delay_apply(const int64_t* codes, int64_t* delayed, int batch, int n_q, int t, int64_t mask_token, void* stream);
int zb200_delay_revert(const int64_t* delayed, int64_t* codes, int batch, int n_q, int t_delayed, void* stream);

/* ---- weight repack (done once in Zonos.from_local, zonos/model.py:65-88) ----------------------------------
 * src: row-major [n_rows, k] 16-bit.  dst: [n_tiles][k/64][tile_n*128 B] shared-memory images (SWIZZLE_128B).
 * row_mode 0: tile row r of tile j <- src row j*tile_n + r (zero beyond n_rows)
 * row_mode 1: gated-MLP pairing (fc1, _torch.py:151-152): first tile_n/2 rows <- value rows j*tile_n/2 + r,
 *             last tile_n/2 rows <- gate rows n_rows/2 + j*tile_n/2 + r
 * row_mode 2: per-head padding (heads, model.py:100-101): src is [n_heads_out][rows_per_head][k], every head is
 *             padded to rows_per_head_padded rows (multiple of tile_n) with zeros. */
int zb200_pack_weight(const void* src, void* dst, int n_rows, int k, int tile_n, int row_mode, int rows_per_head,
                      int rows_per_head_padded, void* stream);
int64_t zb200_packed_weight_bytes(int n_rows, int k, int tile_n, int row_mode, int rows_per_head, int rows_per_head_padded);

/* ---- unit-test hooks for single kernels ------------------------------------------------------------------- */
/* out[m, n] = bf16( sum_k a[m,k] * w[n,k] ), packed weights with tile_n = 64. splits >= 1 (split-K). ws/tickets
 * may be NULL when splits == 1. */
int zb200_gemm_bf16(const void* a, int m, int lda, const void* w_packed, int n, int k, void* out, int ldo, int splits,
                    float* ws, int* tickets, void* stream);
/* Same contract (no split-K) through the persistent many-row kernel of the prefill (csrc/gemm_big.cuh): A by 2-D TMA,
 * two TMEM accumulators.  lda must be a multiple of 8 elements, k a multiple of 64. */
int zb200_gemm_big_bf16(const void* a, int m, int lda, const void* w_packed, int n, int k, void* out, int ldo, void* stream);
/* Same contract for m <= 128 rows through the weight-streaming stream-K kernel of the decode step (csrc/gemm_skinny.cuh)
 * on `grid` CTAs (grid <= 0: planned like the engine's ticketed GEMMs, at most 2 CTAs per SM).  ws: >= grid * 32768
 * floats, tickets: >= ceil(n/64) zeroed ints. */
int zb200_gemm_skinny_bf16(const void* a, int m, int lda, const void* w_packed, int n, int k, void* out, int ldo, int grid,
                           float* ws, int64_t ws_floats, int* tickets, int n_tickets, void* stream);
/* Same contract for m <= 8 rows through the operand-swapped weight-streaming kernel (csrc/gemm_swap.cuh: weights are the
 * 128-row UMMA operand).  k % 128 == 0; ws: >= grid * 2048 floats, tickets: >= ceil(n/128) zeroed ints. */
int zb200_gemm_swap_bf16(const void* a, int m, int lda, const void* w_packed, int n, int k, void* out, int ldo, int grid,
                         float* ws, int64_t ws_floats, int* tickets, int n_tickets, void* stream);
/* The out_proj / fc2 + residual + next LayerNorm group of the decode step (_torch.py:99-102): x <- bf16(x + bf16(a w^T)),
 * xn <- LayerNorm(x) * ln_w + ln_b, as a split-K weight-streaming GEMM whose partial sums are reduced by the fused
 * residual + LayerNorm kernel.  m <= 128, n % 256 == 0 (<= 4096); ws: >= splits*m*n floats. */
int zb200_gemm_resid_ln(const void* a, int m, int lda, const void* w_packed, int n, int k, int splits, float* ws,
                        int64_t ws_floats, void* x, const void* ln_w, const void* ln_b, void* xn, float eps, void* stream);
/* apply_heads + CFG mix of _compute_logits (zonos/model.py:100-116) as one weight-streaming GEMM (the kernel the engine
 * runs at the end of its step): xn bf16 [2*batch][d] (cond rows, then uncond rows), heads packed with row_mode 2 ->
 * logits fp32 [batch][n_q][vocab] = uncond + (cond - uncond) * cfg_scale, slots >= 1025 = -inf.  2*batch <= 128;
 * ws / tickets as zb200_gemm_skinny_bf16. */
int zb200_heads_cfg(const void* xn, int batch, int d, const void* w_heads_packed, int n_q, int vocab, int head_rows_padded,
                    float cfg_scale, float* logits, int grid, float* ws, int64_t ws_floats, int* tickets, int n_tickets,
                    void* stream);
/* Debug (libraries built with ZB200_TRACE=1 only, ZB200_ERR_STATE otherwise): device_buf uint64 [64][160][12] receives the
 * in-kernel globaltimer timeline of the following zb200_gemm_skinny_bf16 launches (NULL: off); returns the number of
 * launches traced since the previous call. */
int zb200_debug_trace(unsigned long long* device_buf);
/* nn.LayerNorm(d, eps) with affine weight+bias on bf16 rows, fp32 statistics (_torch.py:62,88,90). */
int zb200_layernorm(const void* x, const void* w, const void* b, void* out, int rows, int d, float eps, void* stream);
/* Zonos.embed_codes (model.py:97-98): sum over codebooks in order, each add rounded to bf16.
 * ids int32, element (b, q, s) at ids[b*ids_stride_b + q*ids_stride_q + s], s in [0, n_cols); emb bf16 [n_q][vocab][d];
 * out row (b*out_stride_b + out_off + s), written again dup_rows_offset rows further when dup_rows_offset > 0. */
int zb200_embed_sum(const int32_t* ids, int ids_stride_b, int ids_stride_q, int n_cols, const void* emb, void* out,
                    int batch, int n_q, int vocab, int d, int out_stride_b, int out_off, int dup_rows_offset, void* stream);

/* Paged KV pool layout: [layer][page][kv][n_kv_heads][64 positions][128] bf16 (see csrc/kv.cuh).  Test hook that
 * scatters k, v bf16 [n_seq, seq_len, n_kv_heads, 128] to positions pos0.. (the product path writes the cache from
 * the in_proj epilogue; replaces _update_kv_cache, zonos/backbone/_torch.py:33-49). */
int zb200_kv_write(const void* k, const void* v, void* pool, const int* block_table, int pages_per_seq, int num_pages,
                   int n_kv_heads, int layer, int n_seq, int seq_len, int pos0, void* stream);
/* Decode attention, one query per sequence over n_keys cached positions (non-causal over the visible keys,
 * scale 1/sqrt(128), GQA): F.scaled_dot_product_attention at _torch.py:136 with S == 1.  q/out bf16 [n_seq, H*128].
 * ws: n_seq*n_kv_heads*nsplit*520 floats, tickets: n_seq*n_kv_heads zeroed ints (both only for nsplit > 1). */
int zb200_attn_decode(const void* q, const void* pool, const int* block_table, int pages_per_seq, int num_pages,
                      int n_heads, int n_kv_heads, int layer, int n_seq, int n_keys, int nsplit, float* ws, int* tickets,
                      void* out, void* stream);
/* Allocation-free variant of zb200_attn_decode.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  Same computation and arguments as zb200_attn_decode, minus ws / tickets: the split-KV scratch lives in a
 * file-scope `static __device__` array of the library (csrc/attention.cu), 8.14 MiB that every GPU loading the library
 * holds until exit.  The call performs no cudaMalloc / cudaFree / allocating library routine (one cudaMemsetAsync of the
 * merge tickets on `stream`, and on the first call per device one cudaGetSymbolAddress).
 * Limit: n_seq * n_kv_heads * nsplit <= ZB200_ATTN_NOALLOC_MAX_UNITS, otherwise ZB200_ERR_INVALID with a message that
 * names the limit (use zb200_attn_decode with your own scratch then).
 * All calls on a device share the one array, which keeps the last call's contents: calls that OVERLAP on the GPU
 * (different streams, or graph branches without a dependency) race on it -- serialise them. */
#define ZB200_ATTN_NOALLOC_MAX_UNITS 4096
int zb200_attn_decode_noalloc(const void* q, const void* pool, const int* block_table, int pages_per_seq, int num_pages,
                              int n_heads, int n_kv_heads, int layer, int n_seq, int n_keys, int nsplit, void* out,
                              void* stream);
/* Causal prefill attention (_torch.py:136 with S > 1): query s of a sequence sees keys 0..pos0+s.
 * q/out bf16 [n_seq, seq_len, H*128]. */
int zb200_attn_prefill(const void* q, const void* pool, const int* block_table, int pages_per_seq, int num_pages,
                       int n_heads, int n_kv_heads, int layer, int n_seq, int seq_len, int pos0, void* out, void* stream);

/* ---- zonos/sampling.py:232-328 sample_from_logits + the per-step bookkeeping of zonos/model.py:345-424 -------------
 * A decode step calls pass 1 then pass 2 (pass 2 re-draws every row only if some utterance produced its first EOS in
 * pass 1, model.py:380-393; otherwise it only runs the bookkeeping).  All state lives on the device:
 * state[]: int32[16] {0 offset (column written this step), 1 pos, 2 done, 3 final_offset, 4 draws, 5 any_fresh,
 * 6 steps, 7/8 tickets}.  loop_mode 0 = first frame after the prefill (model.py:304-314). */
typedef struct {
  const float* logits;   /* [batch, n_q, vocab] fp32, CFG-mixed, pad slot -inf (model.py:111-115) */
  const float* noise;    /* optional injected Exp(1) noise [n_draws][batch, n_q, vocab]; NULL -> Philox4x32-10 */
  int32_t* delayed;      /* [batch, n_q, t_delayed] delayed codes, -1 = unknown */
  int32_t* tok1;         /* [batch, n_q] */
  int32_t* tok2;         /* [batch, n_q] */
  int32_t* state;        /* [16] */
  int32_t* remaining;    /* [batch] remaining_steps */
  int32_t* grace;        /* [batch] steps_after_eos */
  uint8_t* stopping;     /* [batch] */
  uint8_t* eos_mode;     /* [batch] */
  uint8_t* mask2;        /* [batch] scratch */
  uint8_t* eos_old;      /* [batch] scratch */
  int32_t* host_flags;   /* optional pinned host int32[2] {done, steps} */
  int32_t batch, n_q, vocab, t_delayed;
  int32_t loop_mode, first_offset, row_base;
  uint64_t seed;
  float temperature, top_p, min_p, linear, conf, quad, rep_penalty;
  int32_t top_k, rep_window;
} zb200_sampler_args;
int zb200_sample(const zb200_sampler_args* args, int pass, void* stream);

/* ---- zonos/autoencoder.py:27-28 DACAutoencoder.encode -> HF DacModel.encode (modeling_dac.py:581-605)
 * The reference encodes in fp32 (no autocast) and the result is integer codes picked by an argmax, so these are plain
 * fp32 kernels on channels-first tensors [B][C][T] like torch.
 * zb200_conv1d_f32: nn.Conv1d with stride / dilation / zero padding (DacEncoder convs :442-472, DacEncoderBlock's strided
 *   conv :222-224, DacResidualUnit's convs :181-185); w_ckn = the weight permuted to [C_in][K][C_out]; resid (optional,
 *   [B][C_out][T_out]) is added to the result (the residual unit's `hidden_state + output_tensor`, :207).
 * zb200_snake_f32: Snake1d (:94-99), alpha fp32 [C].
 * zb200_rvq_encode: DacResidualVectorQuantizer.forward (:300-343) in eval mode with all codebooks -- per codebook in_proj,
 *   F.normalize, dist = -(|e|^2 - 2 e.c) + |c|^2 over the normalised codebook, first-index argmax (:154-166), out_proj of
 *   the picked code, residual update.  latent fp32 [B][H][T]; in_w [Q][D][H], in_b [Q][D], codebook [Q][N][D],
 *   codebook_norm [Q][N][D+1] (F.normalize(codebook) and, last, its squared norm), out_w [Q][H][D], out_b [Q][H];
 *   codes int64 [B][Q][T]. */
int zb200_conv1d_f32(const float* x, const float* w_ckn, const float* bias, const float* resid, float* out, int batch, int c_in,
                     int c_out, int t_in, int t_out, int k, int stride, int dilation, int pad, void* stream);
int zb200_snake_f32(const float* x, const float* alpha, float* out, int batch, int channels, int t, void* stream);
int zb200_rvq_encode(const float* latent, const float* in_w, const float* in_b, const float* codebook, const float* codebook_norm,
                     const float* out_w, const float* out_b, int64_t* codes, int batch, int n_q, int hidden, int dim, int n_codes,
                     int t, void* stream);

/* ---- zonos/speaker_cloning.py:387-411 SpeakerEmbeddingLDA.forward: the speaker encoder, fp32 like the
 * reference, channels-first tensors like torch.  Once per new speaker; csrc/speaker.cu.
 * zb200_resample_f32: torchaudio.transforms.Resample as SpeakerEmbedding.prepare_input uses it (:371-380): out[b][j*new + q] =
 *   sum_k kernel[q][k] * x[b][j*orig + k - width], kernel fp32 [new][2*width + orig] from torchaudio's sinc_interp_hann
 *   formula (orig, new already divided by their gcd); n_out <= ceil(n_in * new / orig).
 * zb200_logfbank_f32: logFbankCal.forward (:12-34) = MelSpectrogram(16 kHz, n_fft 512, win 400, hop 160, 80 HTK mels, power 2,
 *   centre + reflect padding) -> log(. + 1e-6) -> minus the mean over time.  window fp32 [400] (periodic Hann), mel_fb fp32
 *   [257][80]; out fp32 [batch][80][n_frames], n_frames = 1 + n / 160.
 * zb200_conv2d_f32: nn.Conv2d 3x3 (padding 1) or 1x1, stride 1 or 2, no bias, followed by an eval-mode BatchNorm2d folded into
 *   scale / shift per output channel (NULL: none) and an optional ReLU (SimAMBasicBlock's conv1+bn1+relu, conv2+bn2 and the
 *   downsample branch, :68-88; ResNet.conv1, :169-171).  w_ckn = the weight permuted to [C_in][k*k][C_out].
 * zb200_simam_residual_f32: SimAM (:90-95) on every (sample, channel) plane of hw values, + shortcut, ReLU (:85-87).
 * zb200_relu_affine_f32: ReLU then a per-channel affine (ASP's ReLU + BatchNorm1d, :46-47) on [batch][channels][t].
 * zb200_asp_pool_f32: softmax over time of the attention logits, weighted mean and standard deviation (:54-57):
 *   x, logits fp32 [batch][channels][t] -> out fp32 [batch][2*channels] = (mu | sg).
 * zb200_linear_f32: nn.Linear on one vector per sample (the bottleneck :213,222 and the LDA :404-411): out[b] = bias + x[b] W^T,
 *   w_t = the weight transposed to [c_in][c_out].
 * ASP's 1x1 Conv1d layers run on zb200_conv1d_f32. */
int zb200_resample_f32(const float* x, const float* kernel, float* out, int batch, int n_in, int n_out, int orig_freq, int new_freq,
                       int width, void* stream);
int zb200_logfbank_f32(const float* wav16k, const float* window, const float* mel_fb, float* out, int batch, int n, int n_frames,
                       void* stream);
int zb200_conv2d_f32(const float* x, const float* w_ckn, const float* scale, const float* shift, float* out, int batch, int c_in,
                     int h, int w, int c_out, int k, int stride, int relu, void* stream);
int zb200_simam_residual_f32(const float* x, const float* shortcut, float* out, int planes, int hw, float lambda_p, void* stream);
int zb200_relu_affine_f32(const float* x, const float* scale, const float* shift, float* out, int batch, int channels, int t,
                          void* stream);
int zb200_asp_pool_f32(const float* x, const float* logits, float* out, int batch, int channels, int t, void* stream);
int zb200_linear_f32(const float* x, const float* w_t, const float* bias, float* out, int batch, int c_in, int c_out, void* stream);

/* ---- zonos/conditioning.py:373-389 PrefixConditioner.forward ---------------------------------------
 * One launch: every conditioner's output row at its position of the prefix (rounded to bf16 like the eager reference),
 * then nn.LayerNorm over d.  The prefix is tiled by `segments` in conditioner order; a segment is one of
 *   ZB200_COND_ROWS    rows of an embedding table picked by ids: EspeakPhonemeConditioner (conditioning.py:295-312; ids =
 *                      tokenised phonemes, text -> phonemes stays on the host) and IntegerConditioner (:340-349, id_offset =
 *                      -min_val).  table bf16 [n][d], x int32 [B][len]
 *   ZB200_COND_VECTOR  a conditioner called with None: its learned uncond_vector (:38-40, :46-48).  table bf16 [d], len 1
 *   ZB200_COND_FOURIER FourierConditioner (:318-337): f = bf16(2 pi bf16((x - min) / (max - min))) @ W^T, [cos f | sin f].
 *                      table = weight bf16 [d/2][in_dim], x fp32 [B][len][in_dim]
 *   ZB200_COND_LINEAR  PassthroughConditioner (:352-358) + the base class's linear `project` (:26-27): bf16(x W^T + bias).
 *                      table = project.weight bf16 [d][in_dim], bias bf16 [d] or NULL, x bf16 [B][len][in_dim]
 * batch_stride: elements between utterances in x (0 = one input broadcast over the batch, like the reference's expand).
 * out bf16 [batch][l_total][d]. */
#define ZB200_MAX_COND_SEGMENTS 12
enum { ZB200_COND_ROWS = 0, ZB200_COND_VECTOR = 1, ZB200_COND_FOURIER = 2, ZB200_COND_LINEAR = 3 };
typedef struct {
  int32_t kind, start, len, in_dim, batch_stride, id_offset;
  float min_val, max_val;
  const void* table;
  const void* bias;
  const void* x;
} zb200_cond_segment;
int zb200_prefix_conditioner(const zb200_cond_segment* segs, int n_segs, int batch, int l_total, int d, const void* ln_w,
                             const void* ln_b, float eps, void* out, void* stream);

/* ---- engine: the transformer backbone + heads behind Zonos._prefill / _decode_one_token / generate ------------------
 * Replaces TorchZonosBackbone.forward (zonos/backbone/_torch.py:73-80), Zonos.embed_codes/apply_heads/_compute_logits
 * (zonos/model.py:97-116), _decode_one_token (:118-179, incl. the CUDA-graph replay) and _prefill (:181-202).
 * All buffers are caller-owned device memory; the handle stores the description and a captured CUDA graph. */
typedef struct {
  int32_t d_model, n_layer, n_heads, n_kv_heads, d_ff, n_q, vocab, head_rows_padded;
  float ln_eps, cfg_scale;
  int32_t batch;                 /* utterances B; the backbone runs 2B rows (cond | uncond) */
  int32_t rows_max;              /* capacity (rows) of x/xn/q/att/h */
  int32_t pages_per_seq, num_pages, t_delayed, attn_nsplit;
  const void* emb;               /* bf16 [n_q][vocab][d_model] */
  const void *ln1_w, *ln1_b, *ln2_w, *ln2_b; /* bf16 [n_layer][d_model] */
  const void *lnf_w, *lnf_b;     /* bf16 [d_model] */
  const void *w_qkv, *w_out, *w_fc1, *w_fc2; /* packed (zb200_pack_weight, tile 64; fc1 row_mode 1), layers contiguous */
  int64_t qkv_stride, out_stride, fc1_stride, fc2_stride; /* bytes between layers */
  const void* w_heads;           /* packed row_mode 2, rows_per_head_padded = head_rows_padded */
  const float* rope;             /* fp32 [n_pos][64][2] (cos, sin), precompute_freqs_cis (_torch.py:9-15) */
  void* kv_pool;                 /* bf16 [n_layer][num_pages][2][n_kv_heads][64][128] */
  const int32_t* block_table;    /* [2B][pages_per_seq] */
  void *x, *xn, *q, *att, *h;    /* bf16 [rows_max][d_model | d_model | H*128 | H*128 | d_ff] */
  float* logits;                 /* fp32 [B][n_q][vocab] */
  float* gemm_ws; int64_t gemm_ws_floats; int32_t* gemm_tickets; int32_t n_gemm_tickets;
  float* attn_ws; int32_t* attn_tickets;
  int32_t* state;                /* int32[16], see zb200_sampler_args */
  int32_t* delayed;              /* int32 [B][n_q][t_delayed] */
  int32_t flags;                 /* ZB200_ENGINE_* bits; 0 = defaults */
} zb200_engine_desc;
/* flags: run the decode step on the fused per-layer chain kernel (csrc/chain.cuh: one persistent launch for everything
 * between two attention kernels, software grid barriers) instead of one launch per GEMM + reduction kernel.  Same
 * rounding points, different split-K summation order.  Measured SLOWER than the default on B200: opt-in. */
#define ZB200_ENGINE_CHAIN 1
/* flags: the engine only serves zb200_engine_forward (the backbone plugin contract); emb / w_heads / logits / state /
 * delayed may be NULL and n_q / vocab are ignored. */
#define ZB200_ENGINE_BACKBONE_ONLY 2
typedef struct zb200_engine zb200_engine;
int zb200_engine_create(const zb200_engine_desc* desc, zb200_engine** out);
void zb200_engine_destroy(zb200_engine* e);
/* The backbone plugin contract on its own -- TorchZonosBackbone.forward(hidden_states, inference_params)
 * (zonos/backbone/_torch.py:73-80): x holds hidden [2B*seq_len, d_model]; positions are pos0 .. pos0+seq_len-1
 * (inference_params.seqlen_offset), causal attention over the cache for seq_len > 1, all cached keys for seq_len == 1;
 * writes the KV cache and norm_f of EVERY row to out (bf16 [2B*seq_len, d_model]).  ZB200_ERR_STATE on cache overflow
 * (the reference's assert, _torch.py:44-45). */
int zb200_engine_forward(zb200_engine* e, int seq_len, int pos0, void* out, void* stream);
/* x holds the hidden states [2B*seq_len, d_model] (prefix conditioning ++ embedded prefix codes); writes the KV cache
 * for positions [0, seq_len) and the CFG logits of the last position. */
int zb200_engine_prefill(zb200_engine* e, int seq_len, void* stream);
/* One decode step up to the logits: embeds column state[offset]-1 of `delayed`, position state[pos]. */
int zb200_engine_decode_logits(zb200_engine* e, void* stream);
/* decode_logits + sampler pass 1 + pass 2 (one loop iteration of model.py:345-424), no host synchronisation. */
int zb200_engine_step(zb200_engine* e, const zb200_sampler_args* smp, void* stream);
/* Captures one step into a CUDA graph (stream capture on `stream`); zb200_engine_run replays it n_steps times. */
int zb200_engine_capture(zb200_engine* e, const zb200_sampler_args* smp, void* stream);
int zb200_engine_run(zb200_engine* e, int n_steps, void* stream);
/* number of kernel launches in one step (for the bench's gpu_launches claim) */
int zb200_engine_launches_per_step(const zb200_engine* e);
/* debug hook (no reference counterpart): globaltimer stamps of the fused chain kernel, uint64
 * [n_layer + 1 launches][SMs][4 phases][8 events] in device memory; pass NULL to switch it off.  Takes effect for launches issued afterwards (not for a captured graph). */
int zb200_engine_set_chain_trace(zb200_engine* e, unsigned long long* device_buf);

/* ---- Hybrid backbone, Mamba2 mixer decode step (zonos/backbone/_mamba_ssm.py:18-31 builds the layers
 * with mamba_ssm's create_block, the per-token arithmetic is mamba_ssm 2.2.4 Mamba2.step).  zxbcdt is the in_proj output
 * [rows][ld] bf16 = [z (d_inner) | x (d_inner) B (G*N) C (G*N) | dt (H)].  States are caller-owned bf16 tensors
 * (what allocate_inference_cache returns, _mamba_ssm.py:37-41). */
/* causal_conv1d_update + SiLU: conv_state bf16 [rows][channels][4] shifted in place, out bf16 [rows][channels]. */
int zb200_mamba2_conv_update(void* conv_state, const void* zxbcdt, int ld, int col0, const void* w, const void* bias,
                             void* out, int rows, int channels, int d_conv, void* stream);
/* selective_state_update (dt_softplus, D skip, no z): ssm_state bf16 [rows][H][P][N] updated in place,
 * xbc = conv output [rows][ld_xbc], dt = zxbcdt[:, dt_col0 + h], y bf16 [rows][H*P]. */
int zb200_mamba2_ssd_step(void* ssm_state, const void* xbc, int ld_xbc, const void* zxbcdt, int ld, int dt_col0,
                          const float* A_log, const float* D, const float* dt_bias, void* y, int rows, int n_heads,
                          int head_dim, int d_state, int n_groups, void* stream);
/* Prefill forms of the two kernels above over n_seq sequences of seq_len tokens (rows are sequence-major: row = seq *
 * seq_len + t): the conv window / SSM state stay in registers over the sequence and the final ones are written to
 * conv_state / ssm_state (use_initial_state = 0: start from zeros, as a fresh allocate_inference_cache does). */
int zb200_mamba2_conv_seq(void* conv_state, const void* zxbcdt, int ld, int col0, const void* w, const void* bias,
                          void* out, int n_seq, int seq_len, int channels, int d_conv, int use_initial_state, void* stream);
int zb200_mamba2_ssd_scan(void* ssm_state, const void* xbc, int ld_xbc, const void* zxbcdt, int ld, int dt_col0,
                          const float* A_log, const float* D, const float* dt_bias, void* y, int n_seq, int seq_len,
                          int n_heads, int head_dim, int d_state, int n_groups, int use_initial_state, void* stream);
/* RMSNormGated(norm_before_gate=False): out = rmsnorm_group(y * silu(z)) * w, z = zxbcdt[:, :d_inner]. */
int zb200_mamba2_gated_rmsnorm(const void* y, const void* zxbcdt, int ld, const void* w, void* out, int rows, int d_inner,
                               int n_groups, float eps, void* stream);

/* Block wiring of the hybrid backbone with fused_add_norm (mamba_ssm layer_norm_fn, prenorm=True, residual_in_fp32=True;
 * _mamba_ssm.py:18-31,46-57): residual fp32 [rows][d] <- x + residual (has_residual = 0: residual <- x), out bf16 =
 * norm(residual) * w (+ b); rms != 0: RMSNorm, else LayerNorm.  b may be NULL. */
int zb200_add_norm(const void* x, float* residual, int has_residual, const void* w, const void* b, void* out, int rows,
                   int d, float eps, int rms, void* stream);
/* The same block wiring with the GEMM in front of it (decode rows, m <= 128): hidden = a * w^T in bf16 (the mixer's out_proj,
 * MHA's out_proj or the GatedMLP's fc2: mamba_ssm Block, _mamba_ssm.py:46-57), then residual <- hidden + residual and
 * out = norm(residual) of the NEXT block (or norm_f).  One deferred split-K weight-streaming GEMM whose fp32 partial sums
 * (ws, >= m*n floats; more capacity allows more splits) are added up, in split order, by the add + norm kernel. */
int zb200_gemm_add_norm(const void* a, int m, int lda, const void* w_packed, int n, int k, float* ws, int64_t ws_floats,
                        float* residual, int has_residual, const void* norm_w, const void* norm_b, void* out, float eps, int rms,
                        void* stream);

/* Hybrid attention layers (mamba_ssm MHA; _mamba_ssm.py:18-31 with attn_layer_idx / attn_cfg): splits the Wqkv output
 * [rows][(H + 2*Hkv)*128] into q [rows][H*128], k and v [rows][Hkv*128] and applies the non-interleaved rotary embedding
 * (pairs (i, i+64), cos / sin rounded to bf16) at position pos0 + (row % seq_len); inv_freq fp32 [64]. */
int zb200_rope_split(const void* qkv, const float* inv_freq, void* q, void* k, void* v, int rows, int seq_len, int pos0,
                     const int* pos_ptr, int n_heads, int n_kv_heads, void* stream);
/* Decode-step forms with the position / column in DEVICE memory, so that a host-sequenced step can be replayed from a
 * CUDA graph (the engine's own kernels take the same pointers): one token per sequence at position *pos_ptr; the column
 * of `ids` embedded is *col_ptr + col_add (out bf16 [batch][d], written again dup_rows_offset rows further if > 0).
 * zb200_rope_split takes pos_ptr too (NULL: pos0). */
int zb200_kv_write_at(const void* k, const void* v, void* pool, const int* block_table, int pages_per_seq, int num_pages,
                      int n_kv_heads, int layer, int n_seq, const int* pos_ptr, void* stream);
int zb200_attn_decode_at(const void* q, const void* pool, const int* block_table, int pages_per_seq, int num_pages,
                         int n_heads, int n_kv_heads, int layer, int n_seq, const int* pos_ptr, int nsplit, float* ws,
                         int* tickets, void* out, void* stream);
int zb200_embed_sum_at(const int32_t* ids, int ids_stride_b, int ids_stride_q, const int* col_ptr, int col_add,
                       const void* emb, void* out, int batch, int n_q, int vocab, int d, int dup_rows_offset, void* stream);
/* mamba_ssm GatedMLP activation: fc1_out bf16 [rows][2*d_ff] -> out[r][c] = fc1_out[r][c] * silu(fc1_out[r][d_ff + c]). */
int zb200_gated_mlp_act(const void* fc1_out, void* out, int rows, int d_ff, void* stream);

/* ---- DACAutoencoder.decode (zonos/autoencoder.py:44-47 -> transformers DacModel.decode, modeling_dac.py:608-638) ---
 * The decoder is described as a list of implicit-GEMM layers over five caller-owned fp16 work regions (channels-last
 * [batch][time][pitch]); the host shim (zonos_b200/dac.py) builds the list from the HF state dict once. */
typedef struct {
  int32_t taps;          /* 7, 1, or 2 (transposed conv) */
  int32_t shift0, shift_step; /* time shift of tap i = shift0 + i*shift_step */
  int32_t up;            /* 0 = Conv1d; s > 0 = ConvTranspose1d with stride s, kernel 2s, padding s/2 */
  int32_t pitch_in;      /* channels == channel pitch of the input region (multiple of 8; k-blocks of 64 are zero-filled) */
  int32_t c_out, pitch_out, tile_n; /* output channels, output pitch, N tile (64 or 96) */
  int32_t t_mul_in;      /* input rows per code frame (T_in = t_mul_in * T) */
  int32_t in_region, resid_region, raw_region, act_region; /* region index or -1 */
  const void* w;         /* packed fp16 [n_total][taps*round_up(pitch_in,64)], n_total = c_out * max(up,1) */
  const float* bias;     /* [n_total] */
  const float* alpha;    /* [c_out] Snake alpha of the consumer (NULL when act_region < 0) */
  const float* inv_alpha;/* [c_out] 1/(alpha + 1e-9) */
} zb200_dac_layer;
typedef struct {
  int32_t n_q, n_codes, latent, n_layers;
  const float* code_table;       /* fp32 [n_q][n_codes][latent]: out_proj_k(codebook_k) + bias_k */
  const zb200_dac_layer* layers; /* host array */
  int32_t z_region;              /* region receiving from_codes output (pitch = latent) */
  int32_t final_region, final_c, final_pitch, final_t_mul; /* input of the last conv (k7 -> 1 channel) + tanh */
  const float* final_w;          /* fp32 [7][final_c] */
  float final_bias;
  /* Optional tensor-core form of the last conv: eight consecutive time steps are one input row of 8*final_c channels,
   * so the k7 conv is a 3-tap conv with 8 output columns (the 8 time phases).  final_wp: packed fp16 (tile 16)
   * [8][3][8*final_c], W'[r][s+1][j*final_c + c] = w[8s + j - r + 3][c] (0 outside the 7 taps); final_bias8: fp32 [8].
   * NULL: scalar kernel. */
  const void* final_wp;
  const float* final_bias8;
} zb200_dac_plan;
/* codes int64 [batch][n_q][t]  ->  wav fp32 [batch][final_t_mul * t].  regions: 5 device buffers of region_halves fp16. */
int zb200_dac_decode(const zb200_dac_plan* plan, const int64_t* codes, int batch, int t, void* const* regions,
                     int64_t region_halves, float* wav, void* stream);

/* ---- DACAutoencoder.codes_to_wavs post-processing (zonos/autoencoder.py:188-245) on a ragged batch of decoded
 * waveforms: normalize_loudness (autoencoder.py:172-186 -> pyloudnorm 0.1.1 Meter.integrated_loudness: K-weighting
 * high-shelf + high-pass biquads, 75 %-overlapped gating blocks, -70 LUFS absolute and -10 LU relative gates),
 * trim_silence (autoencoder.py:49-90, including its quirks: the tail scan never tests the last frame and cuts the
 * frame it detects), the 512-sample linear fade-in and the logarithmic fade-out (autoencoder.py:233-241).
 * Everything the reference derives from the utterance LENGTH on the host (block size 0.4 / 0.1 s, the number of gating
 * blocks and their integer sample bounds, computed with the reference's own float expressions) is passed in as
 * tables (block bounds as indices into one sorted edge list, so that overlapping blocks share their partial sums); the fade curves are torch.linspace / torch.logspace evaluated once on the host so the device multiplies by
 * the same fp32 values. */
typedef struct {
  double shelf_b[3], shelf_a[3]; /* K-weighting stage 1 (a[0] == 1) */
  double hp_b[3], hp_a[3];       /* K-weighting stage 2 */
  int32_t chunk;                 /* samples one thread filters sequentially (multiple of 4) */
  int32_t max_blocks;            /* rows of blk_lo / blk_hi */
  double chunk_transition[16];   /* row-major 4x4: zero-input transition of the cascaded filter state over `chunk` samples */
  const int32_t* edges[2];       /* device int32 [n_edges[k]], k = block size 0.4 s / 0.1 s: the sorted distinct first /
                                    one-past-last samples of all gating blocks (edges[k][0] == 0); consecutive edges
                                    bound a "segment"; a chunk may touch at most 8 segments */
  int32_t n_edges[2];            /* <= 2 * max_blocks + 2 */
  const int32_t* blk_lo[2];      /* device int32 [max_blocks]: gating block j = segments blk_lo[k][j] .. blk_hi[k][j] - 1 */
  const int32_t* blk_hi[2];      /* (the block's samples are not clipped to the utterance in the table; the kernels clip) */
  double inv_norm[2];            /* 1 / (block_size * rate) */
  const float* fade_in;          /* device fp32 [512] = torch.linspace(0, 1, 512) */
  const float* fade_out;         /* device fp32: torch.logspace(0, -10, 512*k) for k = 1..20, concatenated */
  double target_lufs;            /* -23.0 in codes_to_wavs */
  float trim_threshold;          /* 1e-5 */
  int32_t flags;                 /* bit 0: no fades (trim_silence / normalize_loudness called on their own) */
} zb200_postproc_plan;
/* wav fp32 [batch][pitch] (pitch % 4 == 0, 16-byte aligned), n_samples int32 [batch] valid samples per utterance,
 * blk_sel int32 [batch]: 0 / 1 = index into the plan's block tables, -1 = loudness step skipped (the reference's
 * "audio too short" path: gain 1), n_blocks int32 [batch]; all device pointers.  Writes out fp32 [batch][pitch]
 * (utterance b occupies out[b][0 .. bounds[b][1]-bounds[b][0]), zero after), bounds int32 [batch][2] = the [start, end)
 * trim_silence kept, lufs double [batch] (integrated loudness before the gain), gain fp32 [batch].
 * ws: device scratch of at least zb200_postproc_ws_bytes(batch, max_n, chunk, max_blocks) bytes. */
int64_t zb200_postproc_ws_bytes(int batch, int max_n, int chunk, int max_blocks);
int zb200_codes_postprocess(const zb200_postproc_plan* plan, const float* wav, int64_t pitch, const int32_t* n_samples,
                            const int32_t* blk_sel, const int32_t* n_blocks, int batch, int max_n, float* out,
                            int32_t* bounds, double* lufs, float* gain, void* ws, int64_t ws_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ZONOS_B200_H */
