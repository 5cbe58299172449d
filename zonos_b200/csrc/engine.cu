// Session engine: sequences the kernels of the prefill and of one decode step, and owns the captured CUDA graph.
// Reference control flow being replaced: zonos/model.py:103-116 (_compute_logits), :118-179 (_decode_one_token),
// :181-202 (_prefill); zonos/backbone/_torch.py:73-80,99-102 (backbone / block forward).
#include <cstdio>
#include "../../include/zonos_b200.h"
#include "attention.cuh"
#include "chain.cuh"
#include "common.cuh"
#include "epilogues.cuh"
#include "epilogues_t.cuh"
#include "gemm_skinny.cuh"
#include "gemm_big.cuh"
#include "gemm_swap.cuh"
#include "gemm_tc.cuh"
#include "ops.cuh"
#include "sampler.cuh"

#include <algorithm>
#include <type_traits>
#include <cstdlib>

using namespace zb;
#ifdef SK_TRACE
extern unsigned long long* g_trace_buf;  // ops.cu: debug timeline of the weight-streaming GEMMs
extern int g_trace_next;
#endif

struct zb200_engine {
  zb200_engine_desc d;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  int launches = 0;  // kernels launched by the last step
  int sk_sms = 0;    // SMs the weight-streaming GEMM plans for (0: tiled kernel only)
  // L2 weight prefetch: a side stream (forked / joined inside the captured step) runs ahead of the GEMM chain
  int l2_prefetch = 0;
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  bool forked = false;
  unsigned long long* chain_trace = nullptr;  // debug timeline of the chain kernel (zb200_engine_set_chain_trace)
};

namespace {

// Asks the L2 to fetch [base, base+bytes) from HBM (TMA-engine bulk prefetch, no shared memory, no registers held):
// the weights of the NEXT GEMM of the step travel HBM -> L2 while the current kernels run their dependent chains.
__global__ void l2_prefetch_kernel(const uint8_t* base, long long bytes) {
  constexpr long long CHUNK = 16384;
  const long long n = (bytes + CHUNK - 1) / CHUNK;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long off = i * CHUNK;
    const unsigned sz = (unsigned)((bytes - off < CHUNK ? bytes - off : CHUNK) & ~15LL);
    if (sz) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(base + off), "r"(sz) : "memory");
  }
}

constexpr int NT = 64;
constexpr int STAGES = 6;
constexpr int SM_COUNT = 148;

struct SplitPlan {
  int splits, kb_per_split;
};
// Fill the 148 SMs: n_tiles x splits CTAs, each split at least 4 k-blocks deep, workspace permitting.
SplitPlan plan_splits(const zb200_engine_desc& d, int n_tiles, int m_tiles, int k_blocks) {
  int s = 1;
  if (m_tiles == 1 && n_tiles < SM_COUNT) {
    s = SM_COUNT / n_tiles;
    if (s < 1) s = 1;
    while (s > 1 && k_blocks / s < 4) --s;
    while (s > 1 && ((int64_t)n_tiles * s * NT * 128 > d.gemm_ws_floats || n_tiles > d.n_gemm_tickets)) --s;
  }
  SplitPlan p;
  p.kb_per_split = ceil_div(k_blocks, s);
  p.splits = ceil_div(k_blocks, p.kb_per_split);
  return p;
}

GemmCore make_core(const zb200_engine_desc& d, const void* w, int n_tiles, int m_rows, int k) {
  GemmCore g{};
  const int m_tiles = ceil_div(m_rows, GEMM_BM);
  const SplitPlan sp = plan_splits(d, n_tiles, m_tiles, k / 64);
  g.Wp = w, g.k_blocks = k / 64, g.kb_per_split = sp.kb_per_split, g.splits = sp.splits, g.m_valid = m_rows;
  g.partial_ws = d.gemm_ws, g.tickets = d.gemm_tickets, g.stream_weights = (m_tiles == 1);
  return g;
}

KvLayout kv_of(const zb200_engine_desc& d, int layer) {
  KvLayout kv;
  kv.pool = (bf16*)d.kv_pool, kv.block_table = d.block_table, kv.pages_per_seq = d.pages_per_seq;
  kv.num_pages = d.num_pages, kv.n_kv_heads = d.n_kv_heads, kv.layer = layer;
  return kv;
}

// ---- decode-step GEMM plans (weight-streaming kernel, gemm_skinny.cuh) -------------------------------------------------
// Every CTA should be exactly one fragment, so grids are aligned to the group structure:
//   ticketed epilogues (in_proj, fc1, heads): whole groups per CTA when there are enough groups; otherwise each group is
//     shared by at most 3 CTAs (the last one to arrive reduces, so the contributor count is the fix-up latency);
//   deferred epilogues (out_proj, fc2): n_groups * splits CTAs, the split sums are reduced by resid_ln_kernel.
// `sms` CTAs are resident at M > 32 (192 KB of shared memory each), 2 * sms at M <= 32.

SkinnyCore sk_core(const zb200_engine_desc& d, const void* w, int n_tiles, int rows, int k) {
  SkinnyCore g{};
  g.Wp = w, g.n_tiles = n_tiles, g.k_blocks = k / 64, g.m_valid = rows;
  g.partial_ws = d.gemm_ws, g.tickets = d.gemm_tickets, g.n_splits = 1, g.trace = nullptr, g.trace_id = -1;
  g.next = NextW{};
#ifdef SK_TRACE
  g.trace = g_trace_buf, g.trace_id = g_trace_buf ? g_trace_next++ : -1;
#endif
  return g;
}

// Descriptors of "the GEMM after this one" for the in-kernel L2 prefetch of its first weight bytes (gemm_skinny.cuh).
static int nextw_bytes(int rows) {
  static const int kb = zb::tuning_env("ZB200_NEXTW_KB") ? atoi(zb::tuning_env("ZB200_NEXTW_KB")) : -1;
  (void)rows;
  return kb > 0 ? kb * 1024 : 0;  // off by default: measured neutral at batch 64 and batch 1 (experiments)
}
NextW next_skinny(const void* w, int rows, int n_tiles, int k, int grid, bool prefer_k1) {
  const int mode = skinny_mode(rows, n_tiles, k / 64, prefer_k1);
  const int units = (n_tiles / skinny_tg(mode)) * ((k / 64) / skinny_kg(mode));
  NextW nx{(const uint8_t*)w, n_tiles, k / 64, std::min(grid, units), skinny_tg(mode), skinny_kg(mode), nextw_bytes(rows)};
  if (nx.bytes <= 0) nx.wp = nullptr;
  return nx;
}
int swap_grid(const zb200_engine* e, int n_tiles, int k) {
  const int groups = (n_tiles + 1) / 2, kbu = (k / 64) / 2, cap = 2 * e->sk_sms;
  return groups * std::max(1, std::min(std::min(SWP_MAXC, kbu), cap / groups));
}
NextW next_swap(const zb200_engine* e, const void* w, int rows, int n_tiles, int k) {
  NextW nx{(const uint8_t*)w, n_tiles, k / 64, swap_grid(e, n_tiles, k), 2, 2, nextw_bytes(rows)};
  if (nx.bytes <= 0) nx.wp = nullptr;
  return nx;
}

// Ticketed-epilogue GEMM: the weight-streaming kernel for the decode step (rows <= 128), the tiled kernel otherwise.
template <class ALoader, class Epilogue>
cudaError_t gemm_any(const zb200_engine* e, const void* w, int n_tiles, int rows, int k, const ALoader& al,
                     const Epilogue& epi, cudaStream_t st, const NextW& nx = NextW{}, const KvPrefetch& kvpf = KvPrefetch{}) {
  const zb200_engine_desc& d = e->d;
  if (rows <= GEMM_BM && e->sk_sms > 0) {
    const SkPlan p = plan_ticketed(e->sk_sms, rows, n_tiles, k / 64);
    SkinnyCore g = sk_core(d, w, n_tiles, rows, k);
    g.next = nx;
    g.kvpf = kvpf;
    return launch_gemm_skinny(g, al, epi, p.grid, st, p.prefer_k1);
  }
  GemmCore g = make_core(d, w, n_tiles, rows, k);
  if constexpr (std::is_same<ALoader, ARowMajor>::value) {
    // Many row tiles (the batched prefill): the persistent kernel overlaps each tile's epilogue with the next tile's
    // MMAs.  Small problems keep the per-tile kernel (split-K fills the SMs there).
    static const bool no_big = zb::tuning_env("ZB200_NO_BIG") != nullptr;
    const int m_tiles = ceil_div(rows, GEMM_BM);
    if (!no_big && g.splits == 1 && e->sk_sms > 0 && m_tiles * ceil_div(n_tiles, 4) >= e->sk_sms) {
      const cudaError_t rc = launch_gemm_big<1>(al.base, al.row_bytes, rows, k, w, n_tiles, epi, e->sk_sms, st);
      if (rc != cudaErrorNotSupported) return rc;
    }
  }
  return launch_gemm_tc<NT, STAGES, 1>(g, al, epi, n_tiles, ceil_div(rows, GEMM_BM), st);
}

#define ZB_TRY(expr)                                         \
  do {                                                       \
    cudaError_t _e = (expr);                                 \
    if (_e != cudaSuccess) return zb::cuda_fail(_e, #expr);  \
    ++e->launches;                                           \
  } while (0)

// Forks the side stream off `st` (everything launched on `st` so far is a dependency) and queues an L2 prefetch on it.
int prefetch_l2(zb200_engine* e, const void* w, int64_t bytes, cudaStream_t st) {
  if (!e->l2_prefetch || bytes <= 0) return ZB200_OK;
  ZB_CHECK_CUDA(cudaEventRecord(e->ev_fork, st));
  ZB_CHECK_CUDA(cudaStreamWaitEvent(e->side, e->ev_fork, 0));
  l2_prefetch_kernel<<<16, 128, 0, e->side>>>((const uint8_t*)w, (long long)bytes);
  ZB_CHECK_LAUNCH("l2_prefetch_kernel");
  ++e->launches;
  e->forked = true;
  return ZB200_OK;
}
int join_side(zb200_engine* e, cudaStream_t st) {
  if (!e->forked) return ZB200_OK;
  ZB_CHECK_CUDA(cudaEventRecord(e->ev_join, e->side));
  ZB_CHECK_CUDA(cudaStreamWaitEvent(st, e->ev_join, 0));
  e->forked = false;
  return ZB200_OK;
}
#define ZB_TRY_RC(expr)            \
  do {                             \
    const int _rc = (expr);        \
    if (_rc != ZB200_OK) return _rc; \
  } while (0)

template <class ALoader, class Epilogue>
cudaError_t gemm_swap(const zb200_engine* e, const void* w, int n_tiles, int rows, int k, const ALoader& al,
                      const Epilogue& epi, cudaStream_t st, const NextW& nx = NextW{}) {
  SwapCore g{};
  g.Wp = w, g.n_tiles = n_tiles, g.k_blocks = k / 64, g.m_valid = rows;
  g.partial_ws = e->d.gemm_ws, g.tickets = e->d.gemm_tickets, g.trace = nullptr, g.trace_id = -1;
  g.next = nx;
  return launch_gemm_swap(g, al, epi, swap_grid(e, n_tiles, k), st);
}

// Decode step with rows = 2B <= 128: every projection streams its weights through gemm_skinny; out_proj and fc2 leave
// their split-K partial sums to the fused residual + LayerNorm kernel, which also produces the next GEMM's input.
int run_decode_skinny(zb200_engine* e, cudaStream_t st) {
  const zb200_engine_desc& d = e->d;
  const int dm = d.d_model, hd = d.n_heads * ZB200_HEAD_DIM, kvd = d.n_kv_heads * ZB200_HEAD_DIM;
  const int rows = 2 * d.batch, n_seq = rows;
  const int* pos_ptr = d.state + ST_POS;
  bf16 *x = (bf16*)d.x, *xn = (bf16*)d.xn, *q = (bf16*)d.q, *att = (bf16*)d.att, *h = (bf16*)d.h;
  // at <= 8 rows selected projections run on the operand-swapped kernel (bit 0 in_proj, 2 fc1, 3 heads)
  static const int swap_env = zb::tuning_env("ZB200_SWAP_MASK") ? atoi(zb::tuning_env("ZB200_SWAP_MASK")) : 13;
  const int swap_mask = (rows <= 8 && dm % 128 == 0) ? swap_env : 0;
  static const bool no_qkv_def = zb::tuning_env("ZB200_NO_QKV_DEFERRED") != nullptr;
  const bool qkv_deferred = rows > 32 && !no_qkv_def && (hd + 2 * kvd) <= 8192;
  const SkPlan p_qkv = plan_deferred(e->sk_sms, rows, (hd + 2 * kvd) / NT, dm / 64, d.gemm_ws_floats, hd + 2 * kvd);
  const SkPlan p_out = plan_deferred(e->sk_sms, rows, dm / NT, hd / 64, d.gemm_ws_floats, dm);
  const SkPlan p_fc2 = plan_deferred(e->sk_sms, rows, dm / NT, d.d_ff / 64, d.gemm_ws_floats, dm);
  // what follows each GEMM in the step (for the L2 prefetch of its first weight bytes): out_proj -> fc1 -> fc2 -> next
  // layer's in_proj (heads after the last layer).  in_proj -> out_proj is left out: the attention kernel in between
  // streams the KV pages through L2.
  const int qkv_tiles = (hd + 2 * kvd) / NT, fc1_tiles = d.d_ff / (NT / 2), head_tiles = d.n_q * (d.head_rows_padded / NT);
  auto nx_fc1 = [&](int l) {
    const void* w = (const uint8_t*)d.w_fc1 + (size_t)l * d.fc1_stride;
    if (swap_mask & 4) return next_swap(e, w, rows, fc1_tiles, dm);
    const SkPlan p = plan_ticketed(e->sk_sms, rows, fc1_tiles, dm / 64);
    return next_skinny(w, rows, fc1_tiles, dm, p.grid, p.prefer_k1);
  };
  auto nx_fc2 = [&](int l) { return next_skinny((const uint8_t*)d.w_fc2 + (size_t)l * d.fc2_stride, rows, dm / NT, d.d_ff, p_fc2.grid, false); };
  auto nx_qkv = [&](int l) {
    if (l >= d.n_layer) {  // heads
      if (swap_mask & 8) return next_swap(e, d.w_heads, rows, head_tiles, dm);
      const SkPlan p = plan_ticketed(e->sk_sms, rows, head_tiles, dm / 64);
      return next_skinny(d.w_heads, rows, head_tiles, dm, p.grid, p.prefer_k1);
    }
    const void* w = (const uint8_t*)d.w_qkv + (size_t)l * d.qkv_stride;
    if (swap_mask & 1) return next_swap(e, w, rows, qkv_tiles, dm);
    if (qkv_deferred) return next_skinny(w, rows, qkv_tiles, dm, p_qkv.grid, false);
    const SkPlan p = plan_ticketed(e->sk_sms, rows, qkv_tiles, dm / 64);
    return next_skinny(w, rows, qkv_tiles, dm, p.grid, p.prefer_k1);
  };
  // KV pages of the next attention kernel pulled into L2 from the tails of the GEMMs in front of it (gemm_skinny.cuh):
  // pages per CTA for out_proj, fc1, fc2 (layer l + 1's pages) and in_proj (its own layer's).  Measured at batch 64
  // (ms per frame): none 2.97, {1,0,1,0} 2.86, {0,1,1,1} 2.81 (54 MB of the layer's 376 MB), {2,0,2,1} 2.86, {3,0,3,2}
  // 2.91 (141 MB: more than L2 keeps); batch 32: 2.05 -> 1.91.
  static int kvpf_n[4] = {-1, 1, 1, 1};
  if (kvpf_n[0] < 0) {
    kvpf_n[0] = 0;
    if (const char* v = zb::tuning_env("ZB200_KVPF")) sscanf(v, "%d,%d,%d,%d", &kvpf_n[0], &kvpf_n[1], &kvpf_n[2], &kvpf_n[3]);
  }
  const int kv_page_bytes = 2 * d.n_kv_heads * 64 * ZB200_HEAD_DIM * 2;
  const int fc1_grid = plan_ticketed(e->sk_sms, rows, d.d_ff / (NT / 2), dm / 64).grid;
  const int kvpf_grid[4] = {p_out.grid, fc1_grid, p_fc2.grid, p_qkv.grid};
  auto kvpf = [&](int which, int layer, int grid) {
    KvPrefetch k{};
    if (kvpf_n[which] <= 0 || rows <= 32) return k;  // (<= 32 rows: two CTAs per SM already keep HBM busy; measured slower)
    int first = 0;
    for (int j = 0; j < which; ++j) first += kvpf_n[j] * kvpf_grid[j];
    (void)grid;
    k.layer_base = (const uint8_t*)d.kv_pool + (size_t)(layer % d.n_layer) * d.num_pages * kv_page_bytes;
    k.block_table = d.block_table, k.pos_ptr = pos_ptr, k.pos_add = 0, k.pages_per_seq = d.pages_per_seq, k.n_seq = n_seq;
    k.page_bytes = kv_page_bytes, k.first_item = first, k.items_per_cta = kvpf_n[which];
    return k;
  };
  ZB_TRY(launch_layernorm(x, (const bf16*)d.ln1_w, (const bf16*)d.ln1_b, xn, rows, dm, d.ln_eps, 1, 0, st));
  for (int l = 0; l < d.n_layer; ++l) {
    {  // in_proj + RoPE + KV page write
      ZB_TRY_RC(prefetch_l2(e, (const uint8_t*)d.w_out + (size_t)l * d.out_stride, d.out_stride, st));
      ARowMajor al{(const uint8_t*)xn, (size_t)dm * 2};
      if (swap_mask & 1) {
        EpiTQkvRope epi{q, kv_of(d, l), d.rope, pos_ptr, d.n_heads};
        ZB_TRY(gemm_swap(e, (const uint8_t*)d.w_qkv + (size_t)l * d.qkv_stride, (hd + 2 * kvd) / NT, rows, dm, al, epi, st));
      } else if (qkv_deferred) {  // split-K partials, finished (sum + RoPE + q / KV write) by a row x column kernel
        SkinnyCore g = sk_core(d, (const uint8_t*)d.w_qkv + (size_t)l * d.qkv_stride, (hd + 2 * kvd) / NT, rows, dm);
        g.n_splits = p_qkv.splits;
        g.kvpf = kvpf(3, l, p_qkv.grid);
        EpiPartial epi{d.gemm_ws, rows, hd + 2 * kvd};
        ZB_TRY(launch_gemm_skinny(g, al, epi, p_qkv.grid, st));
        ZB_TRY(launch_qkv_finish(d.gemm_ws, p_qkv.splits, rows, hd + 2 * kvd, q, kv_of(d, l), d.rope, pos_ptr, d.n_heads, st));
      } else {
        EpiQkvRope epi{q, kv_of(d, l), d.rope, pos_ptr, 0, 1, d.n_heads};
        ZB_TRY(gemm_any(e, (const uint8_t*)d.w_qkv + (size_t)l * d.qkv_stride, (hd + 2 * kvd) / NT, rows, dm, al, epi, st));
      }
    }
    {
      AttnDecodeArgs a;
      a.q = q, a.out = att, a.kv = kv_of(d, l), a.pos_ptr = pos_ptr, a.pos_add = 0;
      a.n_heads = d.n_heads, a.n_kv_heads = d.n_kv_heads, a.scale = 1.0f / sqrtf((float)ZB200_HEAD_DIM);
      a.partial_ws = d.attn_ws, a.tickets = d.attn_tickets;
      ZB_TRY(launch_attn_decode(a, n_seq, d.attn_nsplit, st));
    }
    {  // out_proj (split-K partials) -> residual + norm2
      ZB_TRY_RC(prefetch_l2(e, (const uint8_t*)d.w_fc1 + (size_t)l * d.fc1_stride, d.fc1_stride, st));
      SkinnyCore g = sk_core(d, (const uint8_t*)d.w_out + (size_t)l * d.out_stride, dm / NT, rows, hd);
      g.n_splits = p_out.splits;
      g.next = nx_fc1(l);
      g.kvpf = kvpf(0, l + 1, p_out.grid);
      ARowMajor al{(const uint8_t*)att, (size_t)hd * 2};
      EpiPartial epi{d.gemm_ws, rows, dm};
      ZB_TRY(launch_gemm_skinny(g, al, epi, p_out.grid, st));
      ZB_TRY(launch_resid_ln(x, d.gemm_ws, p_out.splits, rows, (const bf16*)d.ln2_w + (size_t)l * dm,
                             (const bf16*)d.ln2_b + (size_t)l * dm, xn, dm, d.ln_eps, st));
    }
    {  // fc1 + SiLU gate
      ZB_TRY_RC(prefetch_l2(e, (const uint8_t*)d.w_fc2 + (size_t)l * d.fc2_stride, d.fc2_stride, st));
      ARowMajor al{(const uint8_t*)xn, (size_t)dm * 2};
      if (swap_mask & 4) {
        EpiTSwiGLU epi{h, d.d_ff};
        ZB_TRY(gemm_swap(e, (const uint8_t*)d.w_fc1 + (size_t)l * d.fc1_stride, d.d_ff / (NT / 2), rows, dm, al, epi, st, nx_fc2(l)));
      } else {
        EpiSwiGLU<NT> epi{h, d.d_ff};
        ZB_TRY(gemm_any(e, (const uint8_t*)d.w_fc1 + (size_t)l * d.fc1_stride, d.d_ff / (NT / 2), rows, dm, al, epi, st, nx_fc2(l),
                        kvpf(1, l + 1, fc1_grid)));
      }
    }
    {  // fc2 (split-K partials) -> residual + the next layer's norm (norm_f after the last layer)
      if (l + 1 < d.n_layer)
        ZB_TRY_RC(prefetch_l2(e, (const uint8_t*)d.w_qkv + (size_t)(l + 1) * d.qkv_stride, d.qkv_stride, st));
      else
        ZB_TRY_RC(prefetch_l2(e, d.w_heads, (int64_t)d.n_q * d.head_rows_padded * dm * 2, st));
      SkinnyCore g = sk_core(d, (const uint8_t*)d.w_fc2 + (size_t)l * d.fc2_stride, dm / NT, rows, d.d_ff);
      g.n_splits = p_fc2.splits;
      g.next = nx_qkv(l + 1);
      g.kvpf = kvpf(2, l + 1, p_fc2.grid);
      ARowMajor al{(const uint8_t*)h, (size_t)d.d_ff * 2};
      EpiPartial epi{d.gemm_ws, rows, dm};
      ZB_TRY(launch_gemm_skinny(g, al, epi, p_fc2.grid, st));
      const bool last = (l + 1 == d.n_layer);
      const bf16* nw = last ? (const bf16*)d.lnf_w : (const bf16*)d.ln1_w + (size_t)(l + 1) * dm;
      const bf16* nb = last ? (const bf16*)d.lnf_b : (const bf16*)d.ln1_b + (size_t)(l + 1) * dm;
      ZB_TRY(launch_resid_ln(x, d.gemm_ws, p_fc2.splits, rows, nw, nb, xn, dm, d.ln_eps, st));
    }
  }
  {  // heads + CFG on the normalised rows (model.py:110-115)
    ZB_TRY_RC(prefetch_l2(e, d.w_qkv, d.qkv_stride, st));  // the next frame starts with layer 0's in_proj
    const int tiles_per_head = d.head_rows_padded / NT;
    if (swap_mask & 8) {
      ARowMajor al{(const uint8_t*)xn, (size_t)dm * 2};
      EpiTHeadsCfg epi{d.logits, d.batch, d.n_q, d.vocab, tiles_per_head, d.cfg_scale};
      ZB_TRY(gemm_swap(e, d.w_heads, d.n_q * tiles_per_head, n_seq, dm, al, epi, st));
    } else {
      ARowCfgPairs al{(const uint8_t*)xn, (size_t)dm * 2, d.batch};
      EpiHeadsCfg<NT> epi{d.logits, d.batch, d.n_q, d.vocab, tiles_per_head, d.cfg_scale};
      ZB_TRY(gemm_any(e, d.w_heads, d.n_q * tiles_per_head, n_seq, dm, al, epi, st));
    }
  }
  return join_side(e, st);
}

// Decode step with rows = 2B <= 128 on the fused chain kernel (chain.cuh): one launch for everything between two
// attention kernels.  Per frame: embedding sum, LayerNorm, [in_proj 0], then per layer attention + one chain launch
// (out_proj, residual + norm2, fc1 + SwiGLU, fc2, residual + next norm, next in_proj -- heads + CFG after the last layer).
// The chain launches' grid-barrier counters live in the upper half of gemm_tickets (zero-initialised by the shim): launch
// j uses slot j and clears slot j - 1 (mod the launches of a frame), whose launch has completed by then.
constexpr int CHAIN_SYNC_BASE = 512;  // ints: [0, 512) of gemm_tickets stay the ticketed phases' tile-pair tickets
int run_decode_chain(zb200_engine* e, cudaStream_t st) {
  const zb200_engine_desc& d = e->d;
  const int dm = d.d_model, hd = d.n_heads * ZB200_HEAD_DIM, kvd = d.n_kv_heads * ZB200_HEAD_DIM;
  const int rows = 2 * d.batch, sms = e->sk_sms, L = d.n_layer;
  const int n_launches = L + 1;
  const int* pos_ptr = d.state + ST_POS;
  bf16 *x = (bf16*)d.x, *xn = (bf16*)d.xn, *q = (bf16*)d.q, *att = (bf16*)d.att, *h = (bf16*)d.h;

  ChainArgs base{};
  base.rows = rows, base.ws = d.gemm_ws, base.tickets = d.gemm_tickets, base.err = d.state + ST_ERROR;
  base.x = x, base.xn = xn, base.q = q, base.d_model = dm, base.ln_eps = d.ln_eps;
  base.kv = kv_of(d, 0), base.rope = d.rope, base.pos_ptr = pos_ptr, base.n_heads = d.n_heads;
  base.logits = d.logits, base.batch = d.batch, base.n_q = d.n_q, base.vocab = d.vocab;
  base.tiles_per_head = d.head_rows_padded / NT, base.cfg_scale = d.cfg_scale, base.trace = e->chain_trace;

  auto sync_slot = [&](int j) { return d.gemm_tickets + CHAIN_SYNC_BASE + ((j + n_launches) % n_launches) * CH_SYNC_INTS; };
  auto ph_qkv = [&](int l) {
    ChainPhase p{};
    p.W = (const uint8_t*)d.w_qkv + (size_t)l * d.qkv_stride, p.n_tiles = (hd + 2 * kvd) / NT, p.k_blocks = dm / 64;
    p.epi = CH_EPI_PARTIAL, p.fin = CH_FIN_QKV, p.A = (const uint8_t*)xn, p.a_row_bytes = dm * 2, p.ldn = hd + 2 * kvd;
    p.kv_layer = l;
    chain_plan_partial(p, sms, rows, d.gemm_ws_floats);
    return p;
  };
  auto ph_resid = [&](const void* w, int64_t stride, int l, const bf16* a_ptr, int k, const bf16* lnw, const bf16* lnb) {
    ChainPhase p{};
    p.W = (const uint8_t*)w + (size_t)l * stride, p.n_tiles = dm / NT, p.k_blocks = k / 64;
    p.epi = CH_EPI_PARTIAL, p.fin = CH_FIN_RESID_LN, p.A = (const uint8_t*)a_ptr, p.a_row_bytes = k * 2, p.ldn = dm;
    p.ln_w = lnw, p.ln_b = lnb;
    chain_plan_partial(p, sms, rows, d.gemm_ws_floats);
    return p;
  };
  auto ph_fc1 = [&](int l) {
    ChainPhase p{};
    p.W = (const uint8_t*)d.w_fc1 + (size_t)l * d.fc1_stride, p.n_tiles = d.d_ff / (NT / 2), p.k_blocks = dm / 64;
    p.epi = CH_EPI_SWIGLU, p.fin = CH_FIN_NONE, p.A = (const uint8_t*)xn, p.a_row_bytes = dm * 2, p.h = h, p.ldh = d.d_ff;
    chain_plan_ticketed(p, sms);
    return p;
  };
  auto ph_heads = [&]() {
    ChainPhase p{};
    p.W = (const uint8_t*)d.w_heads, p.n_tiles = d.n_q * (d.head_rows_padded / NT), p.k_blocks = dm / 64;
    p.epi = CH_EPI_HEADS, p.fin = CH_FIN_NONE, p.a_pairs = 1, p.A = (const uint8_t*)xn, p.a_row_bytes = dm * 2;
    chain_plan_ticketed(p, sms);
    return p;
  };

  ZB_TRY(launch_layernorm(x, (const bf16*)d.ln1_w, (const bf16*)d.ln1_b, xn, rows, dm, d.ln_eps, 1, 0, st));
  {
    ChainArgs a = base;
    a.n_phases = 1, a.ph[0] = ph_qkv(0);
    a.sync_mine = sync_slot(0), a.sync_prev = sync_slot(-1);
    ZB_TRY(launch_chain(a, sms, st));
  }
  auto trace_slot = [&](int j) { return e->chain_trace ? e->chain_trace + (size_t)j * sms * CH_MAX_PHASES * CH_TRACE_EVENTS : nullptr; };
  {
  }
  for (int l = 0; l < L; ++l) {
    {
      AttnDecodeArgs at;
      at.q = q, at.out = att, at.kv = kv_of(d, l), at.pos_ptr = pos_ptr, at.pos_add = 0;
      at.n_heads = d.n_heads, at.n_kv_heads = d.n_kv_heads, at.scale = 1.0f / sqrtf((float)ZB200_HEAD_DIM);
      at.partial_ws = d.attn_ws, at.tickets = d.attn_tickets;
      ZB_TRY(launch_attn_decode(at, rows, d.attn_nsplit, st));
    }
    const bool last = (l + 1 == L);
    ChainArgs a = base;
    a.n_phases = 4;
    a.ph[0] = ph_resid(d.w_out, d.out_stride, l, att, hd, (const bf16*)d.ln2_w + (size_t)l * dm, (const bf16*)d.ln2_b + (size_t)l * dm);
    a.ph[1] = ph_fc1(l);
    a.ph[2] = ph_resid(d.w_fc2, d.fc2_stride, l, h, d.d_ff, last ? (const bf16*)d.lnf_w : (const bf16*)d.ln1_w + (size_t)(l + 1) * dm,
                       last ? (const bf16*)d.lnf_b : (const bf16*)d.ln1_b + (size_t)(l + 1) * dm);
    a.ph[3] = last ? ph_heads() : ph_qkv(l + 1);
    a.sync_mine = sync_slot(l + 1), a.sync_prev = sync_slot(l);
    a.trace = trace_slot(l + 1);
    ZB_TRY(launch_chain(a, sms, st));
  }
  return ZB200_OK;
}

// Shapes the chain kernel takes: even tile counts for the deferred phases, counters and tickets inside gemm_tickets,
// at least one split of every deferred phase inside the workspace.
bool chain_supported(const zb200_engine* e) {
  const zb200_engine_desc& d = e->d;
  const int dm = d.d_model, hd = d.n_heads * ZB200_HEAD_DIM, kvd = d.n_kv_heads * ZB200_HEAD_DIM, rows = 2 * d.batch;
  if (!(d.flags & ZB200_ENGINE_CHAIN)) return false;
  if (rows > GEMM_BM || e->sk_sms < 8) return false;
  if ((dm / NT) % 2 || ((hd + 2 * kvd) / NT) % 2 || dm > 4096) return false;
  if (d.n_gemm_tickets < CHAIN_SYNC_BASE + (d.n_layer + 1) * CH_SYNC_INTS) return false;
  const int max_pairs = std::max(d.d_ff / (NT / 2), d.n_q * (d.head_rows_padded / NT)) / 2 + 1;
  if (max_pairs > CHAIN_SYNC_BASE) return false;
  if ((int64_t)rows * std::max(dm, hd + 2 * kvd) > d.gemm_ws_floats) return false;
  if ((int64_t)e->sk_sms * 4 * SK_PARTIAL_F4 * 4 > d.gemm_ws_floats) return false;  // fix-up slots: 2 x 2 sub-tiles per CTA
  return true;
}

// Decode step with rows = 2B <= 8 (gemm_swap.cuh): the weights are the 128-row UMMA operand, thread == output feature.
// Each group of two n_tiles is shared by at most 8 CTAs whose 4 KB partials the last one to arrive sums in one batch of
// loads, so RoPE / residual / SwiGLU / CFG epilogues run in the GEMM itself and no partial-sum pass exists.
int run_decode_small(zb200_engine* e, cudaStream_t st) {
  const zb200_engine_desc& d = e->d;
  const int dm = d.d_model, hd = d.n_heads * ZB200_HEAD_DIM, kvd = d.n_kv_heads * ZB200_HEAD_DIM;
  const int rows = 2 * d.batch;
  const int* pos_ptr = d.state + ST_POS;
  bf16 *x = (bf16*)d.x, *xn = (bf16*)d.xn, *q = (bf16*)d.q, *att = (bf16*)d.att, *h = (bf16*)d.h;
  for (int l = 0; l < d.n_layer; ++l) {
    ZB_TRY(launch_layernorm(x, (const bf16*)d.ln1_w + (size_t)l * dm, (const bf16*)d.ln1_b + (size_t)l * dm, xn, rows, dm,
                            d.ln_eps, 1, 0, st));
    {  // in_proj + RoPE + KV page write
      ARowMajor al{(const uint8_t*)xn, (size_t)dm * 2};
      EpiTQkvRope epi{q, kv_of(d, l), d.rope, pos_ptr, d.n_heads};
      ZB_TRY(gemm_swap(e, (const uint8_t*)d.w_qkv + (size_t)l * d.qkv_stride, (hd + 2 * kvd) / NT, rows, dm, al, epi, st));
    }
    {
      AttnDecodeArgs a;
      a.q = q, a.out = att, a.kv = kv_of(d, l), a.pos_ptr = pos_ptr, a.pos_add = 0;
      a.n_heads = d.n_heads, a.n_kv_heads = d.n_kv_heads, a.scale = 1.0f / sqrtf((float)ZB200_HEAD_DIM);
      a.partial_ws = d.attn_ws, a.tickets = d.attn_tickets;
      ZB_TRY(launch_attn_decode(a, rows, d.attn_nsplit, st));
    }
    {  // out_proj + residual
      ARowMajor al{(const uint8_t*)att, (size_t)hd * 2};
      EpiTResidual epi{x, dm};
      ZB_TRY(gemm_swap(e, (const uint8_t*)d.w_out + (size_t)l * d.out_stride, dm / NT, rows, hd, al, epi, st));
    }
    ZB_TRY(launch_layernorm(x, (const bf16*)d.ln2_w + (size_t)l * dm, (const bf16*)d.ln2_b + (size_t)l * dm, xn, rows, dm,
                            d.ln_eps, 1, 0, st));
    {  // fc1 + SiLU gate
      ARowMajor al{(const uint8_t*)xn, (size_t)dm * 2};
      EpiTSwiGLU epi{h, d.d_ff};
      ZB_TRY(gemm_swap(e, (const uint8_t*)d.w_fc1 + (size_t)l * d.fc1_stride, d.d_ff / (NT / 2), rows, dm, al, epi, st));
    }
    {  // fc2 + residual
      ARowMajor al{(const uint8_t*)h, (size_t)d.d_ff * 2};
      EpiTResidual epi{x, dm};
      ZB_TRY(gemm_swap(e, (const uint8_t*)d.w_fc2 + (size_t)l * d.fc2_stride, dm / NT, rows, d.d_ff, al, epi, st));
    }
  }
  ZB_TRY(launch_layernorm(x, (const bf16*)d.lnf_w, (const bf16*)d.lnf_b, xn, rows, dm, d.ln_eps, 1, 0, st));
  {  // heads + CFG on the normalised rows (model.py:110-115)
    const int tiles_per_head = d.head_rows_padded / NT;
    ARowMajor al{(const uint8_t*)xn, (size_t)dm * 2};
    EpiTHeadsCfg epi{d.logits, d.batch, d.n_q, d.vocab, tiles_per_head, d.cfg_scale};
    ZB_TRY(gemm_swap(e, d.w_heads, d.n_q * tiles_per_head, rows, dm, al, epi, st));
  }
  return ZB200_OK;
}

// All transformer layers on `rows` rows of x (seq_len positions per sequence), then norm_f + heads + CFG.
// decode: one query per sequence at position *state[ST_POS] (+ pos0); otherwise seq_len positions from pos0 (causal).
// hidden_out != nullptr: the backbone-only form (TorchZonosBackbone.forward, _torch.py:73-80) -- norm_f over ALL rows is
// written there and the heads are skipped.
int run_backbone(zb200_engine* e, int rows, int seq_len, bool decode, cudaStream_t st, int pos0 = 0, bf16* hidden_out = nullptr,
                 bool pos_from_state = true) {
  const zb200_engine_desc& d = e->d;
  const int dm = d.d_model, hd = d.n_heads * ZB200_HEAD_DIM, kvd = d.n_kv_heads * ZB200_HEAD_DIM;
  const int m_tiles = ceil_div(rows, GEMM_BM);
  const int n_seq = 2 * d.batch;
  const int* pos_ptr = (decode && pos_from_state) ? d.state + ST_POS : nullptr;
  bf16 *x = (bf16*)d.x, *xn = (bf16*)d.xn, *q = (bf16*)d.q, *att = (bf16*)d.att, *h = (bf16*)d.h;
  for (int l = 0; l < d.n_layer; ++l) {
    const bf16* ln1w = (const bf16*)d.ln1_w + (size_t)l * dm;
    const bf16* ln1b = (const bf16*)d.ln1_b + (size_t)l * dm;
    const bf16* ln2w = (const bf16*)d.ln2_w + (size_t)l * dm;
    const bf16* ln2b = (const bf16*)d.ln2_b + (size_t)l * dm;
    ZB_TRY(launch_layernorm(x, ln1w, ln1b, xn, rows, dm, d.ln_eps, 1, 0, st));
    {  // in_proj + RoPE + KV page write
      const int n_tiles = (hd + 2 * kvd) / NT;
      ARowMajor al{(const uint8_t*)xn, (size_t)dm * 2};
      EpiQkvRope epi{q, kv_of(d, l), d.rope, pos_ptr, pos0, seq_len, d.n_heads};
      ZB_TRY(gemm_any(e, (const uint8_t*)d.w_qkv + (size_t)l * d.qkv_stride, n_tiles, rows, dm, al, epi, st));
    }
    if (decode) {
      AttnDecodeArgs a;
      a.q = q, a.out = att, a.kv = kv_of(d, l), a.pos_ptr = pos_ptr, a.pos_add = pos0;
      a.n_heads = d.n_heads, a.n_kv_heads = d.n_kv_heads, a.scale = 1.0f / sqrtf((float)ZB200_HEAD_DIM);
      a.partial_ws = d.attn_ws, a.tickets = d.attn_tickets;
      ZB_TRY(launch_attn_decode(a, n_seq, d.attn_nsplit, st));
    } else {
      AttnPrefillArgs a;
      a.q = q, a.out = att, a.kv = kv_of(d, l), a.seq_len = seq_len, a.pos0 = pos0, a.n_heads = d.n_heads;
      a.scale = 1.0f / sqrtf((float)ZB200_HEAD_DIM);
      ZB_TRY(launch_attn_prefill(a, n_seq, st));
    }
    {  // out_proj + residual
      const int n_tiles = dm / NT;
      ARowMajor al{(const uint8_t*)att, (size_t)hd * 2};
      EpiResidual<NT> epi{x, dm};
      ZB_TRY(gemm_any(e, (const uint8_t*)d.w_out + (size_t)l * d.out_stride, n_tiles, rows, hd, al, epi, st));
    }
    ZB_TRY(launch_layernorm(x, ln2w, ln2b, xn, rows, dm, d.ln_eps, 1, 0, st));
    {  // fc1 + SiLU gate
      const int n_tiles = d.d_ff / (NT / 2);
      ARowMajor al{(const uint8_t*)xn, (size_t)dm * 2};
      EpiSwiGLU<NT> epi{h, d.d_ff};
      ZB_TRY(gemm_any(e, (const uint8_t*)d.w_fc1 + (size_t)l * d.fc1_stride, n_tiles, rows, dm, al, epi, st));
    }
    {  // fc2 + residual
      const int n_tiles = dm / NT;
      ARowMajor al{(const uint8_t*)h, (size_t)d.d_ff * 2};
      EpiResidual<NT> epi{x, dm};
      ZB_TRY(gemm_any(e, (const uint8_t*)d.w_fc2 + (size_t)l * d.fc2_stride, n_tiles, rows, d.d_ff, al, epi, st));
    }
  }
  if (hidden_out) {  // backbone-only: norm_f over every row, no heads
    ZB_TRY(launch_layernorm(x, (const bf16*)d.lnf_w, (const bf16*)d.lnf_b, hidden_out, rows, dm, d.ln_eps, 1, 0, st));
    return ZB200_OK;
  }
  // norm_f on the last position of every sequence, heads + CFG (model.py:110-115)
  ZB_TRY(launch_layernorm(x, (const bf16*)d.lnf_w, (const bf16*)d.lnf_b, xn, n_seq, dm, d.ln_eps, seq_len, seq_len - 1, st));
  {
    const int tiles_per_head = d.head_rows_padded / NT;
    const int n_tiles = d.n_q * tiles_per_head;
    ARowCfgPairs al{(const uint8_t*)xn, (size_t)dm * 2, d.batch};
    EpiHeadsCfg<NT> epi{d.logits, d.batch, d.n_q, d.vocab, tiles_per_head, d.cfg_scale};
    ZB_TRY(gemm_any(e, d.w_heads, n_tiles, n_seq, dm, al, epi, st));
  }
  return ZB200_OK;
}

}  // namespace

static int validate(const zb200_engine_desc& d) {
  ZB_REQUIRE(d.d_model > 0 && d.d_model % 256 == 0 && d.d_model <= 4096, "engine: d_model=%d must be a multiple of 256 (<= 4096)", d.d_model);
  ZB_REQUIRE(d.n_heads > 0 && d.n_kv_heads > 0 && d.n_heads % d.n_kv_heads == 0 && d.n_heads / d.n_kv_heads <= 4,
             "engine: unsupported head layout H=%d Hkv=%d (GQA group <= 4)", d.n_heads, d.n_kv_heads);
  ZB_REQUIRE(d.d_model == d.n_heads * ZB200_HEAD_DIM, "engine: head_dim must be 128 (d_model=%d, H=%d)", d.d_model, d.n_heads);
  ZB_REQUIRE(d.d_ff > 0 && d.d_ff % 64 == 0, "engine: d_ff=%d must be a multiple of 64", d.d_ff);
  const bool bb_only = (d.flags & ZB200_ENGINE_BACKBONE_ONLY) != 0;
  ZB_REQUIRE(bb_only || (d.n_q > 0 && d.n_q <= ZB200_MAX_CODEBOOKS && d.vocab > 1025 && d.vocab <= d.head_rows_padded &&
                         d.head_rows_padded % 64 == 0),
             "engine: bad codebook/vocab layout");
  ZB_REQUIRE(d.batch > 0 && d.rows_max >= 2 * d.batch, "engine: rows_max=%d < 2*batch", d.rows_max);
  ZB_REQUIRE(d.attn_nsplit >= 1 && d.pages_per_seq > 0 && d.num_pages > 0, "engine: bad KV paging");
  ZB_REQUIRE(d.ln1_w && d.ln1_b && d.ln2_w && d.ln2_b && d.lnf_w && d.lnf_b && d.w_qkv && d.w_out && d.w_fc1 && d.w_fc2 &&
                 d.rope && d.kv_pool && d.block_table && d.x && d.xn && d.q && d.att && d.h && d.gemm_ws && d.gemm_tickets,
             "engine: null buffer in descriptor");
  ZB_REQUIRE(bb_only || (d.emb && d.w_heads && d.logits && d.state && d.delayed),
             "engine: null embedding / head / logits / loop-state buffer (only a ZB200_ENGINE_BACKBONE_ONLY engine may omit them)");
  ZB_REQUIRE(d.attn_nsplit == 1 || (d.attn_ws && d.attn_tickets), "engine: split-KV attention needs workspace");
  return ZB200_OK;
}

extern "C" int zb200_engine_create(const zb200_engine_desc* desc, zb200_engine** out) {
  ZB_REQUIRE(desc && out, "engine_create: null argument");
  const int rc = validate(*desc);
  if (rc != ZB200_OK) return rc;
  zb200_engine* e = new zb200_engine();
  e->d = *desc;
  {
    int sms = SM_COUNT, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (const char* v = zb::tuning_env("ZB200_SK_SMS")) sms = atoi(v);  // 0 switches the decode step back to the tiled kernel
    // fix-up workspace per SM: two CTAs x two slots x one 32 KB sub-tile (M <= 32) or one CTA x two slots x two sub-tiles
    const int64_t cap = desc->gemm_ws_floats / (4 * SK_PARTIAL_F4 * 4);
    e->sk_sms = (int)std::max<int64_t>(0, std::min<int64_t>(sms, cap));
    if (const char* v = zb::tuning_env("ZB200_L2_PREFETCH")) e->l2_prefetch = atoi(v);
    if (e->l2_prefetch) {
      cudaError_t ce = cudaStreamCreateWithFlags(&e->side, cudaStreamNonBlocking);
      if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming);
      if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming);
      if (ce != cudaSuccess) {
        zb200_engine_destroy(e);
        return cuda_fail(ce, "engine_create: side stream");
      }
    }
  }
  *out = e;
  return ZB200_OK;
}

extern "C" void zb200_engine_destroy(zb200_engine* e) {
  if (!e) return;
  if (e->exec) cudaGraphExecDestroy(e->exec);
  if (e->graph) cudaGraphDestroy(e->graph);
  if (e->ev_fork) cudaEventDestroy(e->ev_fork);
  if (e->ev_join) cudaEventDestroy(e->ev_join);
  if (e->side) cudaStreamDestroy(e->side);
  delete e;
}

/* TorchZonosBackbone.forward (zonos/backbone/_torch.py:73-80) on its own: x holds hidden [2B * seq_len, d_model]; the KV
 * cache is written at positions [pos0, pos0 + seq_len) and norm_f of every row goes to `out` (bf16, same shape). */
extern "C" int zb200_engine_forward(zb200_engine* e, int seq_len, int pos0, void* out, void* stream) {
  ZB_REQUIRE(e && seq_len > 0 && pos0 >= 0 && out, "engine_forward: bad arguments");
  const int rows = 2 * e->d.batch * seq_len;
  ZB_REQUIRE(rows <= e->d.rows_max, "engine_forward: %d rows exceed the workspace (%d)", rows, e->d.rows_max);
  if (pos0 + seq_len > e->d.pages_per_seq * ZB200_PAGE) {  // the reference asserts on cache overflow (_torch.py:44-45)
    set_error("engine_forward: positions up to %d exceed the KV cache (%d)", pos0 + seq_len, e->d.pages_per_seq * ZB200_PAGE);
    return ZB200_ERR_STATE;
  }
  e->launches = 0;
  return run_backbone(e, rows, seq_len, seq_len == 1, (cudaStream_t)stream, pos0, (bf16*)out, false);
}

extern "C" int zb200_engine_prefill(zb200_engine* e, int seq_len, void* stream) {
  ZB_REQUIRE(e && seq_len > 0, "engine_prefill: bad arguments");
  ZB_REQUIRE(!(e->d.flags & ZB200_ENGINE_BACKBONE_ONLY), "engine_prefill: backbone-only engine (use zb200_engine_forward)");
  const int rows = 2 * e->d.batch * seq_len;
  ZB_REQUIRE(rows <= e->d.rows_max, "engine_prefill: %d rows exceed the workspace (%d)", rows, e->d.rows_max);
  if (seq_len > e->d.pages_per_seq * ZB200_PAGE) {
    set_error("engine_prefill: sequence length %d exceeds the KV cache (%d)", seq_len, e->d.pages_per_seq * ZB200_PAGE);
    return ZB200_ERR_STATE;
  }
  e->launches = 0;
  return run_backbone(e, rows, seq_len, false, (cudaStream_t)stream);
}

static int decode_logits(zb200_engine* e, cudaStream_t st) {
  const zb200_engine_desc& d = e->d;
  // embed the previous frame (column offset-1) for the B conditional rows and duplicate for the unconditional ones
  ZB_TRY(launch_embed_sum(d.delayed, d.n_q * d.t_delayed, d.t_delayed, d.state + ST_OFFSET, -1, 1, (const bf16*)d.emb,
                          (bf16*)d.x, d.batch, d.n_q, d.vocab, d.d_model, 1, 0, d.batch, st));
  static const bool all_swap = zb::tuning_env("ZB200_ALL_SWAP") != nullptr;
  if (2 * d.batch <= 8 && e->sk_sms > 0 && all_swap && d.d_model % 128 == 0 && d.d_ff % 128 == 0)
    return run_decode_small(e, st);
  if (chain_supported(e)) return run_decode_chain(e, st);
  if (2 * d.batch <= GEMM_BM && e->sk_sms > 0) return run_decode_skinny(e, st);
  return run_backbone(e, 2 * d.batch, 1, true, st);
}

extern "C" int zb200_engine_decode_logits(zb200_engine* e, void* stream) {
  ZB_REQUIRE(e, "engine_decode_logits: null engine");
  ZB_REQUIRE(!(e->d.flags & ZB200_ENGINE_BACKBONE_ONLY), "engine_decode_logits: backbone-only engine (use zb200_engine_forward)");
  e->launches = 0;
  return decode_logits(e, (cudaStream_t)stream);
}

static SamplerArgs to_args(const zb200_sampler_args* u) {
  SamplerArgs a;
  a.logits = u->logits, a.noise = u->noise, a.delayed = u->delayed, a.tok1 = u->tok1, a.tok2 = u->tok2;
  a.state = u->state, a.remaining = u->remaining, a.grace = u->grace;
  a.stopping = u->stopping, a.eos_mode = u->eos_mode, a.mask2 = u->mask2, a.eos_old = u->eos_old;
  a.host_flags = u->host_flags;
  a.batch = u->batch, a.n_q = u->n_q, a.vocab = u->vocab, a.t_delayed = u->t_delayed;
  a.loop_mode = u->loop_mode, a.first_offset = u->first_offset, a.row_base = u->row_base;
  a.seed = u->seed, a.temperature = u->temperature, a.top_p = u->top_p, a.top_k = u->top_k, a.min_p = u->min_p;
  a.linear = u->linear, a.conf = u->conf, a.quad = u->quad, a.rep_penalty = u->rep_penalty, a.rep_window = u->rep_window;
  return a;
}

static int step(zb200_engine* e, const zb200_sampler_args* smp, cudaStream_t st) {
  e->launches = 0;
  const int rc = decode_logits(e, st);
  if (rc != ZB200_OK) return rc;
  const SamplerArgs a = to_args(smp);
  ZB_TRY(launch_sampler(a, 1, st));
  ZB_TRY(launch_sampler(a, 2, st));
  return ZB200_OK;
}

extern "C" int zb200_engine_step(zb200_engine* e, const zb200_sampler_args* smp, void* stream) {
  ZB_REQUIRE(e && smp && smp->loop_mode == 1, "engine_step: needs loop-mode sampler arguments");
  ZB_REQUIRE(!(e->d.flags & ZB200_ENGINE_BACKBONE_ONLY), "engine_step: backbone-only engine");
  ZB_REQUIRE(smp->state == e->d.state && smp->delayed == e->d.delayed && smp->logits == e->d.logits,
             "engine_step: sampler buffers differ from the engine's");
  return step(e, smp, (cudaStream_t)stream);
}

extern "C" int zb200_engine_capture(zb200_engine* e, const zb200_sampler_args* smp, void* stream) {
  ZB_REQUIRE(e && smp && smp->loop_mode == 1, "engine_capture: needs loop-mode sampler arguments");
  ZB_REQUIRE(!(e->d.flags & ZB200_ENGINE_BACKBONE_ONLY), "engine_capture: backbone-only engine");
  cudaStream_t st = (cudaStream_t)stream;
  if (e->exec) cudaGraphExecDestroy(e->exec), e->exec = nullptr;
  if (e->graph) cudaGraphDestroy(e->graph), e->graph = nullptr;
  ZB_CHECK_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
  const int rc = step(e, smp, st);
  cudaGraph_t g = nullptr;
  const cudaError_t ce = cudaStreamEndCapture(st, &g);
  if (rc != ZB200_OK) {
    if (g) cudaGraphDestroy(g);
    return rc;
  }
  if (ce != cudaSuccess) return cuda_fail(ce, "cudaStreamEndCapture");
  e->graph = g;
  ZB_CHECK_CUDA(cudaGraphInstantiate(&e->exec, e->graph, 0));
  return ZB200_OK;
}

extern "C" int zb200_engine_run(zb200_engine* e, int n_steps, void* stream) {
  ZB_REQUIRE(e && n_steps >= 0, "engine_run: bad arguments");
  if (!e->exec) {
    set_error("engine_run: no captured graph (call zb200_engine_capture first)");
    return ZB200_ERR_STATE;
  }
  for (int i = 0; i < n_steps; ++i) ZB_CHECK_CUDA(cudaGraphLaunch(e->exec, (cudaStream_t)stream));
  return ZB200_OK;
}

extern "C" int zb200_engine_launches_per_step(const zb200_engine* e) { return e ? e->launches : 0; }

/* debug: device buffer [n_layer + 1 launches][SMs][4 phases][8 events] uint64 receiving the chain kernel's in-kernel
 * globaltimer stamps; nullptr switches it off. */
extern "C" int zb200_engine_set_chain_trace(zb200_engine* e, unsigned long long* device_buf) {
  ZB_REQUIRE(e, "engine_set_chain_trace: null engine");
  e->chain_trace = device_buf;
  return ZB200_OK;
}
