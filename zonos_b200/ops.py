"""Thin torch-tensor wrappers over the single-kernel entry points of the C ABI (include/zonos_b200.h).

Used by the public API in zonos_b200/ (delay pattern, weight repack) and by the parity tests.  Nothing here computes
on the host: every function launches CUDA kernels of libzonos_b200.so on the current torch stream.
"""
import ctypes

import torch

from ._lib import check, lib, ptr, stream_ptr

TILE_N = 64
HEAD_ROWS_PADDED = 1088  # 1026 logits per head padded to 17 tiles of 64


def _cuda(t: torch.Tensor, dtype=None) -> torch.Tensor:
    if not t.is_cuda:
        raise ValueError("zonos_b200 kernels need CUDA tensors (there is no CPU fallback)")
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


def apply_delay_pattern(codes: torch.Tensor, mask_token: int) -> torch.Tensor:
    """zonos/codebook_pattern.py:5-7.  int64 [B,Q,T] -> [B,Q,T+Q]."""
    codes = _cuda(codes, torch.int64)
    B, Q, T = codes.shape
    out = torch.empty(B, Q, T + Q, dtype=torch.int64, device=codes.device)
    check(lib().zb200_delay_apply(ptr(codes), ptr(out), B, Q, T, ctypes.c_int64(mask_token), stream_ptr()), "delay_apply")
    return out


def revert_delay_pattern(delayed: torch.Tensor) -> torch.Tensor:
    """zonos/codebook_pattern.py:10-12.  int64 [B,Q,Tp] -> [B,Q,Tp-Q]."""
    delayed = _cuda(delayed, torch.int64)
    B, Q, Tp = delayed.shape
    out = torch.empty(B, Q, Tp - Q, dtype=torch.int64, device=delayed.device)
    check(lib().zb200_delay_revert(ptr(delayed), ptr(out), B, Q, Tp, stream_ptr()), "delay_revert")
    return out


def pack_weight(w: torch.Tensor, row_mode: int = 0, tile_n: int = TILE_N, rows_per_head: int = 0,
                rows_per_head_padded: int = 0) -> torch.Tensor:
    """Repack a 16-bit [N, K] (or [heads, rows, K] for row_mode 2) weight into per-stage shared-memory images."""
    w = _cuda(w)
    assert w.dtype in (torch.bfloat16, torch.float16)
    K = w.shape[-1]
    n_rows = w.numel() // K
    nbytes = lib().zb200_packed_weight_bytes(n_rows, K, tile_n, row_mode, rows_per_head, rows_per_head_padded)
    if nbytes <= 0:
        raise ValueError(f"cannot pack weight of shape {tuple(w.shape)} (K must be a multiple of 64)")
    out = torch.empty(nbytes // 2, dtype=w.dtype, device=w.device)
    check(lib().zb200_pack_weight(ptr(w), ptr(out), n_rows, K, tile_n, row_mode, rows_per_head, rows_per_head_padded,
                                  stream_ptr()), "pack_weight")
    return out


def gemm_bf16(a: torch.Tensor, w_packed: torch.Tensor, n: int, splits: int = 1) -> torch.Tensor:
    """Test hook: bf16 [M,K] x packed [N,K]^T -> bf16 [M,N] through the tcgen05 kernel."""
    a = _cuda(a, torch.bfloat16)
    M, K = a.shape
    out = torch.empty(M, n, dtype=torch.bfloat16, device=a.device)
    ws = tickets = None
    if splits > 1:
        tiles = ((n + 63) // 64) * ((M + 127) // 128)
        ws = torch.empty(tiles * splits * 64 * 128, dtype=torch.float32, device=a.device)
        tickets = torch.zeros(tiles, dtype=torch.int32, device=a.device)
    check(lib().zb200_gemm_bf16(ptr(a), M, K, ptr(w_packed), n, K, ptr(out), n, splits, ptr(ws), ptr(tickets),
                                stream_ptr()), "gemm_bf16")
    return out


def gemm_big_bf16(a: torch.Tensor, w_packed: torch.Tensor, n: int) -> torch.Tensor:
    """Test hook: bf16 [M,K] x packed [N,K]^T -> bf16 [M,N] through the persistent many-row kernel of the prefill."""
    a = _cuda(a, torch.bfloat16)
    M, K = a.shape
    out = torch.empty(M, n, dtype=torch.bfloat16, device=a.device)
    check(lib().zb200_gemm_big_bf16(ptr(a), M, K, ptr(w_packed), n, K, ptr(out), n, stream_ptr()), "gemm_big_bf16")
    return out


def gemm_skinny_bf16(a: torch.Tensor, w_packed: torch.Tensor, n: int, grid: int = 148, ws=None, tickets=None) -> torch.Tensor:
    """Test hook: bf16 [M<=128,K] x packed [N,K]^T -> bf16 [M,N] through the weight-streaming stream-K kernel."""
    a = _cuda(a, torch.bfloat16)
    M, K = a.shape
    n_tiles = (n + 63) // 64
    grid = min(grid, 1024)  # <= 0: the library plans the grid like the engine (needs ws for 296 CTAs)
    out = torch.empty(M, n, dtype=torch.bfloat16, device=a.device)
    if ws is None:
        ws = torch.empty((grid if grid > 0 else 296) * 32768, dtype=torch.float32, device=a.device)
    if tickets is None:
        tickets = torch.zeros(n_tiles, dtype=torch.int32, device=a.device)
    check(lib().zb200_gemm_skinny_bf16(ptr(a), M, K, ptr(w_packed), n, K, ptr(out), n, grid, ptr(ws),
                                       ctypes.c_int64(ws.numel()), ptr(tickets), tickets.numel(), stream_ptr()),
          "gemm_skinny_bf16")
    return out


def gemm_swap_bf16(a: torch.Tensor, w_packed: torch.Tensor, n: int, grid: int = 148) -> torch.Tensor:
    """Test hook: bf16 [M<=8,K] x packed [N,K]^T -> bf16 [M,N] through the operand-swapped weight-streaming kernel."""
    a = _cuda(a, torch.bfloat16)
    M, K = a.shape
    grid = max(1, min(grid, 1024))
    out = torch.empty(M, n, dtype=torch.bfloat16, device=a.device)
    ws = torch.full((grid * 2048,), float("nan"), dtype=torch.float32, device=a.device)
    tickets = torch.zeros((n + 127) // 128, dtype=torch.int32, device=a.device)
    check(lib().zb200_gemm_swap_bf16(ptr(a), M, K, ptr(w_packed), n, K, ptr(out), n, grid, ptr(ws),
                                     ctypes.c_int64(ws.numel()), ptr(tickets), tickets.numel(), stream_ptr()), "gemm_swap_bf16")
    assert int(tickets.abs().sum()) == 0
    return out


def gemm_resid_ln(a, w_packed, x, ln_w, ln_b, splits: int = 4, eps: float = 1e-5):
    """Test hook: returns (bf16(x + bf16(a @ w^T)), LayerNorm of that) through the deferred split-K GEMM + resid_ln kernel."""
    a, x = _cuda(a, torch.bfloat16), _cuda(x, torch.bfloat16).clone()
    M, K = a.shape
    n = x.shape[1]
    ws = torch.full((splits * M * n,), float("nan"), dtype=torch.float32, device=a.device)
    xn = torch.empty_like(x)
    check(lib().zb200_gemm_resid_ln(ptr(a), M, K, ptr(w_packed), n, K, splits, ptr(ws), ctypes.c_int64(ws.numel()), ptr(x),
                                    ptr(_cuda(ln_w, torch.bfloat16)), ptr(_cuda(ln_b, torch.bfloat16)), ptr(xn),
                                    ctypes.c_float(eps), stream_ptr()), "gemm_resid_ln")
    return x, xn


def layernorm(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, eps: float = 1e-5) -> torch.Tensor:
    x = _cuda(x, torch.bfloat16)
    d = x.shape[-1]
    out = torch.empty_like(x)
    check(lib().zb200_layernorm(ptr(x), ptr(_cuda(w, torch.bfloat16)), ptr(_cuda(b, torch.bfloat16)), ptr(out),
                                x.numel() // d, d, ctypes.c_float(eps), stream_ptr()), "layernorm")
    return out


def embed_codes(ids: torch.Tensor, emb: torch.Tensor) -> torch.Tensor:
    """Zonos.embed_codes (zonos/model.py:97-98): ids [B,Q,S] -> bf16 [B,S,d]; emb bf16 [Q,V,d]."""
    ids = _cuda(ids, torch.int32)
    B, Q, S = ids.shape
    _, V, d = emb.shape
    out = torch.empty(B, S, d, dtype=torch.bfloat16, device=ids.device)
    check(lib().zb200_embed_sum(ptr(ids), Q * S, S, S, ptr(_cuda(emb)), ptr(out), B, Q, V, d, S, 0, 0, stream_ptr()),
          "embed_sum")
    return out


class PagedKV:
    """Paged KV pool + block table (replaces InferenceParams.key_value_memory_dict, zonos/config.py:8-25).

    pool bf16 [n_layer, num_pages, 2, n_kv_heads, 64, 128]; block_table int32 [n_seq, pages_per_seq].  Pages are handed
    out from a free list so that a sequence's pages are in general NOT contiguous.
    """

    PAGE = 64

    def __init__(self, n_layer, n_seq, max_seq, n_kv_heads, device, shuffle_seed=None):
        self.n_layer, self.n_seq, self.n_kv_heads = n_layer, n_seq, n_kv_heads
        self.pages_per_seq = (max_seq + self.PAGE - 1) // self.PAGE
        self.num_pages = n_seq * self.pages_per_seq
        self.max_seq = self.pages_per_seq * self.PAGE
        # zero-filled: masked-out tail positions of a page are multiplied by P == 0 and must be finite
        self.pool = torch.zeros(n_layer, self.num_pages, 2, n_kv_heads, self.PAGE, 128, dtype=torch.bfloat16, device=device)
        if shuffle_seed is None:
            # interleave sequences so consecutive pages of one sequence are far apart (a realistic allocation order)
            ids = torch.arange(self.num_pages, dtype=torch.int32).view(self.pages_per_seq, n_seq).t().contiguous()
        else:
            g = torch.Generator().manual_seed(shuffle_seed)
            ids = torch.randperm(self.num_pages, generator=g).to(torch.int32).view(n_seq, self.pages_per_seq)
        self.block_table = ids.to(device)


def kv_write(kv: PagedKV, layer: int, k: torch.Tensor, v: torch.Tensor, pos0: int):
    """Test hook: scatter k, v bf16 [n_seq, S, n_kv, 128] into the pool at positions pos0.."""
    k, v = _cuda(k, torch.bfloat16), _cuda(v, torch.bfloat16)
    n_seq, S = k.shape[:2]
    check(lib().zb200_kv_write(ptr(k), ptr(v), ptr(kv.pool), ptr(kv.block_table), kv.pages_per_seq, kv.num_pages,
                               kv.n_kv_heads, layer, n_seq, S, pos0, stream_ptr()), "kv_write")


def attn_decode(q: torch.Tensor, kv: PagedKV, layer: int, n_keys: int, n_heads: int, nsplit: int = 1) -> torch.Tensor:
    q = _cuda(q, torch.bfloat16)
    n_seq = q.shape[0]
    out = torch.empty_like(q)
    ws = tickets = None
    if nsplit > 1:
        ws = torch.empty(n_seq * kv.n_kv_heads * nsplit * 520, dtype=torch.float32, device=q.device)
        tickets = torch.zeros(n_seq * kv.n_kv_heads, dtype=torch.int32, device=q.device)
    check(lib().zb200_attn_decode(ptr(q), ptr(kv.pool), ptr(kv.block_table), kv.pages_per_seq, kv.num_pages, n_heads,
                                  kv.n_kv_heads, layer, n_seq, n_keys, nsplit, ptr(ws), ptr(tickets), ptr(out),
                                  stream_ptr()), "attn_decode")
    return out


ATTN_NOALLOC_MAX_UNITS = 4096  # ZB200_ATTN_NOALLOC_MAX_UNITS (include/zonos_b200.h)


def attn_decode_noalloc(q: torch.Tensor, kv: PagedKV, layer: int, n_keys: int, n_heads: int, nsplit: int, out: torch.Tensor):
    """zb200_attn_decode_noalloc: split-KV scratch in the library's built-in device array; this wrapper allocates nothing
    either (q and out are the caller's CUDA bf16 tensors).  Calls that overlap on the GPU race on the shared scratch."""
    assert q.is_cuda and out.is_cuda and q.dtype == out.dtype == torch.bfloat16 and q.is_contiguous() and out.is_contiguous()
    check(lib().zb200_attn_decode_noalloc(ptr(q), ptr(kv.pool), ptr(kv.block_table), kv.pages_per_seq, kv.num_pages, n_heads,
                                          kv.n_kv_heads, layer, q.shape[0], n_keys, nsplit, ptr(out), stream_ptr()),
          "attn_decode_noalloc")
    return out


def attn_prefill(q: torch.Tensor, kv: PagedKV, layer: int, pos0: int, n_heads: int) -> torch.Tensor:
    q = _cuda(q, torch.bfloat16)
    n_seq, S = q.shape[:2]
    out = torch.empty_like(q)
    check(lib().zb200_attn_prefill(ptr(q), ptr(kv.pool), ptr(kv.block_table), kv.pages_per_seq, kv.num_pages, n_heads,
                                   kv.n_kv_heads, layer, n_seq, S, pos0, ptr(out), stream_ptr()), "attn_prefill")
    return out


# ---- Mamba2 mixer decode step (hybrid backbone) -------------------------------------------------------
def mamba2_conv_update(conv_state: torch.Tensor, zxbcdt: torch.Tensor, col0: int, w: torch.Tensor, bias) -> torch.Tensor:
    """causal_conv1d_update + SiLU.  conv_state bf16 [rows, C, 4] (updated in place), zxbcdt bf16 [rows, ld]."""
    rows, C, K = conv_state.shape
    assert conv_state.is_cuda and conv_state.dtype == torch.bfloat16 and conv_state.is_contiguous()
    assert zxbcdt.dtype == torch.bfloat16 and zxbcdt.is_contiguous() and zxbcdt.shape[0] == rows
    out = torch.empty(rows, C, dtype=torch.bfloat16, device=conv_state.device)
    check(lib().zb200_mamba2_conv_update(ptr(conv_state), ptr(zxbcdt), zxbcdt.shape[1], col0, ptr(_cuda(w, torch.bfloat16)),
                                         ptr(None if bias is None else _cuda(bias, torch.bfloat16)), ptr(out), rows, C, K,
                                         stream_ptr()), "mamba2_conv_update")
    return out


def mamba2_ssd_step(ssm_state: torch.Tensor, xbc: torch.Tensor, zxbcdt: torch.Tensor, dt_col0: int, A_log, D, dt_bias,
                    n_groups: int = 1) -> torch.Tensor:
    """selective_state_update for one token.  ssm_state bf16 [rows, H, P, N] (updated in place) -> y bf16 [rows, H*P]."""
    rows, H, P, N = ssm_state.shape
    assert ssm_state.is_cuda and ssm_state.dtype == torch.bfloat16 and ssm_state.is_contiguous()
    assert xbc.dtype == torch.bfloat16 and xbc.is_contiguous() and zxbcdt.is_contiguous()
    y = torch.empty(rows, H * P, dtype=torch.bfloat16, device=ssm_state.device)
    f32 = lambda t: _cuda(t, torch.float32)
    a, d, b = f32(A_log), f32(D), f32(dt_bias)
    check(lib().zb200_mamba2_ssd_step(ptr(ssm_state), ptr(xbc), xbc.shape[1], ptr(zxbcdt), zxbcdt.shape[1], dt_col0, ptr(a),
                                      ptr(d), ptr(b), ptr(y), rows, H, P, N, n_groups, stream_ptr()), "mamba2_ssd_step")
    return y


def mamba2_gated_rmsnorm(y: torch.Tensor, zxbcdt: torch.Tensor, w: torch.Tensor, eps: float, n_groups: int = 1) -> torch.Tensor:
    """RMSNormGated(norm_before_gate=False) with z = zxbcdt[:, :d_inner]."""
    rows, Di = y.shape
    assert y.is_cuda and y.dtype == torch.bfloat16 and y.is_contiguous() and zxbcdt.is_contiguous()
    out = torch.empty_like(y)
    check(lib().zb200_mamba2_gated_rmsnorm(ptr(y), ptr(zxbcdt), zxbcdt.shape[1], ptr(_cuda(w, torch.bfloat16)), ptr(out), rows,
                                           Di, n_groups, ctypes.c_float(eps), stream_ptr()), "mamba2_gated_rmsnorm")
    return out


def mamba2_conv_seq(conv_state: torch.Tensor, zxbcdt: torch.Tensor, seq_len: int, col0: int, w, bias, use_initial_state=False):
    """Prefill form of mamba2_conv_update: zxbcdt bf16 [n_seq*seq_len, ld] (sequence-major) -> bf16 [n_seq*seq_len, C];
    conv_state [n_seq, C, 4] receives the final window."""
    n_seq, C, K = conv_state.shape
    assert conv_state.dtype == torch.bfloat16 and conv_state.is_contiguous() and zxbcdt.is_contiguous()
    assert zxbcdt.shape[0] == n_seq * seq_len
    out = torch.empty(n_seq * seq_len, C, dtype=torch.bfloat16, device=conv_state.device)
    check(lib().zb200_mamba2_conv_seq(ptr(conv_state), ptr(zxbcdt), zxbcdt.shape[1], col0, ptr(_cuda(w, torch.bfloat16)),
                                      ptr(None if bias is None else _cuda(bias, torch.bfloat16)), ptr(out), n_seq, seq_len, C, K,
                                      int(use_initial_state), stream_ptr()), "mamba2_conv_seq")
    return out


def mamba2_ssd_scan(ssm_state: torch.Tensor, xbc: torch.Tensor, zxbcdt: torch.Tensor, seq_len: int, dt_col0: int, A_log, D,
                    dt_bias, n_groups: int = 1, use_initial_state=False) -> torch.Tensor:
    """Prefill form of mamba2_ssd_step (sequential scan, state in registers): -> y bf16 [n_seq*seq_len, H*P]."""
    n_seq, H, P, N = ssm_state.shape
    assert ssm_state.dtype == torch.bfloat16 and ssm_state.is_contiguous() and xbc.is_contiguous() and zxbcdt.is_contiguous()
    y = torch.empty(n_seq * seq_len, H * P, dtype=torch.bfloat16, device=ssm_state.device)
    f32 = lambda t: _cuda(t, torch.float32)
    a, d, b = f32(A_log), f32(D), f32(dt_bias)
    check(lib().zb200_mamba2_ssd_scan(ptr(ssm_state), ptr(xbc), xbc.shape[1], ptr(zxbcdt), zxbcdt.shape[1], dt_col0, ptr(a), ptr(d),
                                      ptr(b), ptr(y), n_seq, seq_len, H, P, N, n_groups, int(use_initial_state), stream_ptr()),
          "mamba2_ssd_scan")
    return y


def add_norm(x: torch.Tensor, residual, w: torch.Tensor, b, eps: float, rms: bool):
    """Fused residual add + (RMS|Layer)Norm of the hybrid blocks (fp32 residual stream).  x bf16 [rows, d]; residual fp32
    [rows, d] updated in place, or None for the first layer.  Returns (normed bf16, residual fp32)."""
    x = _cuda(x, torch.bfloat16)
    rows, d = x.shape
    has = residual is not None
    if not has:
        residual = torch.empty(rows, d, dtype=torch.float32, device=x.device)
    assert residual.dtype == torch.float32 and residual.is_contiguous()
    out = torch.empty_like(x)
    check(lib().zb200_add_norm(ptr(x), ptr(residual), int(has), ptr(_cuda(w, torch.bfloat16)),
                               ptr(None if b is None else _cuda(b, torch.bfloat16)), ptr(out), rows, d, ctypes.c_float(eps),
                               int(rms), stream_ptr()), "add_norm")
    return out, residual


def gemm_add_norm(a: torch.Tensor, w_packed: torch.Tensor, n: int, residual: torch.Tensor, w: torch.Tensor, b, eps: float,
                  rms: bool, ws: torch.Tensor):
    """Decode rows of the hybrid blocks: (a @ W^T in bf16) -> residual += . (fp32, in place) -> norm of the next block, as a
    deferred split-K weight-streaming GEMM + the add-norm kernel reducing its partial sums.  Returns the normed bf16 rows."""
    a = _cuda(a, torch.bfloat16)
    m, k = a.shape
    assert residual.dtype == torch.float32 and residual.is_contiguous() and residual.shape == (m, n)
    assert ws.dtype == torch.float32 and ws.numel() >= m * n
    out = torch.empty(m, n, dtype=torch.bfloat16, device=a.device)
    check(lib().zb200_gemm_add_norm(ptr(a), m, k, ptr(w_packed), n, k, ptr(ws), ctypes.c_int64(ws.numel()), ptr(residual), 1,
                                    ptr(_cuda(w, torch.bfloat16)), ptr(None if b is None else _cuda(b, torch.bfloat16)), ptr(out),
                                    ctypes.c_float(eps), int(rms), stream_ptr()), "gemm_add_norm")
    return out


def rope_split(qkv: torch.Tensor, inv_freq: torch.Tensor, seq_len: int, pos0: int, n_heads: int, n_kv_heads: int, pos_ptr=None):
    """Hybrid attention layers: Wqkv output bf16 [rows, (H+2Hkv)*128] -> rotated q [rows, H*128], rotated k and v
    [rows, Hkv*128] (non-interleaved rotary at position pos0 + row % seq_len)."""
    qkv = _cuda(qkv, torch.bfloat16)
    rows = qkv.shape[0]
    assert qkv.shape[1] == (n_heads + 2 * n_kv_heads) * 128
    q = torch.empty(rows, n_heads * 128, dtype=torch.bfloat16, device=qkv.device)
    k = torch.empty(rows, n_kv_heads * 128, dtype=torch.bfloat16, device=qkv.device)
    v = torch.empty_like(k)
    check(lib().zb200_rope_split(ptr(qkv), ptr(_cuda(inv_freq, torch.float32)), ptr(q), ptr(k), ptr(v), rows, seq_len, pos0,
                                 ctypes.c_void_p(0 if pos_ptr is None else pos_ptr), n_heads, n_kv_heads, stream_ptr()), "rope_split")
    return q, k, v


def gated_mlp_act(fc1_out: torch.Tensor) -> torch.Tensor:
    """mamba_ssm GatedMLP: y, gate = fc1_out.chunk(2, -1) -> y * silu(gate)."""
    fc1_out = _cuda(fc1_out, torch.bfloat16)
    rows, two_f = fc1_out.shape
    out = torch.empty(rows, two_f // 2, dtype=torch.bfloat16, device=fc1_out.device)
    check(lib().zb200_gated_mlp_act(ptr(fc1_out), ptr(out), rows, two_f // 2, stream_ptr()), "gated_mlp_act")
    return out


def heads_cfg(xn: torch.Tensor, w_heads_packed: torch.Tensor, batch: int, n_q: int, vocab: int, cfg_scale: float, ws=None,
              tickets=None) -> torch.Tensor:
    """apply_heads + CFG mix (zonos/model.py:100-116): xn bf16 [2B, d] -> fp32 logits [B, n_q, vocab]."""
    xn = _cuda(xn, torch.bfloat16)
    d = xn.shape[1]
    n_tiles = n_q * (HEAD_ROWS_PADDED // 64)
    logits = torch.empty(batch, n_q, vocab, dtype=torch.float32, device=xn.device)
    if ws is None:
        ws = torch.empty(148 * 32768, dtype=torch.float32, device=xn.device)
    if tickets is None:
        tickets = torch.zeros(n_tiles, dtype=torch.int32, device=xn.device)
    check(lib().zb200_heads_cfg(ptr(xn), batch, d, ptr(w_heads_packed), n_q, vocab, HEAD_ROWS_PADDED, ctypes.c_float(cfg_scale),
                                ptr(logits), 148, ptr(ws), ctypes.c_int64(ws.numel()), ptr(tickets), tickets.numel(), stream_ptr()),
          "heads_cfg")
    return logits


# ---- decode-step forms with device-side position / column (host-sequenced steps replayed from a CUDA graph) -----------------
def kv_write_at(kv: PagedKV, layer: int, k: torch.Tensor, v: torch.Tensor, pos_ptr: int):
    """k, v bf16 [n_seq, n_kv*128] -> the pool at position *pos_ptr (device int32 address)."""
    check(lib().zb200_kv_write_at(ptr(k), ptr(v), ptr(kv.pool), ptr(kv.block_table), kv.pages_per_seq, kv.num_pages,
                                  kv.n_kv_heads, layer, k.shape[0], ctypes.c_void_p(pos_ptr), stream_ptr()), "kv_write_at")


def attn_decode_at(q: torch.Tensor, kv: PagedKV, layer: int, pos_ptr: int, n_heads: int) -> torch.Tensor:
    out = torch.empty_like(q)
    check(lib().zb200_attn_decode_at(ptr(q), ptr(kv.pool), ptr(kv.block_table), kv.pages_per_seq, kv.num_pages, n_heads,
                                     kv.n_kv_heads, layer, q.shape[0], ctypes.c_void_p(pos_ptr), 1, ptr(None), ptr(None), ptr(out),
                                     stream_ptr()), "attn_decode_at")
    return out


def embed_codes_at(delayed_i32: torch.Tensor, col_ptr: int, col_add: int, emb: torch.Tensor, dup: bool = True) -> torch.Tensor:
    """Embedding sum of column *col_ptr + col_add of delayed int32 [B, Q, T] -> bf16 [2B or B, d] (rows duplicated for CFG)."""
    B, Q, T = delayed_i32.shape
    _, V, d = emb.shape
    out = torch.empty((2 * B if dup else B), d, dtype=torch.bfloat16, device=delayed_i32.device)
    check(lib().zb200_embed_sum_at(ptr(delayed_i32), Q * T, T, ctypes.c_void_p(col_ptr), col_add, ptr(emb), ptr(out), B, Q, V, d,
                                   B if dup else 0, stream_ptr()), "embed_sum_at")
    return out
