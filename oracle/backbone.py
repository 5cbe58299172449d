"""Transformer backbone, embeddings, heads and CFG logits on CPU (test infrastructure; see oracle/__init__.py).

Functional restatement (weights are a plain dict keyed like the reference state-dict) of
  zonos/backbone/_torch.py:9-15   precompute_freqs_cis      -> rope_table
  zonos/backbone/_torch.py:18-30  apply_rotary_emb          -> rope_interleaved
  zonos/backbone/_torch.py:33-49  _update_kv_cache          -> DenseKV.append
  zonos/backbone/_torch.py:117-141 Attention.forward        -> _attention
  zonos/backbone/_torch.py:150-152 FeedForward.forward      -> _mlp
  zonos/backbone/_torch.py:73-80,99-102 backbone/block fwd  -> backbone_forward
  zonos/model.py:97-98            embed_codes               -> embed_codes
  zonos/model.py:100-116          apply_heads/_compute_logits -> compute_logits
Every op rounds to the tensor dtype (bf16) exactly where the eager reference does.
"""
import torch
import torch.nn.functional as F

from .shapes import SHAPES, N_CODEBOOKS


def rope_table(n_pos: int, head_dim: int, base: float = 10000.0, device=None) -> torch.Tensor:
    """fp32 [n_pos, head_dim/2, 2] of (cos, sin);  theta_i = base^(-2i/head_dim)."""
    inv = 1.0 / (base ** (torch.arange(0, head_dim, 2, device=device)[: head_dim // 2].float() / head_dim))
    ang = torch.outer(torch.arange(n_pos, device=device), inv)
    unit = torch.polar(torch.ones_like(ang), ang)
    return torch.stack([unit.real, unit.imag], dim=-1)


def rope_interleaved(x: torch.Tensor, cs: torch.Tensor) -> torch.Tensor:
    """x [B,S,H,D] (any float dtype), cs [B or 1, S, D/2, 2] fp32. Pairs (x[2i], x[2i+1]); fp32 math;
    products rounded separately (no fused multiply-add), result cast back to x.dtype."""
    xf = x.float().reshape(*x.shape[:-1], -1, 2)
    c = cs[:, :, None, :, 0]
    s = cs[:, :, None, :, 1]
    a, b = xf[..., 0], xf[..., 1]
    out = torch.stack([a * c - b * s, b * c + a * s], dim=-1)
    return out.flatten(3).to(x.dtype)


class DenseKV:
    """Per-layer dense cache [B2, max_seq, 2, Hkv, D] (the reference layout, _torch.py:96-97)."""

    def __init__(self, n_layer, batch2, max_seq, n_kv, head_dim, dtype=torch.bfloat16, device=None):
        self.buf = [torch.zeros(batch2, max_seq, 2, n_kv, head_dim, dtype=dtype, device=device) for _ in range(n_layer)]
        self.len = 0  # uniform over the batch

    def append(self, layer, k, v):
        S = k.shape[1]
        assert self.len + S <= self.buf[layer].shape[1], "KV cache overflow"
        self.buf[layer][:, self.len:self.len + S, 0] = k
        self.buf[layer][:, self.len:self.len + S, 1] = v
        return self.buf[layer][:, : self.len + S]


def _attention(W, p, shape, x, kv: DenseKV, layer, cs):
    H, Hkv = shape["n_heads"], shape["n_kv_heads"]
    D = shape["d_model"] // H
    B2, S, _ = x.shape
    qkv = F.linear(x, W[p + "mixer.in_proj.weight"])
    q, k, v = qkv.split([H * D, Hkv * D, Hkv * D], dim=-1)
    q = rope_interleaved(q.view(B2, S, H, D), cs)
    k = rope_interleaved(k.view(B2, S, Hkv, D), cs)
    v = v.view(B2, S, Hkv, D)
    cache = kv.append(layer, k, v)
    kk, vv = cache.unbind(dim=-3)
    y = F.scaled_dot_product_attention(
        q.transpose(1, 2), kk.transpose(1, 2), vv.transpose(1, 2), is_causal=S > 1, enable_gqa=True)
    y = y.transpose(1, 2).contiguous().view(B2, S, H * D)
    return F.linear(y, W[p + "mixer.out_proj.weight"])


def _mlp(W, p, x):
    h = F.linear(x, W[p + "mlp.fc1.weight"])
    val, gate = h.chunk(2, dim=-1)  # gate is the SECOND half (Appendix B.9)
    return F.linear(val * F.silu(gate), W[p + "mlp.fc2.weight"])


def backbone_forward(W, shape_name, hidden, kv: DenseKV, table, eps=1e-5):
    """hidden [B2,S,d] -> [B2,S,d].  Does NOT advance kv.len (the caller does, like model.py:317-319,420-421)."""
    shape = SHAPES[shape_name] if isinstance(shape_name, str) else shape_name
    d = shape["d_model"]
    S = hidden.shape[1]
    cs = table[kv.len: kv.len + S][None]
    x = hidden
    for i in range(shape["n_layer"]):
        p = f"backbone.layers.{i}."
        x = x + _attention(W, p, shape, F.layer_norm(x, (d,), W[p + "norm.weight"], W[p + "norm.bias"], eps),
                           kv, i, cs)
        x = x + _mlp(W, p, F.layer_norm(x, (d,), W[p + "norm2.weight"], W[p + "norm2.bias"], eps))
    return F.layer_norm(x, (d,), W["backbone.norm_f.weight"], W["backbone.norm_f.bias"], eps)


def embed_codes(W, codes):
    """codes int64 [B,9,S] -> [B,S,d]; sequential adds in codebook order, each rounded to the weight dtype."""
    acc = None
    for k in range(N_CODEBOOKS):
        e = F.embedding(codes[:, k], W[f"embeddings.{k}.weight"])
        acc = e if acc is None else acc + e
    return acc


def compute_logits(W, shape_name, hidden, kv, table, cfg_scale: float):
    """hidden [2B,S,d] -> fp32 [B,9,V]; last position only; CFG mix (rows [0,B) cond, [B,2B) uncond);
    pad slots (index >= 1025) set to -inf."""
    last = backbone_forward(W, shape_name, hidden, kv, table)[:, -1, :]
    logits = torch.stack([F.linear(last, W[f"heads.{k}.weight"]) for k in range(N_CODEBOOKS)], dim=1).float()
    cond, uncond = logits.chunk(2)
    out = uncond + (cond - uncond) * cfg_scale
    out[..., 1025:] = -torch.inf
    return out
