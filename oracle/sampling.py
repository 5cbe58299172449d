"""Logit post-processing and sampling on CPU (test infrastructure; see oracle/__init__.py).

Restates zonos/sampling.py:
  :131-169 modify_logit_for_repetition_penalty -> repetition_penalty
  :54-75   apply_unified                       -> unified
  :96-111  apply_top_p                         -> top_p_filter
  :77-93   apply_top_k                         -> top_k_filter
  :114-128 apply_min_p                         -> min_p_filter
  :11-33   multinomial (num_samples == 1)      -> exp_race
  :232-328 sample_from_logits                  -> sample
The debug printers (:35-52,171-226) are out of scope.  `noise` lets a test inject the Exp(1) draw.
"""
import torch


def repetition_penalty(logits, history, rp, window: int):
    """logits fp32 [B,Q,V]; history int64 [B,Q,T] (delayed codes so far); rp fp32 [B] or scalar.
    factor[v] = rp ** (#occurrences of v in the last `window` columns), ids clamped to V-1;
    logit <= 0 ? logit*factor : logit/factor."""
    V = logits.shape[-1]
    recent = history[..., -window:].clamp_max(V - 1).to(torch.int64)
    rp = torch.as_tensor(rp, dtype=logits.dtype)
    if rp.ndim == 1:
        rp = rp[:, None, None]
    factor = torch.ones_like(logits)
    per_tok = (torch.ones_like(logits) * rp)
    # product over duplicates, in column order (multiplication by the same value commutes exactly)
    for j in range(recent.shape[-1]):
        idx = recent[..., j:j + 1]
        factor.scatter_(2, idx, factor.gather(2, idx) * per_tok.gather(2, idx))
    return torch.where(logits <= 0, logits * factor, logits / factor)


def unified(probs, linear: float, conf: float, quad: float):
    logp = torch.log(probs.clamp_min(1e-20))
    ent = -(probs * logp).sum(-1, keepdim=True)
    return (logp * (linear + ent * conf) - logp ** 2 * quad).softmax(-1)


def top_p_filter(probs, p: float):
    srt, idx = torch.sort(probs, dim=-1, descending=True)
    before = torch.cumsum(srt, dim=-1) - srt  # mass strictly before each token
    srt = srt * (~(before > p)).float()
    out = probs.scatter(-1, idx, srt)
    return out / out.sum(-1, keepdim=True)


def top_k_filter(probs, k: int):
    kth = torch.topk(probs, min(k, probs.shape[-1])).values[..., -1:]
    out = torch.where(probs < kth, 0.0, probs)  # ties at the k-th value are kept
    return out / out.sum(-1, keepdim=True)


def min_p_filter(probs, min_p: float):
    top = probs.max(-1, keepdim=True).values
    out = probs.masked_fill(probs < min_p * top, 0.0)
    return out / out.sum(-1, keepdim=True)


def exp_race(probs, noise=None):
    """argmax(p / q), q ~ Exp(1)."""
    q = torch.empty_like(probs).exponential_(1) if noise is None else noise
    return torch.argmax(probs / q, dim=-1, keepdim=True).to(torch.int64)


def final_probs(logits, temperature=1.0, top_p=0.0, top_k=0, min_p=0.0, linear=0.0, conf=0.0, quad=0.0):
    """The distribution the draw is taken from (temperature > 0 branch)."""
    probs = torch.softmax(logits / temperature, dim=-1)
    if linear > 0:
        probs = unified(probs, linear, conf, quad)
    if top_p > 0:
        probs = top_p_filter(probs, top_p)
    if top_k > 0:
        probs = top_k_filter(probs, top_k)
    if min_p > 0:
        probs = min_p_filter(probs, min_p)
    return probs


def sample(logits, temperature=1.0, top_p=0.0, top_k=0, min_p=0.0, linear=0.0, conf=0.0, quad=0.0,
           history=None, rp=3.0, rp_window=2, noise=None):
    """-> int64 [B,Q,1].  Order: repetition penalty -> (T>0: softmax(l/T) -> unified -> top-p -> top-k ->
    min-p -> exponential race) | (T<=0: argmax of the penalised logits)."""
    rp_t = torch.as_tensor(rp, dtype=logits.dtype)
    if history is not None and bool((rp_t != 1.0).any()):
        logits = repetition_penalty(logits, history, rp_t, rp_window)
    if temperature > 0:
        return exp_race(final_probs(logits, temperature, top_p, top_k, min_p, linear, conf, quad), noise)
    return torch.argmax(logits, dim=-1, keepdim=True)
