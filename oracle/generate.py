"""The autoregressive generate loop with the fork's EOS policy, on CPU (test infrastructure; see oracle/__init__.py).

Restates zonos/model.py:224-457 (Zonos.generate), :181-202 (_prefill), :118-142 (_decode_one_token eager path),
:204-208 (setup_cache) and zonos/utils.py:7-10 (find_multiple).  Row-wise tensor code instead of the reference's
Python `for i in range(B)` (model.py:410-414); the whole-batch re-draw on a new EOS (model.py:380-393) is kept.
"""
import math

import torch

from . import backbone as bb
from . import sampling as smp
from .delay import apply_delay_pattern, revert_delay_pattern
from .shapes import SHAPES, EOS_ID, MASK_ID, N_CODEBOOKS

UNKNOWN = -1
GRACE_STEPS = 6  # model.py:336
ROPE_ROWS = 16384  # _torch.py:67


def ceil_to(n: int, k: int) -> int:
    return n if k == 0 or n % k == 0 else n + k - n % k


class ModelLogits:
    """Produces the per-step CFG logits from the transformer (prefill, then one frame at a time)."""

    def __init__(self, W, shape_name, prefix_cond, max_positions, cfg_scale):
        self.W, self.shape_name, self.cfg = W, shape_name, cfg_scale
        s = SHAPES[shape_name]
        self.device = prefix_cond.device  # everything follows the conditioning's device (CPU in the tests; bench.py's
        # eager-GPU baseline leg passes CUDA tensors)
        self.kv = bb.DenseKV(s["n_layer"], prefix_cond.shape[0], ceil_to(max_positions, 8), s["n_kv_heads"],
                             s["d_model"] // s["n_heads"], dtype=prefix_cond.dtype, device=self.device)
        self.table = bb.rope_table(ROPE_ROWS, s["d_model"] // s["n_heads"], device=self.device)
        self.prefix = prefix_cond

    def prefill(self, ids):  # ids [B,9,P+1]
        rep = self.prefix.shape[0] // ids.shape[0]
        hidden = torch.cat([self.prefix, bb.embed_codes(self.W, ids.repeat(rep, 1, 1))], dim=1)
        out = bb.compute_logits(self.W, self.shape_name, hidden, self.kv, self.table, self.cfg)
        self.kv.len += hidden.shape[1]
        return out

    def step(self, ids):  # ids [B,9,1]
        hidden = bb.embed_codes(self.W, ids).repeat(2, 1, 1)
        out = bb.compute_logits(self.W, self.shape_name, hidden, self.kv, self.table, self.cfg)
        self.kv.len += 1
        return out


class ScriptedLogits:
    """Feeds pre-baked logits (tests of the EOS state machine): script[0] = prefill, script[s] = step s."""

    def __init__(self, script):
        self.script, self.i = script, 0

    def prefill(self, ids):
        self.i = 1
        return self.script[0].clone()

    def step(self, ids):
        out = self.script[self.i].clone()
        self.i += 1
        return out


def generate(source, batch: int, l_cond: int, audio_prefix=None, max_new_tokens: int = 86 * 30,
             sampling_params=None, noise_fn=None, callback=None, return_state=False):
    """source: ModelLogits | ScriptedLogits.  noise_fn(shape) -> Exp(1) tensor (None: torch RNG).
    Returns list of B int64 [9, T_i] (and the final delayed-code matrix / offset when return_state)."""
    if sampling_params is None:  # generate()'s default dict (model.py:232-241)
        sampling_params = dict(top_p=0, top_k=0, min_p=0, linear=0.55, conf=0.4, quad=0.0,
                               repetition_penalty=3.0, repetition_penalty_window=2, temperature=1.0)
    # a user dict replaces the default wholesale; missing keys fall back to sample_from_logits' own
    # defaults (sampling.py:232-245), except repetition_penalty which generate() indexes (model.py:342).
    sp = dict(temperature=1.0, top_p=0.0, top_k=0, min_p=0.0, linear=0.0, conf=0.0, quad=0.0,
              repetition_penalty_window=2)
    sp.update(sampling_params)
    rp0 = float(sp.pop("repetition_penalty"))  # KeyError if absent, like the reference
    win = int(sp.pop("repetition_penalty_window"))

    def draw(logits, history, rp):
        noise = None
        if noise_fn is not None and sp["temperature"] > 0:
            noise = noise_fn(tuple(logits.shape))
        return smp.sample(logits, history=history, rp=rp, rp_window=win, noise=noise, **sp)

    B, Q = batch, N_CODEBOOKS
    dev = getattr(source, "device", None)
    P = 0 if audio_prefix is None else audio_prefix.shape[2]
    T = P + max_new_tokens
    codes = torch.full((B, Q, T), UNKNOWN, dtype=torch.int64, device=dev)
    if audio_prefix is not None:
        codes[..., :P] = audio_prefix
    delayed = apply_delay_pattern(codes, MASK_ID)  # [B,Q,T+Q]

    logits = source.prefill(delayed[..., : P + 1])
    tok = draw(logits, None, rp0)  # no repetition penalty on the first frame (model.py:304)
    offset = P + 1
    frame = delayed[..., offset:offset + 1]
    delayed[..., offset:offset + 1] = torch.where(frame == UNKNOWN, tok, frame)

    V = logits.shape[-1]
    bias = torch.zeros(B, Q, V, device=dev)
    bias[:, 1:, EOS_ID] = -torch.inf
    bias[:, 0, EOS_ID] -= torch.log(torch.tensor(1024.0))

    max_steps = delayed.shape[2] - offset
    remaining = torch.full((B,), max_steps, dtype=torch.int64, device=dev)
    stopping = torch.zeros(B, dtype=torch.bool, device=dev)
    eos_mode = torch.zeros(B, dtype=torch.bool, device=dev)
    grace = torch.full((B,), GRACE_STEPS, dtype=torch.int64, device=dev)
    rp = torch.full((B,), rp0, device=dev)
    kq = torch.arange(Q, device=dev)[None, :, None]
    step = 0

    while int(remaining.max()) > 0:
        offset += 1
        logits = source.step(delayed[..., offset - 1:offset]) + bias
        rp[eos_mode] = 1.0
        active = eos_mode & (grace > 0)
        logits[active, 0, EOS_ID] = -torch.inf
        grace[active] -= 1

        history = delayed[..., :offset]
        tok = draw(logits, history, rp)
        hit = tok[:, 0, 0] == EOS_ID
        fresh = hit & ~eos_mode
        if bool(fresh.any()):
            eos_mode |= fresh
            grace[fresh] = GRACE_STEPS
            logits[fresh, 0, EOS_ID] = -torch.inf
            tok = draw(logits, history, rp)  # every row is re-drawn (model.py:388-393)
            hit = tok[:, 0, 0] == EOS_ID

        remaining = torch.where(hit, remaining.clamp_max(Q), remaining)
        stopping |= hit
        diag = (Q - remaining).clamp_max(Q - 1)[:, None, None]
        st = stopping[:, None, None]
        tok = torch.where(st & (kq < diag), MASK_ID, torch.where(st & (kq == diag), EOS_ID, tok))

        frame = delayed[..., offset:offset + 1]
        delayed[..., offset:offset + 1] = torch.where(frame == UNKNOWN, tok, frame)
        remaining = remaining - 1
        step += 1
        if callback is not None and not callback(delayed[..., offset:offset + 1], step, max_steps):
            break

    out = revert_delay_pattern(delayed)
    eos_pos = (out[:, 0, :] == EOS_ID).int().argmax(dim=-1)
    eos_pos[eos_pos == 0] = out.shape[2]
    out = out[..., : offset - Q].clone()
    out[out >= 1024] = 0
    result = [out[i, :, P:int(eos_pos[i])].clone() for i in range(B)]
    if return_state:
        return result, delayed, offset
    return result


def generate_with_model(W, shape_name, prefix_cond, audio_prefix=None, max_new_tokens=86 * 30, cfg_scale=2.0,
                        sampling_params=None, noise_fn=None, callback=None, return_state=False):
    assert cfg_scale != 1
    B = prefix_cond.shape[0] // 2
    P = 0 if audio_prefix is None else audio_prefix.shape[2]
    src = ModelLogits(W, shape_name, prefix_cond, prefix_cond.shape[1] + P + max_new_tokens + N_CODEBOOKS, cfg_scale)
    return generate(src, B, prefix_cond.shape[1], audio_prefix, max_new_tokens, sampling_params, noise_fn,
                    callback, return_state)
