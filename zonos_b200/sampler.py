"""Host side of the fused sampler / EOS state machine (csrc/sampler.cu): owns the device-resident loop state.

Mirrors the sampling-related part of Zonos.generate (zonos/model.py:288-342): the delayed-code matrix, the
per-utterance `remaining_steps / stopping / eos_mode / steps_after_eos` tensors, and the sampling parameters.
"""
import ctypes

import torch

from ._lib import check, lib, stream_ptr

EOS_ID, MASK_ID, UNKNOWN_ID, GRACE_STEPS = 1024, 1025, -1, 6
ST_OFFSET, ST_POS, ST_DONE, ST_FINAL_OFFSET, ST_DRAW, ST_ANY_FRESH, ST_STEP = range(7)


class SamplerArgs(ctypes.Structure):
    _fields_ = [
        ("logits", ctypes.c_void_p), ("noise", ctypes.c_void_p), ("delayed", ctypes.c_void_p),
        ("tok1", ctypes.c_void_p), ("tok2", ctypes.c_void_p), ("state", ctypes.c_void_p),
        ("remaining", ctypes.c_void_p), ("grace", ctypes.c_void_p), ("stopping", ctypes.c_void_p),
        ("eos_mode", ctypes.c_void_p), ("mask2", ctypes.c_void_p), ("eos_old", ctypes.c_void_p),
        ("host_flags", ctypes.c_void_p),
        ("batch", ctypes.c_int32), ("n_q", ctypes.c_int32), ("vocab", ctypes.c_int32), ("t_delayed", ctypes.c_int32),
        ("loop_mode", ctypes.c_int32), ("first_offset", ctypes.c_int32), ("row_base", ctypes.c_int32),
        ("seed", ctypes.c_uint64),
        ("temperature", ctypes.c_float), ("top_p", ctypes.c_float), ("min_p", ctypes.c_float),
        ("linear", ctypes.c_float), ("conf", ctypes.c_float), ("quad", ctypes.c_float), ("rep_penalty", ctypes.c_float),
        ("top_k", ctypes.c_int32), ("rep_window", ctypes.c_int32),
    ]


def resolve_sampling_params(sampling_params):
    """Same defaulting as the reference: generate()'s default dict when None (model.py:232-241); otherwise the
    user's dict as-is, missing keys falling back to sample_from_logits' defaults (sampling.py:232-245), except
    `repetition_penalty`, which generate() indexes unconditionally (model.py:342 -> KeyError)."""
    if sampling_params is None:
        sampling_params = dict(top_p=0, top_k=0, min_p=0, linear=0.55, conf=0.4, quad=0.0, repetition_penalty=3.0,
                               repetition_penalty_window=2, temperature=1.0)
    sp = dict(temperature=1.0, top_p=0.0, top_k=0, min_p=0.0, linear=0.0, conf=0.0, quad=0.0,
              repetition_penalty_window=2)
    sp.update(sampling_params)
    sp["repetition_penalty"] = float(sampling_params["repetition_penalty"])
    return sp


class LoopState:
    """Device-side loop state of one generate() call for `batch` utterances."""

    def __init__(self, batch, n_q, vocab, delayed_i32, prefix_len, l_cond, sampling_params, seed=0, row_base=0,
                 noise=None, device="cuda"):
        self.batch, self.n_q, self.vocab = batch, n_q, vocab
        self.delayed = delayed_i32  # int32 [B, Q, T+Q]
        self.t_delayed = delayed_i32.shape[2]
        self.prefix_len = prefix_len
        self.first_offset = prefix_len + 1
        self.max_steps = self.t_delayed - self.first_offset
        i32 = dict(dtype=torch.int32, device=device)
        u8 = dict(dtype=torch.uint8, device=device)
        self.state = torch.zeros(16, **i32)
        self.state[ST_OFFSET] = self.first_offset + 1
        self.state[ST_POS] = l_cond + prefix_len + 1
        self.remaining = torch.full((batch,), self.max_steps, **i32)
        self.grace = torch.full((batch,), GRACE_STEPS, **i32)
        self.stopping, self.eos_mode = torch.zeros(batch, **u8), torch.zeros(batch, **u8)
        self.mask2, self.eos_old = torch.zeros(batch, **u8), torch.zeros(batch, **u8)
        self.tok1, self.tok2 = torch.zeros(batch, n_q, **i32), torch.zeros(batch, n_q, **i32)
        self.host_flags = torch.zeros(2, dtype=torch.int32).pin_memory()
        self.noise = noise  # optional fp32 [n_draws, B, Q, V] (tests)
        self.sp = resolve_sampling_params(sampling_params)
        self.seed, self.row_base = seed, row_base

    def args(self, logits, loop_mode):
        sp = self.sp
        a = SamplerArgs()
        a.logits, a.noise = logits.data_ptr(), (0 if self.noise is None else self.noise.data_ptr())
        a.delayed, a.tok1, a.tok2, a.state = (t.data_ptr() for t in (self.delayed, self.tok1, self.tok2, self.state))
        a.remaining, a.grace = self.remaining.data_ptr(), self.grace.data_ptr()
        a.stopping, a.eos_mode = self.stopping.data_ptr(), self.eos_mode.data_ptr()
        a.mask2, a.eos_old = self.mask2.data_ptr(), self.eos_old.data_ptr()
        a.host_flags = self.host_flags.data_ptr()
        a.batch, a.n_q, a.vocab, a.t_delayed = self.batch, self.n_q, self.vocab, self.t_delayed
        a.loop_mode, a.first_offset, a.row_base, a.seed = int(loop_mode), self.first_offset, self.row_base, self.seed
        a.temperature, a.top_p, a.min_p = float(sp["temperature"]), float(sp["top_p"]), float(sp["min_p"])
        a.linear, a.conf, a.quad = float(sp["linear"]), float(sp["conf"]), float(sp["quad"])
        a.rep_penalty, a.top_k, a.rep_window = sp["repetition_penalty"], int(sp["top_k"]), int(sp["repetition_penalty_window"])
        return a

    def sample_first(self, logits):
        """First frame after the prefill: no bias, no repetition penalty (model.py:304-314)."""
        check(lib().zb200_sample(ctypes.byref(self.args(logits, False)), 1, stream_ptr()), "sample(first)")

    def sample_step(self, logits):
        """One loop iteration's sampling + bookkeeping (model.py:353-424): two launches, no host sync."""
        a = self.args(logits, True)
        check(lib().zb200_sample(ctypes.byref(a), 1, stream_ptr()), "sample(pass 1)")
        check(lib().zb200_sample(ctypes.byref(a), 2, stream_ptr()), "sample(pass 2)")

    def finish(self, offset=None):
        """Epilogue of generate (model.py:437-451): revert the delay pattern, trim at the first EOS, zero codes >= 1024.
        Returns the list of int64 [Q, T_i]."""
        from .ops import revert_delay_pattern

        st = self.state.cpu()
        if offset is None:
            offset = int(st[ST_FINAL_OFFSET]) if int(st[ST_DONE]) else int(st[ST_OFFSET]) - 1
        out = revert_delay_pattern(self.delayed.to(torch.int64))
        eos_pos = (out[:, 0, :] == EOS_ID).int().argmax(dim=-1)
        eos_pos[eos_pos == 0] = out.shape[2]
        out = out[..., : offset - self.n_q]
        out.masked_fill_(out >= 1024, 0)
        ends = eos_pos.tolist()
        return [out[i, :, self.prefix_len:ends[i]].clone() for i in range(self.batch)]


def run_scripted(script, batch, prefix_codes, max_new_tokens, sampling_params, noise=None, callback=None):
    """Drives the sampler + state machine with pre-baked per-step logits (script[0] = prefill output) instead of the
    transformer.  Used by the parity tests of the EOS policy; same control flow as generate()."""
    from .ops import apply_delay_pattern

    dev = script.device
    Q, V = script.shape[2], script.shape[3]
    P = 0 if prefix_codes is None else prefix_codes.shape[2]
    codes = torch.full((batch, Q, P + max_new_tokens), UNKNOWN_ID, dtype=torch.int64, device=dev)
    if prefix_codes is not None:
        codes[..., :P] = prefix_codes.to(dev)
    delayed = apply_delay_pattern(codes, MASK_ID).to(torch.int32)
    st = LoopState(batch, Q, V, delayed, P, 4, sampling_params, noise=noise, device=dev)
    st.sample_first(script[0].contiguous())
    steps = 0
    for i in range(1, st.max_steps + 1):
        st.sample_step(script[i].contiguous())
        steps += 1
        torch.cuda.synchronize()
        if callback is not None:
            col = int(st.state[ST_OFFSET]) - 1
            if not callback(st.delayed[..., col:col + 1].to(torch.int64), steps, st.max_steps):
                break
        if int(st.host_flags[0]):
            break
    return st.finish(), st, steps
