"""Seeded test scenarios shared by oracle/gen_golden.py and tests/ (test infrastructure; see oracle/__init__.py).

Scripted-logit scenarios pin the EOS state machine of zonos/model.py:335-424 (first EOS -> grace period with
re-draw; second EOS -> 9-step diagonal wind-down; early loop exit; ragged trimming :437-451) independently of
the transformer: the per-step CFG logits are replaced by a seeded script with the codebook-0 EOS logit boosted
at chosen (row, step) pairs.
"""
import torch

from .shapes import EOS_ID, V_PAD, N_CODEBOOKS

# name -> dict(B, P, max_new, seed, boosts=[(row, first_step, last_step)], sampling)
GREEDY = dict(temperature=0.0, repetition_penalty=1.0)
GREEDY_RP = dict(temperature=0.0, repetition_penalty=3.0, repetition_penalty_window=2)
MINP = dict(top_p=0, top_k=0, min_p=0.1, linear=0, conf=0, quad=0, repetition_penalty=3.0,
            repetition_penalty_window=2, temperature=1.0)
UNIFIED = dict(top_p=0, top_k=0, min_p=0, linear=0.55, conf=0.4, quad=0.0, repetition_penalty=3.0,
               repetition_penalty_window=2, temperature=1.0)
KITCHEN = dict(top_p=0.9, top_k=40, min_p=0.02, linear=0.3, conf=0.2, quad=0.1, repetition_penalty=2.0,
               repetition_penalty_window=8, temperature=0.8)

EOS_SCENARIOS = {
    # row 0: EOS at step 10 (grace, re-draw), again from step 20 -> wind-down; row 1: EOS at 5, blocked 6..11,
    # allowed at 13; row 2 never stops -> loop runs to max_steps.
    "mixed_greedy": dict(B=3, P=0, max_new=40, seed=11, boosts=[(0, 10, 10), (0, 20, 48), (1, 5, 11), (1, 13, 48)],
                         sampling=GREEDY),
    # all rows stop early -> the while loop exits before max_steps
    "all_stop_greedy_rp": dict(B=2, P=3, max_new=60, seed=12, boosts=[(0, 4, 4), (0, 12, 70), (1, 7, 70)],
                               sampling=GREEDY_RP),
    # EOS immediately after prefill's first draw is impossible (bias only applies in the loop); EOS at step 1
    "eos_at_first_step": dict(B=2, P=0, max_new=30, seed=13, boosts=[(0, 1, 1), (0, 9, 40), (1, 2, 40)],
                              sampling=GREEDY),
    # stochastic draws with injected noise, incl. the whole-batch re-draw
    "mixed_minp": dict(B=4, P=2, max_new=36, seed=14, boosts=[(0, 6, 6), (0, 15, 50), (2, 8, 50), (3, 30, 50)],
                       sampling=MINP),
    "mixed_unified": dict(B=3, P=0, max_new=32, seed=15, boosts=[(1, 3, 3), (1, 11, 50), (2, 20, 20)],
                          sampling=UNIFIED),
    # second EOS so late that the wind-down is cut by max_steps
    "late_eos": dict(B=2, P=0, max_new=24, seed=16, boosts=[(0, 5, 5), (0, 28, 40), (1, 31, 40)], sampling=GREEDY),
}


def scripted_logits(sc: dict, scale: float = 3.0, boost: float = 40.0) -> torch.Tensor:
    """fp32 [n_steps+1, B, 9, V]; index 0 is the prefill output.  Pad slot (1025) = -inf as _compute_logits
    leaves it (model.py:115)."""
    n = sc["max_new"] + N_CODEBOOKS  # >= max_steps + 1
    g = torch.Generator().manual_seed(sc["seed"])
    x = torch.randn(n + 1, sc["B"], N_CODEBOOKS, V_PAD, generator=g) * scale
    x[..., EOS_ID] -= 5.0
    for row, a, b in sc["boosts"]:
        x[a:b + 1, row, 0, EOS_ID] += boost
    x[..., 1025:] = -torch.inf
    return x


def scripted_prefix(sc: dict):
    if sc["P"] == 0:
        return None
    g = torch.Generator().manual_seed(sc["seed"] + 1000)
    return torch.randint(0, 1024, (sc["B"], N_CODEBOOKS, sc["P"]), generator=g)


class NoiseTape:
    """Deterministic Exp(1) noise source: draw i uses seed (base, i).  Same tape feeds reference, oracle, CUDA."""

    def __init__(self, seed: int):
        self.seed, self.i = seed, 0

    def __call__(self, shape):
        g = torch.Generator().manual_seed(self.seed * 100003 + self.i)
        self.i += 1
        # -log(U) with U in (0,1]; not Tensor.exponential_, which gen_golden patches to read this tape
        return (-torch.log(1.0 - torch.rand(shape, generator=g))).clamp_min(1e-10)


# sampler unit cases: name -> kwargs for sample()
SAMPLER_CASES = {
    "greedy": dict(temperature=0.0),
    "plain_t1": dict(temperature=1.0),
    "temp07_topp": dict(temperature=0.7, top_p=0.9),
    "topk50": dict(top_k=50),
    "minp": dict(min_p=0.1),
    "unified": dict(linear=0.55, conf=0.4, quad=0.0),
    "kitchen": dict(temperature=0.8, linear=0.3, conf=0.2, quad=0.1, top_p=0.8, top_k=20, min_p=0.05),
}


def sampler_inputs(seed: int = 21, B: int = 4):
    g = torch.Generator().manual_seed(seed)
    logits = torch.randn(B, N_CODEBOOKS, V_PAD, generator=g) * 3
    logits[:, 1:, EOS_ID] = -torch.inf
    logits[..., 1025:] = -torch.inf
    hist = torch.randint(0, 1026, (B, N_CODEBOOKS, 20), generator=g)
    hist[:, :, -2] = hist[:, :, -1]  # duplicates inside the window -> factor rp^2
    hist[0, :, -1] = 1025  # masked token in the window
    rp = torch.tensor([3.0, 1.0, 2.0, 1.5][:B])
    noise = torch.empty(B, N_CODEBOOKS, V_PAD).exponential_(1, generator=g)
    return logits, hist, rp, noise


# codes_to_wavs post-processing cases: (frames of 512 samples, leading silent frames, trailing silent frames, amplitude)
POSTPROC_CASES = {
    "one_frame": (1, 0, 0, 0.3),
    "three_frames": (3, 0, 0, 0.3),
    "four_frames": (4, 0, 0, 0.2),
    "short_block": (9, 1, 0, 0.25),  # 4608 samples: loudness on 0.1 s blocks
    "lead_tail": (40, 3, 2, 0.1),
    "tail_only": (100, 0, 5, 0.5),
    "long_lead": (180, 20, 1, 0.05),  # > 2 s: 0.4 s blocks; the silence outlasts the 16 frames trim_silence looks at
    "quiet_part": (240, 2, 3, 0.4),  # a passage below the relative gate
    "silent": (64, 0, 0, 0.0),
}
POSTPROC_GOLDEN = ("one_frame", "three_frames", "four_frames", "short_block", "lead_tail", "tail_only")


def postproc_wav(name: str) -> torch.Tensor:
    """fp32 [1, 512 * frames]: a few partials under a slow envelope plus noise, near-silent head / tail."""
    frames, lead, tail, amp = POSTPROC_CASES[name]
    n = 512 * frames
    g = torch.Generator().manual_seed(1000 + frames)
    t = torch.arange(n, dtype=torch.float64) / 44100.0
    x = torch.zeros(n, dtype=torch.float64)
    for f0, a in ((140.0, 1.0), (420.0, 0.5), (1900.0, 0.3), (5200.0, 0.15)):
        x += a * torch.sin(2 * torch.pi * f0 * t + float(torch.rand((), generator=g)) * 6.28)
    x *= 0.6 + 0.4 * torch.sin(2 * torch.pi * 3.1 * t)
    x = amp * (x / 2.0 + 0.05 * torch.randn(n, generator=g, dtype=torch.float64))
    if name == "quiet_part":
        x[n // 3 : n // 2] *= 0.01
    x += 1e-5 * torch.randn(n, generator=g, dtype=torch.float64)
    if lead:
        x[: 512 * lead] *= 1e-4
    if tail:
        x[n - 512 * tail :] *= 1e-4
    return x.to(torch.float32).unsqueeze(0)
