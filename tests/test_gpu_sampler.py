"""Sampler + EOS state machine parity (GPU): fused CUDA kernel vs golden vectors from the reference / the oracle."""
import math

import pytest
import torch

from oracle import scenarios as sc

pytestmark = pytest.mark.gpu


def _run_pass1(logits, hist, rp_scalar, rp_one_rows, params, noise):
    """One pass-1 call in loop mode on arbitrary logits: returns tokens [B,9,1] (int64, cpu)."""
    from zonos_b200.sampler import LoopState, ST_OFFSET

    B, Q, V = logits.shape
    T = hist.shape[2]
    delayed = torch.full((B, Q, T + 4), -1, dtype=torch.int32, device="cuda")
    delayed[..., :T] = hist.to(torch.int32).cuda()
    sp = dict(params)
    sp.update(repetition_penalty=rp_scalar, repetition_penalty_window=2)
    st = LoopState(B, Q, V, delayed, 0, 0, sp, noise=noise.cuda()[None].contiguous())
    st.state[ST_OFFSET] = T
    for r in rp_one_rows:  # rows whose penalty is 1.0: emulate through eos_mode (grace 0 -> EOS not masked)
        st.eos_mode[r] = 1
        st.grace[r] = 0
    lg = logits.clone()
    lg[:, 0, 1024] += math.log(1024.0)  # the loop-mode kernel subtracts ln(1024) on codebook 0 (model.py:324)
    import ctypes
    from zonos_b200._lib import check, lib, stream_ptr

    check(lib().zb200_sample(ctypes.byref(st.args(lg.cuda().contiguous(), True)), 1, stream_ptr()))
    torch.cuda.synchronize()
    return st.tok1.cpu().to(torch.int64)[..., None]


@pytest.mark.parametrize("name", list(sc.SAMPLER_CASES))
def test_sampler_cases_vs_reference(golden, name):
    g = golden("sampler.pt")
    logits, hist, rp, noise = sc.sampler_inputs()
    ref = g["tokens"][name]
    ones = [i for i, r in enumerate(rp.tolist()) if r == 1.0]
    checked = 0
    for value in sorted(set(rp.tolist()) - {1.0}):
        tok = _run_pass1(logits, hist, value, ones, sc.SAMPLER_CASES[name], noise)
        rows = [i for i, r in enumerate(rp.tolist()) if r == value or r == 1.0]
        assert torch.equal(tok[rows], ref[rows]), (name, value)
        checked += len(rows)
    assert checked >= logits.shape[0]


def _noise_stack(seed, n, shape):
    tape = sc.NoiseTape(seed)
    return torch.stack([tape(shape) for _ in range(n)]).cuda()


@pytest.mark.parametrize("name", list(sc.EOS_SCENARIOS))
def test_eos_state_machine_vs_reference(golden, name):
    from zonos_b200.sampler import run_scripted

    g = golden("eos_scenarios.pt")[name]
    s = sc.EOS_SCENARIOS[name]
    script = sc.scripted_logits(s).cuda()
    noise = None
    if s["sampling"].get("temperature", 1.0) > 0:
        noise = _noise_stack(s["seed"], 2 * (s["max_new"] + 9) + 2, (s["B"], 9, script.shape[-1]))
    frames = []
    res, st, steps = run_scripted(script, s["B"], sc.scripted_prefix(s), s["max_new"], dict(s["sampling"]), noise=noise,
                                  callback=lambda f, i, m: frames.append(f.cpu()) or True)
    assert steps == g["steps_run"]
    assert torch.equal(torch.cat(frames, -1), g["frames"])
    for a, b in zip(res, g["codes"]):
        assert torch.equal(a.cpu(), b)


def test_philox_draws_follow_the_distribution():
    """Without injected noise the Philox exponential race must sample from softmax(logits): chi-square-ish check."""
    from zonos_b200.sampler import LoopState

    B, Q, V = 64, 9, 1026
    g = torch.Generator().manual_seed(0)
    base = torch.full((V,), -30.0)
    base[:8] = torch.tensor([2.0, 1.5, 1.0, 0.5, 0.0, -0.5, -1.0, -1.5])
    logits = base.expand(B, Q, V).contiguous().cuda()
    p = torch.softmax(base, -1)[:8]
    counts = torch.zeros(8)
    n = 0
    for seed in range(8):
        delayed = torch.full((B, Q, 12), -1, dtype=torch.int32, device="cuda")
        st = LoopState(B, Q, V, delayed, 0, 0, dict(repetition_penalty=1.0, temperature=1.0), seed=seed)
        st.sample_first(logits)
        torch.cuda.synchronize()
        tok = st.tok1.cpu().flatten()
        assert int(tok.max()) < 8
        counts += torch.bincount(tok, minlength=8)[:8].float()
        n += tok.numel()
        assert torch.equal(delayed[..., 1].cpu(), st.tok1.cpu())  # first frame written at column P+1
    freq = counts / n
    assert (freq - p).abs().max() < 0.02, (freq, p)
    # different utterances / codebooks use different streams
    assert len(set(st.tok1.cpu().flatten().tolist())) > 3


def test_zero_noise_on_a_zero_probability_token_never_wins():
    """p / q with p == 0 (masked EOS on codebooks 1-8, the pad slot 1025, tokens cut by min-p) and q == 0
    would be NaN and could win the race.  Injected noise is 0 exactly on those entries; the draw must still be a token
    with non-zero probability, for every row."""
    B, Q, V = 3, 9, 1026
    g = torch.Generator().manual_seed(5)
    logits = torch.randn(B, Q, V, generator=g) * 2.0
    logits[..., 1025] = -torch.inf
    hist = torch.randint(0, 1024, (B, Q, 6), generator=g)
    noise = torch.empty(B, Q, V).exponential_(generator=g)
    noise[:, 1:, 1024] = 0.0   # EOS is masked to -inf on codebooks 1..8 in loop mode (model.py:322-323)
    noise[..., 1025] = 0.0     # pad slot
    lowest = logits[..., :1024].argmin(-1, keepdim=True)  # far below min_p * max: removed by the filter
    noise.scatter_(-1, lowest, 0.0)
    tok = _run_pass1(logits, hist, 1.0, [], dict(min_p=0.1, temperature=1.0), noise)[..., 0]
    assert int(tok.max()) <= 1024 and bool((tok[:, 1:] != 1024).all()) and bool((tok != lowest[..., 0]).all())
    # and the Philox path: never the pad slot, never a masked EOS, finite state
    from zonos_b200.sampler import LoopState, ST_OFFSET
    import ctypes
    from zonos_b200._lib import check, lib, stream_ptr

    delayed = torch.full((B, Q, 12), -1, dtype=torch.int32, device="cuda")
    delayed[..., :6] = hist.to(torch.int32).cuda()
    st = LoopState(B, Q, V, delayed, 0, 0, dict(min_p=0.1, temperature=1.0, repetition_penalty=1.0), seed=3)
    st.state[ST_OFFSET] = 6
    for draw in range(50):
        st.state[4] = draw  # ST_DRAW
        check(lib().zb200_sample(ctypes.byref(st.args(logits.cuda().contiguous(), True)), 1, stream_ptr()))
        t = st.tok1.cpu()
        assert int(t.max()) <= 1024 and int(t.min()) >= 0 and bool((t[:, 1:] != 1024).all())
