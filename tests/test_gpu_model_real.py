"""Parity of the CUDA engine at the REAL shape (Zonos-v0.1-transformer: 26 layers, d = 2048, 16 q / 4 kv heads, d_ff 8192)
-- the shape every bench number is quoted on -- against (1) logits the unmodified reference produced on the same seeded
weights (tests/golden/model_real.pt) and (2) the CPU oracle, teacher-forced on the GPU's own tokens, through every decode
path the engine picks by batch size (B = 1: <= 8 rows; B = 16: <= 32 rows; B = 64: 128 rows, deferred split-K).

Gates.  The tiny-shape tests use max|dlogit| <= (|s| + |s-1|) * ulp_bf16 (head outputs one bf16 ulp apart, mixed by the
fp32 CFG formula u + (c - u) * s, zonos/model.py:111-114) and rel-RMS <= 1e-2 (a guess made before anything
was measured).  Both assume that the two implementations feed (nearly) the same hidden state to the heads; over 26 layers
of bf16 rounding that no longer holds -- for the REFERENCE ITSELF: its bf16 logits deviate from an fp32 run of the same
weights by rel-RMS 2.5e-2 and max 0.27 (tests/golden/model_real.pt, `logits_fp32_prefill`, made by the unmodified
reference / its bit-exact port).  Two bf16 evaluation orders can therefore differ from each other by up to about twice that.
Gates at this shape, anchored on that measured noise floor:
    rel-RMS(engine - reference) <= 2.5e-2      (observed 1.7e-2)
    max|engine - reference|     <= 0.40        (1.5 x the reference's own max deviation from fp32; observed 0.16-0.25)
    rel-RMS(engine - fp32)      <= 1.25 x rel-RMS(reference - fp32)   (the engine is not further from the truth)
A wrong layer, rotation, mask or rounding point gives errors of order 1-10 / rel-RMS of order 1.  Observed values are
printed per case.
"""
import math

import pytest
import torch

from oracle import delay as odelay
from oracle import generate as ogen
from oracle import sampling as osmp
from oracle import scenarios as sc
from oracle import shapes

pytestmark = pytest.mark.gpu

LOGIT_REL_RMS = 2.5e-2
LOGIT_MAX_ABS = 0.40


def logit_max_abs(cfg_scale, ref_logits):
    return LOGIT_MAX_ABS


@pytest.fixture(scope="module")
def real_weights():
    return shapes.synth_transformer_weights("real", seed=0)


@pytest.fixture(scope="module")
def real_model(real_weights):
    from zonos_b200.config import ZonosConfig
    from zonos_b200.model import Zonos

    cfg = ZonosConfig.from_dict(shapes.reference_config_dict("real"))
    m = Zonos.from_state_dict(cfg, real_weights, "cuda")
    m.packed()
    return m


def _compare(tag, got, want, cfg):
    fin = torch.isfinite(want)
    assert torch.equal(fin, torch.isfinite(got)), f"{tag}: -inf masks differ"
    d = got[fin] - want[fin]
    worst, rel = float(d.abs().max()), float(d.pow(2).mean().sqrt() / want[fin].pow(2).mean().sqrt())
    tol = logit_max_abs(cfg, want)
    print(f"{tag}: max|dlogit| {worst:.4f} (gate {tol:.4f}), rel-RMS {rel:.2e} (gate {LOGIT_REL_RMS:.1e})")
    assert worst <= tol + 1e-6, f"{tag}: max|dlogit| {worst} > {tol}"
    assert rel <= LOGIT_REL_RMS, f"{tag}: rel-RMS {rel}"
    return worst, rel


def test_real_shape_logits_vs_reference_golden(real_model, golden):
    """Prefill + 5 decode frames against the REFERENCE's own logits (golden file), teacher-forced with the reference's
    frames, so the 26-layer engine is pinned to the reference and not only to the port."""
    from oracle.gen_golden import MODEL_CASES_REAL

    g = golden("model_real.pt")
    (name, (B, Lc, P, N, sp, cfg)), = MODEL_CASES_REAL.items()
    case = g["cases"][name]
    cond = shapes.synth_conditioning(B, Lc, 2048, seed=1)
    pre = torch.randint(0, 1024, (B, 9, P), generator=torch.Generator().manual_seed(2))
    got = real_model._debug_prefill_logits(cond.cuda(), pre.cuda(), cfg).cpu()
    _compare(f"{name} prefill vs reference", got, case["logits"][0], cfg)
    l32, ref = case["logits_fp32_prefill"], case["logits"][0]
    fin = torch.isfinite(ref)
    rr = lambda a: float((a[fin] - l32[fin]).pow(2).mean().sqrt() / l32[fin].pow(2).mean().sqrt())
    print(f"{name} prefill vs fp32: engine rel-RMS {rr(got):.2e}, reference (bf16) rel-RMS {rr(ref):.2e}")
    assert rr(got) <= 1.25 * rr(ref)
    # decode frames: force the reference's frames through the engine's step (the sampler is bypassed)
    ref_frames = case["frames"]
    got_steps = real_model._debug_forced_decode_logits(cond.cuda(), pre.cuda(), cfg, case["logits"][0].argmax(-1, keepdim=True),
                                                       ref_frames[..., :4], N)
    for s, lg in enumerate(got_steps):
        _compare(f"{name} decode frame {s + 1} vs reference", lg.cpu(), case["logits"][s + 1], cfg)


@pytest.mark.parametrize("B,Lc,N,chain", [(1, 23, 8, False), (16, 16, 6, False), (64, 16, 6, False), (64, 16, 6, True), (2, 16, 6, True)])
def test_real_shape_teacher_forced_vs_oracle(real_model, real_weights, B, Lc, N, chain, monkeypatch):
    """Free-running greedy generate on the GPU; every frame's CFG logits vs the CPU oracle on the same token history.
    chain: the opt-in fused chain kernel (csrc/chain.cuh) instead of the default per-GEMM launches."""
    if chain:
        monkeypatch.setenv("ZB200_CHAIN", "1")
    cond = shapes.synth_conditioning(B, Lc, 2048, seed=21 + B)
    taps = []
    out = real_model.generate(cond.cuda(), None, max_new_tokens=N, batch_size=B, progress_bar=False,
                              sampling_params=dict(temperature=0.0, repetition_penalty=1.0),
                              _tap=lambda s, l, f: taps.append((l.cpu(), f.cpu())))
    assert len(out) == B and all(o.shape == (9, N) for o in out)
    n_cmp = min(len(taps), 7)  # prefill + 6 decode frames (the CPU oracle at 128 rows costs ~1.5 s per frame)
    delayed = odelay.apply_delay_pattern(torch.full((B, 9, N), -1, dtype=torch.int64), 1025)
    src = ogen.ModelLogits(real_weights, "real", cond, Lc + N + 9, 2.0)
    ref = [src.prefill(delayed[..., :1])]
    ids = taps[0][1]
    for s in range(1, n_cmp):
        ref.append(src.step(ids))
        ids = taps[s][1]
    worst = 0.0
    for s in range(n_cmp):
        w, _ = _compare(f"B={B} frame {s}", taps[s][0], ref[s], 2.0)
        worst = max(worst, w)
    stats = real_model.last_generate_stats
    print(f"B={B}: {n_cmp} frames compared, worst max|dlogit| {worst:.4f}, launches/frame {stats['launches_per_step']}")


def test_real_shape_c3_sampling_tokens(real_model, real_weights):
    """BASELINE configs[2] sampling (min-p 0.1, repetition penalty 3.0 over a window of 2, temperature 1, cfg 2) with
    INJECTED Exp(1) noise: every token the GPU drew must be the oracle's draw on the oracle's own (teacher-forced) logits
    with the same noise, or lose the exponential race by less than the logit tolerance (margin rule)."""
    B, Lc, N = 8, 16, 6
    sp = dict(top_p=0, top_k=0, min_p=0.1, linear=0, conf=0, quad=0, repetition_penalty=3.0, repetition_penalty_window=2,
              temperature=1.0)
    cond = shapes.synth_conditioning(B, Lc, 2048, seed=5)
    tape = sc.NoiseTape(91)
    n_draws = 2 * (N + 9) + 2
    noise = torch.stack([tape((B, 9, 1026)) for _ in range(n_draws)])
    taps = []
    real_model.generate(cond.cuda(), None, max_new_tokens=N, batch_size=B, progress_bar=False, sampling_params=dict(sp),
                        _noise=noise.cuda(), _tap=lambda s, l, f: taps.append((l.cpu(), f.cpu())))
    delayed = odelay.apply_delay_pattern(torch.full((B, 9, N), -1, dtype=torch.int64), 1025)
    src = ogen.ModelLogits(real_weights, "real", cond, Lc + N + 9, 2.0)
    ref = [src.prefill(delayed[..., :1])]
    ids = taps[0][1]
    for s in range(1, len(taps)):
        ref.append(src.step(ids))
        ids = taps[s][1]
    tol = logit_max_abs(2.0, torch.stack(ref))
    # frame 0 (no repetition penalty, no EOS bias: zonos/model.py:304), then the loop frames.  With random-init weights and
    # the cb0 EOS row zeroed no row reaches EOS here, so draw s consumes noise[s] (one draw per frame, no re-draw).
    hist = delayed.clone()
    flips = total = 0
    for s in range(min(len(taps), delayed.shape[-1] - 1)):  # the loop's last iteration addresses an empty slice
        lg = ref[s].clone()
        col = 1 + s
        if s > 0:
            lg[:, 1:, 1024] = -torch.inf
            lg[:, 0, 1024] -= math.log(1024.0)
            lg = osmp.repetition_penalty(lg, hist[..., :col], torch.full((B,), 3.0), 2)
        logp = torch.log_softmax(lg, -1)
        keep = logp >= logp.max(-1, keepdim=True).values + math.log(0.1)
        score = torch.where(keep, logp, torch.full_like(logp, -torch.inf)) - noise[s].log()  # argmax(p / q)
        want = score.argmax(-1)
        cell = hist[..., col]
        free = cell == -1
        got = taps[s][1][..., 0]
        bad = (want != got) & free
        total += int(free.sum())
        if bad.any():
            # the GPU's pick must be within tolerance of the winner of the race, or sit within tolerance of the min-p edge
            gi = got.clamp(0, 1025)[..., None]
            raw = logp - noise[s].log()
            margin = (raw.gather(-1, want[..., None]) - raw.gather(-1, gi)).squeeze(-1)
            edge = (logp.gather(-1, gi).squeeze(-1) - (logp.max(-1).values + math.log(0.1))).abs()
            edge_w = (logp.gather(-1, want[..., None]).squeeze(-1) - (logp.max(-1).values + math.log(0.1))).abs()
            ok = (margin.abs() <= 2 * tol) | (edge <= 2 * tol) | (edge_w <= 2 * tol)
            assert bool(ok[bad].all()), f"frame {s}: token differs outside the tolerance (margin {float(margin[bad].abs().max())})"
            flips += int(bad.sum())
        hist[..., col] = torch.where(free, got, cell)
    print(f"C3 sampling, injected noise: {total - flips}/{total} tokens identical to the oracle's draw, {flips} near-tie flips "
          f"(each within the margin rule)")
    # Every flip above was inside the margin rule.  How many there may be: a flip needs the top two of an exponential race
    # among the 2-3 min-p survivors to be closer than the logit perturbation (|dlogit| rms ~0.07, max ~0.25 at this shape,
    # the reference's own bf16 noise), which happens for ~5 % of the draws (observed 22 / 432); gate at 10 %.
    assert flips <= total // 10

