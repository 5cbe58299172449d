"""Pins oracle/ (the CPU restatement) to golden vectors produced by the unmodified reference
(oracle/gen_golden.py).  CPU only."""
import pytest
import torch

from oracle import dac as odac
from oracle import delay as odelay
from oracle import generate as ogen
from oracle import sampling as osmp
from oracle import scenarios as sc
from oracle import shapes
from oracle.gen_golden import MODEL_CASES, weights_checksum


def test_delay_pattern_bit_exact(golden):
    for case in golden("delay.pt"):
        d = odelay.apply_delay_pattern(case["codes"], 1025)
        assert torch.equal(d, case["delayed"])
        assert torch.equal(odelay.revert_delay_pattern(d), case["reverted"])


def test_repetition_penalty_and_filters(golden):
    g = golden("sampler.pt")
    logits, hist, rp, noise = sc.sampler_inputs()
    for w, ref in g["rep"].items():
        assert torch.equal(osmp.repetition_penalty(logits.clone(), hist, rp, w), ref)
    probs = g["filters"]["softmax"]
    assert torch.equal(osmp.unified(probs.clone(), 0.55, 0.4, 0.1), g["filters"]["unified"])
    assert torch.equal(osmp.top_p_filter(probs.clone(), 0.9), g["filters"]["top_p"])
    assert torch.equal(osmp.top_k_filter(probs.clone(), 50), g["filters"]["top_k"])
    assert torch.equal(osmp.min_p_filter(probs.clone(), 0.1), g["filters"]["min_p"])


@pytest.mark.parametrize("name", list(sc.SAMPLER_CASES))
def test_sample_tokens(golden, name):
    g = golden("sampler.pt")
    logits, hist, rp, noise = sc.sampler_inputs()
    tok = osmp.sample(logits.clone(), history=hist, rp=rp, rp_window=2, noise=noise, **sc.SAMPLER_CASES[name])
    assert torch.equal(tok, g["tokens"][name])


@pytest.mark.parametrize("name", list(sc.EOS_SCENARIOS))
def test_eos_state_machine(golden, name):
    g = golden("eos_scenarios.pt")[name]
    s = sc.EOS_SCENARIOS[name]
    frames = []
    res = ogen.generate(ogen.ScriptedLogits(sc.scripted_logits(s)), s["B"], 4, sc.scripted_prefix(s), s["max_new"],
                        dict(s["sampling"]), noise_fn=sc.NoiseTape(s["seed"]),
                        callback=lambda f, st, m: frames.append(f.clone()) or True)
    assert len(frames) == g["steps_run"]
    assert torch.equal(torch.cat(frames, -1), g["frames"])
    assert len(res) == len(g["codes"])
    for a, b in zip(res, g["codes"]):
        assert torch.equal(a, b)


def test_eos_scenarios_cover_the_policy(golden):
    """The fixtures must actually exercise: early exit, ragged lengths, and full-length rows."""
    g = golden("eos_scenarios.pt")
    s = sc.EOS_SCENARIOS["all_stop_greedy_rp"]
    assert g["all_stop_greedy_rp"]["steps_run"] < s["max_new"] + 9 - 1
    lens = [c.shape[1] for c in g["mixed_greedy"]["codes"]]
    assert len(set(lens)) == 3 and max(lens) == 40


@pytest.mark.parametrize("name", list(MODEL_CASES))
def test_generate_tiny_model(golden, name):
    g = golden("model_tiny.pt")
    W = shapes.synth_transformer_weights("tiny", seed=0)
    assert weights_checksum(W) == pytest.approx(g["weights_checksum"], rel=1e-9)
    B, Lc, P, N, sp, cfg = MODEL_CASES[name]
    cond = shapes.synth_conditioning(B, Lc, shapes.SHAPES["tiny"]["d_model"], seed=1)
    pre = None if P == 0 else torch.randint(0, 1024, (B, 9, P), generator=torch.Generator().manual_seed(2))
    res = ogen.generate_with_model(W, "tiny", cond, pre, N, cfg, dict(sp), noise_fn=sc.NoiseTape(77))
    for a, b in zip(res, g["cases"][name]["codes"]):
        assert torch.equal(a, b)


def test_embed_and_logits_tiny(golden):
    from oracle import backbone as ob

    g = golden("model_tiny.pt")
    W = shapes.synth_transformer_weights("tiny", seed=0)
    assert torch.equal(ob.embed_codes(W, g["embed"]["ids"]), g["embed"]["out"])
    # teacher-forced logits: prefill + 5 decode steps of case b3_greedy_rp_prefix, fed with the golden frames
    B, Lc, P, N, sp, cfg = MODEL_CASES["b3_greedy_rp_prefix"]
    case = g["cases"]["b3_greedy_rp_prefix"]
    cond = shapes.synth_conditioning(B, Lc, 512, seed=1)
    pre = torch.randint(0, 1024, (B, 9, P), generator=torch.Generator().manual_seed(2))
    codes = torch.full((B, 9, P + N), -1, dtype=torch.int64)
    codes[..., :P] = pre
    delayed = odelay.apply_delay_pattern(codes, 1025)
    src = ogen.ModelLogits(W, "tiny", cond, Lc + P + N + 9, cfg)
    assert torch.equal(src.prefill(delayed[..., :P + 1]), case["logits"][0])
    # frames[.., s] is the frame written at loop step s+1 (offset P+2+s); step s+1 consumes the frame at P+1+s
    first = torch.where(delayed[..., P + 1:P + 2] == -1, case["logits"][0].argmax(-1, keepdim=True),
                        delayed[..., P + 1:P + 2])
    ids = first
    for s in range(1, 6):
        assert torch.equal(src.step(ids), case["logits"][s])
        ids = case["frames"][..., s - 1:s]


def test_logits_real_shape_vs_reference(golden):
    """The oracle port at the REAL shape (26 layers, d = 2048, 16/4 heads) against logits the unmodified reference
    produced on the same seeded weights (tests/golden/model_real.pt, `python -m oracle.gen_golden --real`): prefill +
    5 teacher-forced decode frames, bit-exact like the tiny cases."""
    from oracle.gen_golden import MODEL_CASES_REAL

    g = golden("model_real.pt")
    W = shapes.synth_transformer_weights("real", seed=0)
    assert weights_checksum(W) == pytest.approx(g["weights_checksum"], rel=1e-9)
    (name, (B, Lc, P, N, sp, cfg)), = MODEL_CASES_REAL.items()
    case = g["cases"][name]
    cond = shapes.synth_conditioning(B, Lc, 2048, seed=1)
    pre = torch.randint(0, 1024, (B, 9, P), generator=torch.Generator().manual_seed(2))
    codes = torch.full((B, 9, P + N), -1, dtype=torch.int64)
    codes[..., :P] = pre
    delayed = odelay.apply_delay_pattern(codes, 1025)
    src = ogen.ModelLogits(W, "real", cond, Lc + P + N + 9, cfg)
    assert torch.equal(src.prefill(delayed[..., :P + 1]), case["logits"][0])
    ids = torch.where(delayed[..., P + 1:P + 2] == -1, case["logits"][0].argmax(-1, keepdim=True), delayed[..., P + 1:P + 2])
    for s in range(1, 6):
        assert torch.equal(src.step(ids), case["logits"][s])
        ids = case["frames"][..., s - 1:s]


def test_dac_decode(golden):
    g = golden("dac.pt")
    for arch_name, arch in (("tiny", shapes.DAC_TINY), ("real", shapes.DAC)):
        W = shapes.synth_dac_weights(seed=7, arch=arch)
        assert weights_checksum(W) == pytest.approx(g[arch_name]["weights_checksum"], rel=1e-9)
        for case in g[arch_name]["cases"]:
            wav = odac.decode(W, case["codes"])
            assert wav.shape == case["wav"].shape == (case["codes"].shape[0], 1, 512 * case["codes"].shape[2])
            assert float((wav - case["wav"]).abs().max()) < 2e-5  # fp32 conv summation order varies with threading
            assert 0.05 < float(case["wav"].std()) < 0.9  # the fixture is not degenerate / saturated


def test_dac_encode(golden):
    """oracle.dac.encode (encoder + RVQ, fp32) against HF DacModel.encode's codes (tests/golden/dac_encode.pt): bit-exact."""
    g = golden("dac_encode.pt")
    for name, enc, dec in (("tiny", shapes.DAC_ENC_TINY, shapes.DAC_TINY), ("real", shapes.DAC_ENC, shapes.DAC)):
        W = dict(shapes.synth_dac_weights(seed=7, arch=dec))
        W.update(shapes.synth_dac_encoder_weights(seed=8, arch=enc))
        assert weights_checksum(W) == pytest.approx(g[name]["weights_checksum"], rel=1e-9)
        for case in g[name]["cases"][:2]:
            codes = odac.encode(W, case["wav"])
            assert torch.equal(codes, case["codes"]) and len(torch.unique(codes)) > codes.shape[2]  # not degenerate


def test_speaker_encoder(golden):
    """oracle.speaker (resample, log mel filterbank, ResNet293 of SimAM blocks, ASP, bottleneck, LDA; fp32) against the
    reference's own modules (tests/golden/speaker.pt): two narrow stacks -- one from 44.1 kHz stereo, through the restated
    torchaudio resampler -- and the shipped ResNet293 architecture."""
    from oracle import speaker as osp
    from oracle.gen_golden import SPEAKER_CASES, speaker_wav

    g = golden("speaker.pt")
    for name, (ip, blocks, seed, seconds, sr, ch) in SPEAKER_CASES.items():
        c = g[name]
        sd, lda = osp.synth_speaker_weights(ip, blocks, seed)
        assert float(sum(float(v.double().abs().sum()) for v in sd.values())) == pytest.approx(c["weights_checksum"], rel=1e-9)
        wav = speaker_wav(seed, seconds, sr, ch)
        x = osp.resample(wav.mean(0, keepdim=True), sr, 16000)
        assert x.shape == c["wav16k"].shape and float((x - c["wav16k"]).abs().max()) < 2e-6
        # the log of small mel energies amplifies the 1e-6 of the resampler's summation order: features are compared on the
        # golden file's own 16 kHz signal, the whole path (from the raw signal) with the tolerance that noise needs
        feat = osp.log_fbank(c["wav16k"])
        assert float((feat - c["feat"]).abs().max()) < 2e-5
        emb = osp.embedding_from_features(feat, sd, blocks)
        scale = float(c["emb"].abs().max())
        assert float((emb - c["emb"]).abs().max()) < 2e-4 * scale, name
        emb, out = osp.speaker_embedding(wav, sr, sd, lda, blocks)
        tol = 2e-4 if sr == 16000 else 5e-3
        assert float((emb - c["emb"]).abs().max()) < tol * scale, name
        assert float((out - c["lda"]).abs().max()) < tol * float(c["lda"].abs().max()), name
        assert float(c["emb"].std()) > 0.05  # the fixture is not degenerate


def test_mamba2_oracle_pinned_against_transformers_mixer():
    """oracle/mamba2.py (restating mamba_ssm's Mamba2.step, which is absent here) against the pure-torch statement of the
    same mixer that transformers ships (Mamba2Mixer.torch_forward): chunked prefill of 11 tokens, then 5 cached decode
    steps, fp32 on the CPU.  n_groups = 1 (the hybrid's setting; transformers' gated norm ignores groups)."""
    transformers = pytest.importorskip("transformers")
    from transformers.cache_utils import DynamicCache
    from transformers.models.mamba2.configuration_mamba2 import Mamba2Config
    from transformers.models.mamba2.modeling_mamba2 import Mamba2Mixer

    from oracle import mamba2 as om

    p = om.synth_mixer_params(d_model=64, expand=2, head_dim=16, d_state=32, n_groups=1, seed=3)
    s = p["shape"]
    cfg = Mamba2Config(num_heads=s["n_heads"], head_dim=s["head_dim"], hidden_size=64, state_size=32, expand=2, n_groups=1,
                       conv_kernel=4, num_hidden_layers=1, chunk_size=8, use_bias=False, use_conv_bias=True,
                       layer_norm_epsilon=1e-5, vocab_size=16)
    m = Mamba2Mixer(cfg, layer_idx=0).float().eval()
    with torch.no_grad():
        m.in_proj.weight.copy_(p["in_proj.weight"])
        m.conv1d.weight.copy_(p["conv1d.weight"].unsqueeze(1))
        m.conv1d.bias.copy_(p["conv1d.bias"])
        m.dt_bias.copy_(p["dt_bias"])
        m.A_log.copy_(p["A_log"])
        m.D.copy_(p["D"])
        m.norm.weight.copy_(p["norm.weight"])
        m.out_proj.weight.copy_(p["out_proj.weight"])
    B, S, steps = 3, 11, 5
    u = torch.randn(B, S + steps, 64, generator=torch.Generator().manual_seed(0))
    cache = DynamicCache(config=cfg)
    with torch.no_grad():
        pre = m.torch_forward(u[:, :S], cache_params=cache)
        dec = [m.torch_forward(u[:, S + i:S + i + 1], cache_params=cache)[:, 0] for i in range(steps)]
    st = om.new_state(p, B)
    mine = [om.mixer_step(p, st, u[:, t]) for t in range(S + steps)]
    assert float((torch.stack(mine[:S], 1) - pre).abs().max()) < 2e-5
    for i in range(steps):
        assert float((mine[S + i] - dec[i]).abs().max()) < 2e-5, i
