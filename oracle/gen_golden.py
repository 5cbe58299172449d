"""Generates tests/golden/*.pt by running the UNMODIFIED reference on CPU (test infrastructure).

Run in the build container only (needs /root/reference):   python -m oracle.gen_golden
The reference ships no golden vectors of its own, so these files are the parity pin for
oracle/ and for the CUDA path.  Inputs are regenerated from seeds (oracle/shapes.py, oracle/scenarios.py); only
outputs (and small inputs) are stored, plus a checksum of the synthetic weights to catch RNG drift.
"""
import os
import sys

import torch

from . import scenarios as sc
from . import shapes
from .ref_import import import_reference

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def weights_checksum(W) -> float:
    return float(sum(v.float().abs().sum().item() for v in W.values()))


def build_reference_model(shape):
    import zonos.config as zc
    import zonos.model as zm
    from zonos.backbone import BACKBONES

    cfg = zc.ZonosConfig.from_dict(shapes.reference_config_dict(shape))
    m = zm.Zonos(cfg, BACKBONES["torch"]).to("cpu", torch.bfloat16)
    sd = m.state_dict()
    sd.update(shapes.synth_transformer_weights(shape, seed=0, padded_heads=False))
    m.load_state_dict(sd)  # post-hook pads heads 1025 -> 1026 (zonos/model.py:46-51)
    return m.eval()


def gen_delay():
    import zonos.codebook_pattern as zcp

    g = torch.Generator().manual_seed(3)
    cases = []
    for B, T in [(1, 1), (2, 5), (3, 9), (2, 40), (1, 300)]:
        codes = torch.randint(-1, 1026, (B, 9, T), generator=g)
        d = zcp.apply_delay_pattern(codes, 1025)
        cases.append(dict(codes=codes, delayed=d, reverted=zcp.revert_delay_pattern(d)))
    torch.save(cases, os.path.join(OUT, "delay.pt"))


def gen_sampler():
    import zonos.sampling as zs

    logits, hist, rp, noise = sc.sampler_inputs()
    out = dict(rep={}, tokens={}, filters={})
    for w in (2, 8, 16):
        out["rep"][w] = zs.modify_logit_for_repetition_penalty(logits.clone(), hist, rp.clone(), w)
    probs = torch.softmax(out["rep"][2], dim=-1)
    out["filters"]["softmax"] = probs
    out["filters"]["unified"] = zs.apply_unified(probs.clone(), 0.55, 0.4, 0.1)
    out["filters"]["top_p"] = zs.apply_top_p(probs.clone(), 0.9)
    out["filters"]["top_k"] = zs.apply_top_k(probs.clone(), 50)
    out["filters"]["min_p"] = zs.apply_min_p(probs.clone(), 0.1)
    orig = torch.Tensor.exponential_

    def patched(self, *a, **k):
        self.copy_(noise)
        return self

    for name, kw in sc.SAMPLER_CASES.items():
        torch.Tensor.exponential_ = patched
        try:
            out["tokens"][name] = zs.sample_from_logits(
                logits.clone(), generated_tokens=hist, repetition_penalty=rp.clone(), repetition_penalty_window=2, **kw)
        finally:
            torch.Tensor.exponential_ = orig
    torch.save(out, os.path.join(OUT, "sampler.pt"))


def _with_noise_tape(tape, fn):
    orig = torch.Tensor.exponential_

    def patched(self, *a, **k):
        self.copy_(tape(tuple(self.shape)))
        return self

    torch.Tensor.exponential_ = patched
    try:
        return fn()
    finally:
        torch.Tensor.exponential_ = orig


def gen_eos(model):
    out = {}
    for name, s in sc.EOS_SCENARIOS.items():
        script = sc.scripted_logits(s)
        it = {"i": 0}

        def fake_prefill(*a, **k):
            it["i"] = 1
            return script[0].clone()

        def fake_step(*a, **k):
            r = script[it["i"]].clone()
            it["i"] += 1
            return r

        model._prefill, model._decode_one_token = fake_prefill, fake_step
        frames = []

        def cb(frame, step, max_steps):
            frames.append(frame.clone())
            return True

        d = model.config.backbone.d_model
        cond = torch.zeros(2 * s["B"], 4, d, dtype=torch.bfloat16)
        res = _with_noise_tape(sc.NoiseTape(s["seed"]), lambda: model.generate(
            cond, sc.scripted_prefix(s), max_new_tokens=s["max_new"], batch_size=s["B"],
            sampling_params=dict(s["sampling"]), progress_bar=False, disable_torch_compile=True, callback=cb))
        out[name] = dict(codes=res, frames=torch.cat(frames, dim=-1), steps_run=len(frames), script_used=it["i"])
        del model._prefill, model._decode_one_token
    torch.save(out, os.path.join(OUT, "eos_scenarios.pt"))


MODEL_CASES = {
    # name: (B, L_cond, P, max_new, sampling, cfg_scale)
    "b1_greedy": (1, 23, 0, 24, sc.GREEDY, 2.0),
    "b3_greedy_rp_prefix": (3, 17, 5, 24, sc.GREEDY_RP, 2.0),
    "b2_minp": (2, 9, 9, 20, sc.MINP, 2.0),
    "b2_unified_cfg3": (2, 12, 0, 20, sc.UNIFIED, 3.0),
}


# The real (Zonos-v0.1-transformer) shape: one short case, so that the 26-layer d=2048 engine is pinned to the reference
# itself and not only to the oracle port (the port is then checked against this file on CPU, the engine on the GPU).
MODEL_CASES_REAL = {
    "real_b2_greedy_prefix": (2, 9, 3, 6, sc.GREEDY, 2.0),
}


def gen_model(model, shape, cases=None):
    """G2-G4 + G7: embed, prefill logits, teacher-forced decode logits, free-running generate."""
    W = shapes.synth_transformer_weights(shape, seed=0)
    d = shapes.SHAPES[shape]["d_model"]
    out = dict(weights_checksum=weights_checksum(W), cases={})
    g = torch.Generator().manual_seed(5)
    ids = torch.randint(0, 1026, (2, 9, 6), generator=g)
    out["embed"] = dict(ids=ids, out=model.embed_codes(ids))

    for name, (B, Lc, P, N, sp, cfg) in (MODEL_CASES if cases is None else cases).items():
        cond = shapes.synth_conditioning(B, Lc, d, seed=1)
        pre = None if P == 0 else torch.randint(0, 1024, (B, 9, P), generator=torch.Generator().manual_seed(2))
        logit_log = []
        orig_cl = model._compute_logits

        def spy(hidden, params, scale):
            r = orig_cl(hidden, params, scale)
            logit_log.append(r.clone())
            return r

        model._compute_logits = spy
        frames = []

        def cb(frame, step, max_steps):
            frames.append(frame.clone())
            return True

        res = _with_noise_tape(sc.NoiseTape(77), lambda: model.generate(
            cond, pre, max_new_tokens=N, cfg_scale=cfg, batch_size=B, sampling_params=dict(sp),
            progress_bar=False, disable_torch_compile=True, callback=cb))
        del model._compute_logits
        out["cases"][name] = dict(codes=res, frames=torch.cat(frames, dim=-1), logits=torch.stack(logit_log[:6]))
    torch.save(out, os.path.join(OUT, f"model_{shape}.pt"))


def gen_dac():
    from transformers.models.dac import DacConfig, DacModel

    out = {}
    for arch_name, arch in (("tiny", shapes.DAC_TINY), ("real", shapes.DAC)):
        cfg = DacConfig(sampling_rate=44100, hidden_size=arch["latent"], decoder_hidden_size=arch["dec_ch"])
        m = DacModel(cfg).eval()
        W = shapes.synth_dac_weights(seed=7, arch=arch)
        sd = m.state_dict()
        for k, v in W.items():
            assert sd[k].shape == v.shape, (k, sd[k].shape, v.shape)
        sd.update(W)
        m.load_state_dict(sd)
        g = torch.Generator().manual_seed(9)
        cases = [(1, 1), (3, 7), (2, 86)] if arch_name == "tiny" else [(1, 1), (1, 40)]
        res = []
        for B, T in cases:
            codes = torch.randint(0, 1024, (B, 9, T), generator=g)
            with torch.no_grad():
                wav = m.decode(audio_codes=codes).audio_values.unsqueeze(1).float()
            res.append(dict(codes=codes, wav=wav))
        out[arch_name] = dict(cases=res, weights_checksum=weights_checksum(W))
    torch.save(out, os.path.join(OUT, "dac.pt"))


def gen_dac_encode():
    """DacModel.encode (HF, fp32 CPU: the reference's DACAutoencoder.encode, zonos/autoencoder.py:27-28) on seeded
    waveforms; the codes are the golden.  Tiny and real architecture."""
    from transformers.models.dac import DacConfig, DacModel

    out = {}
    for arch_name, enc, dec in (("tiny", shapes.DAC_ENC_TINY, shapes.DAC_TINY), ("real", shapes.DAC_ENC, shapes.DAC)):
        cfg = DacConfig(sampling_rate=44100, encoder_hidden_size=enc["enc_ch"], hidden_size=enc["latent"],
                        decoder_hidden_size=dec["dec_ch"])
        m = DacModel(cfg).eval()
        W = dict(shapes.synth_dac_weights(seed=7, arch=dec))
        W.update(shapes.synth_dac_encoder_weights(seed=8, arch=enc))
        sd = m.state_dict()
        for k, v in W.items():
            assert sd[k].shape == v.shape, (k, sd[k].shape, v.shape)
        sd.update(W)
        m.load_state_dict(sd)
        g = torch.Generator().manual_seed(12)
        cases = [(1, 1), (2, 9), (3, 40)] if arch_name == "tiny" else [(1, 9), (2, 20)]
        res = []
        for B, T in cases:
            wav = 0.3 * torch.randn(B, 1, 512 * T, generator=g)
            with torch.no_grad():
                codes = m.encode(wav).audio_codes
            res.append(dict(wav=wav, codes=codes))
        out[arch_name] = dict(cases=res, weights_checksum=weights_checksum(W))
    torch.save(out, os.path.join(OUT, "dac_encode.pt"))


def gen_postproc():
    """The reference's codes_to_wavs with decode patched to hand back the test waveform.  pyloudnorm is not installed
    here, so its normalize_loudness takes its own except-branch (audio unchanged): these fixtures pin the trim and the
    fades (see oracle/postproc.py's header)."""
    import logging

    import zonos.autoencoder as zae

    zae.logger.setLevel(logging.ERROR)
    ae = object.__new__(zae.DACAutoencoder)
    ae.sampling_rate, ae.num_codebooks = 44100, 9
    out = {}
    for name in sc.POSTPROC_GOLDEN:
        wav = sc.postproc_wav(name)
        ae.decode = lambda c, w=wav: w.clone().unsqueeze(0)
        (res,) = ae.codes_to_wavs(torch.zeros(1, 9, wav.shape[1] // 512, dtype=torch.long))
        trimmed = ae.trim_silence(wav)
        out[name] = dict(wav=res, trimmed_len=trimmed.shape[1], in_sum=float(wav.double().sum()))
    torch.save(out, os.path.join(OUT, "postproc.pt"))


# The conditioner list of the Zonos-v0.1 checkpoints (SURVEY Appendix A / CONDITIONING_README.md) at a small width.
COND_D = 512
COND_LIST = [
    dict(type="EspeakPhonemeConditioner", name="espeak"),
    dict(type="PassthroughConditioner", name="speaker", cond_dim=128, projection="linear", uncond_type="learned"),
    dict(type="FourierConditioner", name="emotion", input_dim=8, uncond_type="learned"),
    dict(type="FourierConditioner", name="fmax", min_val=0, max_val=24000, uncond_type="learned"),
    dict(type="FourierConditioner", name="pitch_std", min_val=0, max_val=400, uncond_type="learned"),
    dict(type="FourierConditioner", name="speaking_rate", min_val=0, max_val=40, uncond_type="learned"),
    dict(type="IntegerConditioner", name="language_id", min_val=-1, max_val=126, uncond_type="learned"),
]


def cond_inputs():
    """Seeded cond_dict in the reference's layout (make_cond_dict, zonos/conditioning.py:484-494): every tensor [1, 1, n],
    broadcast over the batch -- Conditioner.forward star-unpacks its input (:52) -- and phoneme ids [2, 11] in place of the
    text (eSpeak is not installed)."""
    g = torch.Generator().manual_seed(21)
    ids = torch.randint(0, 189, (2, 11), generator=g)
    ids[1, :3] = 0  # left padding like tokenize_phonemes (zonos/conditioning.py:190)
    emo = torch.rand(1, 1, 8, generator=g)
    return dict(espeak=ids, speaker=torch.randn(1, 1, 128, generator=g).to(torch.bfloat16), emotion=emo / emo.sum(-1, keepdim=True),
                fmax=torch.tensor(22050.0).view(1, 1, 1), pitch_std=torch.tensor(45.0).view(1, 1, 1),
                speaking_rate=torch.tensor(15.0).view(1, 1, 1), language_id=torch.tensor(24).view(1, 1, 1))


def gen_conditioning():
    """PrefixConditioner.forward of the unmodified reference (zonos/conditioning.py:373-389), bf16 like from_local makes
    it, for the full cond_dict and for the unconditional dict prepare_conditioning builds (zonos/model.py:216-217)."""
    import zonos.conditioning as zcond
    import zonos.config as zc

    assert len(zcond.SPECIAL_TOKEN_IDS) + len(zcond.symbols) == 189
    torch.manual_seed(11)
    pc = zcond.PrefixConditioner(zc.PrefixConditionerConfig(conditioners=[dict(c) for c in COND_LIST], projection="none"), COND_D)
    with torch.no_grad():
        for c in pc.conditioners:
            if c.uncond_vector is not None:
                c.uncond_vector.normal_()
        pc.norm.weight.normal_(1.0, 0.1)
        pc.norm.bias.normal_(0.0, 0.05)
    pc = pc.to(torch.bfloat16).eval()
    inp = cond_inputs()
    orig = (zcond.phonemize, zcond.tokenize_phonemes)
    zcond.phonemize = lambda texts, languages: texts
    zcond.tokenize_phonemes = lambda ph: (ph, [ph.shape[1]] * ph.shape[0])
    try:
        cd = {k: v for k, v in inp.items()}
        cd["espeak"] = (inp["espeak"], ["en-us", "en-us"])
        full = pc(cd)
        uncond = pc({k: cd[k] for k in pc.required_keys})
        partial = pc({k: cd[k] for k in ("espeak", "speaker", "language_id")})
    finally:
        zcond.phonemize, zcond.tokenize_phonemes = orig
    torch.save(dict(state_dict=pc.state_dict(), required_keys=sorted(pc.required_keys), full=full, uncond=uncond, partial=partial),
               os.path.join(OUT, "conditioning.pt"))


SPEAKER_CASES = {  # name: (in_planes, blocks per stage, seed, seconds, sample rate of the input, channels)
    "tiny_16k": (8, (1, 2, 2, 1), 11, 0.64, 16000, 1),
    "tiny_44k_stereo": (8, (1, 2, 2, 1), 12, 0.5, 44100, 2),
    "resnet293_16k": (64, (10, 20, 64, 3), 13, 1.0, 16000, 1),
}


def speaker_wav(seed, seconds, sr, channels):
    """Seeded test signal: a few decaying partials + noise, [channels, n] in [-1, 1]."""
    g = torch.Generator().manual_seed(seed)
    n = int(seconds * sr)
    t = torch.arange(n, dtype=torch.float64) / sr
    wav = torch.zeros(channels, n, dtype=torch.float64)
    for c in range(channels):
        for f0 in (110.0 * (c + 1), 440.0, 1730.0, 5200.0):
            wav[c] += float(torch.rand(1, generator=g)) * torch.sin(2 * torch.pi * f0 * t + float(torch.rand(1, generator=g)))
        wav[c] += 0.1 * torch.randn(n, generator=g, dtype=torch.float64)
    return (0.25 * wav).float()


def gen_speaker():
    """Speaker encoder (zonos/speaker_cloning.py): the reference's own modules on the seeded weights of
    oracle.speaker.synth_speaker_weights.  The tiny cases compose the reference's ResNet / ASP classes with fewer, narrower
    blocks through the reference's own ResNet293_based.forward; the third case is ResNet293_based as shipped."""
    from zonos import speaker_cloning as ref

    from . import speaker as osp

    out = {}
    for name, (ip, blocks, seed, seconds, sr, ch) in SPEAKER_CASES.items():
        sd, lda = osp.synth_speaker_weights(ip, blocks, seed)
        if blocks == osp.RESNET293_BLOCKS and ip == 64:
            m = ref.ResNet293_based()
        else:
            m = ref.ResNet293_based.__new__(ref.ResNet293_based)
            torch.nn.Module.__init__(m)
            m.front = ref.ResNet(ip, ref.SimAMBasicBlock, list(blocks))
            m.pooling = ref.ASP(ip * ref.SimAMBasicBlock.expansion, 80)
            m.bottleneck = torch.nn.Linear(m.pooling.out_dim, 256)
            m.drop = None
        m.featCal = None
        m.load_state_dict(sd, strict=True)
        m.featCal = ref.logFbankCal()
        m.eval()
        lin = torch.nn.Linear(256, 128)
        lin.load_state_dict(lda)
        wav = speaker_wav(seed, seconds, sr, ch)
        # SpeakerEmbedding.prepare_input + forward, SpeakerEmbeddingLDA.forward (speaker_cloning.py:375-384, 409-411)
        x = wav.mean(0, keepdim=True)
        if sr != 16000:
            import torchaudio

            x = torchaudio.transforms.Resample(sr, 16000)(x)
        feat = m.featCal(x)
        emb = m(x).to(torch.float32)
        out[name] = dict(in_planes=ip, blocks=tuple(blocks), seed=seed, seconds=seconds, sample_rate=sr, channels=ch,
                         wav16k=x.clone(), feat=feat.clone(), emb=emb.clone(), lda=lin(emb).clone(),
                         weights_checksum=float(sum(float(v.double().abs().sum()) for v in sd.values())))
        print(name, "feat", tuple(feat.shape), "emb |max|", float(emb.abs().max()))
    torch.save(out, os.path.join(OUT, "speaker.pt"))


def gen_model_real():
    """Real-shape model file: the reference's bf16 logits, plus the same prefill in fp32 (oracle port, which the CPU test
    pins bit-exactly to the reference at this shape): |bf16 reference - fp32| is the reference's own rounding noise, the
    yardstick for the 26-layer gates."""
    from . import delay as odelay
    from . import generate as ogen

    gen_model(build_reference_model("real"), "real", MODEL_CASES_REAL)
    path = os.path.join(OUT, "model_real.pt")
    g = torch.load(path, weights_only=False)
    for name, (B, Lc, P, N, sp, cfg) in MODEL_CASES_REAL.items():
        W32 = {k: v.float() for k, v in shapes.synth_transformer_weights("real", seed=0).items()}
        cond = shapes.synth_conditioning(B, Lc, 2048, seed=1).float()
        codes = torch.full((B, 9, P + N), -1, dtype=torch.int64)
        codes[..., :P] = torch.randint(0, 1024, (B, 9, P), generator=torch.Generator().manual_seed(2))
        src = ogen.ModelLogits(W32, "real", cond, Lc + P + N + 9, cfg)
        g["cases"][name]["logits_fp32_prefill"] = src.prefill(odelay.apply_delay_pattern(codes, 1025)[..., :P + 1])
    torch.save(g, path)


def main():
    os.makedirs(OUT, exist_ok=True)
    import_reference()
    torch.set_grad_enabled(False)
    if "--dac-encode" in sys.argv:
        gen_dac_encode()
        print("dac_encode.pt", os.path.getsize(os.path.join(OUT, "dac_encode.pt")))
        return
    if "--speaker" in sys.argv:
        gen_speaker()
        print("speaker.pt", os.path.getsize(os.path.join(OUT, "speaker.pt")))
        return
    if "--conditioning" in sys.argv:
        gen_conditioning()
        print("conditioning.pt", os.path.getsize(os.path.join(OUT, "conditioning.pt")))
        return
    if "--real" in sys.argv:  # only the real-shape model file (builds the 1.6 B-parameter reference model on the CPU)
        gen_model_real()
        print("model_real.pt", os.path.getsize(os.path.join(OUT, "model_real.pt")))
        return
    gen_delay()
    gen_sampler()
    model = build_reference_model("tiny")
    gen_eos(model)
    gen_model(model, "tiny")
    gen_dac()
    gen_postproc()
    gen_conditioning()
    gen_dac_encode()
    gen_speaker()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    sys.exit(main())
