"""codes_to_wavs post-processing parity (GPU): zb200_codes_postprocess through the C ABI against oracle/postproc.py
(pinned to the reference for trim + fades, to BS.1770 for the loudness) on ragged batches."""
import numpy as np
import pytest
import torch

from oracle import postproc as opp
from oracle import scenarios as sc
from oracle import shapes

pytestmark = pytest.mark.gpu

# Loudness is a floating-point path: the oracle sums float32 squares pairwise (numpy), the kernel in fp64, and the
# K-weighted signal is stitched from 1024-sample chunks; both effects are ~1e-7 relative on the block energies.
LUFS_TOL_DB = 1e-4
# The gain enters every output sample once: 1e-4 dB is 1.2e-5 relative; measured deviations are a few fp32 ulps.
WAV_RTOL = 2e-5


def _post(device="cuda"):
    from zonos_b200.postproc import WavPostprocessor

    return WavPostprocessor(device)


def _batch(names):
    wavs = [sc.postproc_wav(n) for n in names]
    n = [w.shape[1] for w in wavs]
    buf = torch.zeros(len(wavs), (max(n) + 3) // 4 * 4)
    for i, w in enumerate(wavs):
        buf[i, : n[i]] = w[0]
    return wavs, n, buf.cuda()


def test_trim_and_fades_bit_exact_against_reference_golden(golden):
    """No loudness step (what the reference itself does without pyloudnorm): integer bounds and every sample exact."""
    g = golden("postproc.pt")
    names = list(sc.POSTPROC_GOLDEN)
    _, n, buf = _batch(names)
    out, bounds, lufs, gain = _post().run(buf, n, loudness=False)
    torch.cuda.synchronize()
    for i, name in enumerate(names):
        s, e = bounds[i].tolist()
        assert e - s == g[name]["trimmed_len"] == g[name]["wav"].shape[1]
        assert torch.equal(out[i, : e - s].cpu(), g[name]["wav"][0])
        assert not out[i, e - s :].any()
        assert float(gain[i]) == 1.0 and torch.isnan(lufs[i])


def test_full_chain_against_oracle_on_a_ragged_batch():
    names = [k for k in sc.POSTPROC_CASES if k != "silent"]
    wavs, n, buf = _batch(names)
    out, bounds, lufs, gain = _post().run(buf, n)
    torch.cuda.synchronize()
    for i, name in enumerate(names):
        ref = opp.post(wavs[i])
        assert bounds[i].tolist() == [ref["start"], ref["end"]], name
        if ref["lufs"] is None:
            assert torch.isnan(lufs[i]) and float(gain[i]) == 1.0
            assert torch.equal(out[i, : ref["end"] - ref["start"]].cpu(), ref["wav"][0]), name
        else:
            assert abs(float(lufs[i]) - ref["lufs"]) < LUFS_TOL_DB, (name, float(lufs[i]), ref["lufs"])
            torch.testing.assert_close(out[i, : ref["end"] - ref["start"]].cpu(), ref["wav"][0], rtol=WAV_RTOL,
                                       atol=1e-9)
        assert not out[i, ref["end"] - ref["start"] :].any()


def test_silent_utterance_follows_the_reference_into_infinity():
    """All gating blocks below -70 LUFS: pyloudnorm returns -inf, the reference's gain is +inf."""
    wavs, n, buf = _batch(["silent", "lead_tail"])
    out, bounds, lufs, gain = _post().run(buf, n)
    ref = opp.post(wavs[0])
    assert ref["lufs"] == -np.inf and float(lufs[0]) == -np.inf and float(gain[0]) == np.inf
    assert bounds[0].tolist() == [ref["start"], ref["end"]]
    o = out[0, : ref["end"] - ref["start"]].cpu()
    assert torch.equal(torch.isnan(o), torch.isnan(ref["wav"][0]))
    assert torch.equal(torch.nan_to_num(o), torch.nan_to_num(ref["wav"][0]))
    ref1 = opp.post(wavs[1])  # the neighbour in the batch is untouched by it
    assert abs(float(lufs[1]) - ref1["lufs"]) < LUFS_TOL_DB


def test_long_utterance_many_chunks():
    """12 s of audio: 517 chunks of 1024 per utterance, > 100 gating blocks; loudness within 1e-4 dB of the oracle."""
    g = torch.Generator().manual_seed(5)
    n = [44100 * 12 + 123, 44100 * 7]
    t = torch.arange(n[0], dtype=torch.float64) / 44100
    a = (0.2 * torch.sin(2 * torch.pi * 220 * t) * (0.5 + 0.5 * torch.sin(2 * torch.pi * 0.7 * t))).float()
    a += 0.02 * torch.randn(n[0], generator=g)
    b = 0.5 * torch.randn(n[1], generator=g)
    buf = torch.zeros(2, (n[0] + 3) // 4 * 4)
    buf[0, : n[0]], buf[1, : n[1]] = a, b
    out, bounds, lufs, gain = _post().run(buf.cuda(), n)
    for i, w in enumerate((a, b)):
        ref = opp.post(w.unsqueeze(0))
        assert abs(float(lufs[i]) - ref["lufs"]) < LUFS_TOL_DB
        assert bounds[i].tolist() == [ref["start"], ref["end"]]
        torch.testing.assert_close(out[i, : ref["end"] - ref["start"]].cpu(), ref["wav"][0], rtol=WAV_RTOL, atol=1e-9)


def test_997hz_known_answer_on_device():
    t = np.arange(44100 * 10) / 44100
    x = torch.from_numpy(np.sin(2 * np.pi * 997.0 * t).astype(np.float32)).unsqueeze(0)
    _, _, lufs, gain = _post().run(x.cuda().contiguous(), [x.shape[1]])
    assert abs(float(lufs[0]) - (-3.01 + 20 * np.log10(0.995))) < 5e-3
    assert abs(float(gain[0]) - 10 ** ((-23.0 - float(lufs[0])) / 20)) < 1e-6


def test_codes_to_wavs_api_matches_decode_plus_oracle():
    """The public call: ragged list of codes -> decode (grouped by length) -> post-processing, against
    autoencoder.decode of each utterance on its own followed by the oracle chain."""
    from zonos_b200.autoencoder import DACAutoencoder

    ae = DACAutoencoder().to("cuda")
    ae.load_dac_state_dict(shapes.synth_dac_weights(seed=7, arch=shapes.DAC_TINY))
    g = torch.Generator().manual_seed(11)
    codes = [torch.randint(0, 1024, (1, 9, t), generator=g) for t in (12, 30, 12, 0, 200)]
    codes[1] = codes[1][0]  # [9, T] form
    wavs = ae.codes_to_wavs(codes)
    assert len(wavs) == 4 and all(w.device.type == "cpu" and w.dim() == 2 and w.shape[0] == 1 for w in wavs)
    kept = [c if c.dim() == 3 else c.unsqueeze(0) for c in codes if c.shape[-1] > 0]
    for w, c in zip(wavs, kept):
        single = ae.decode(c.cuda()).cpu().squeeze(0)
        ref = opp.post(single)
        assert w.shape == ref["wav"].shape
        torch.testing.assert_close(w, ref["wav"], rtol=WAV_RTOL, atol=1e-9)
    # trim_silence / normalize_loudness on their own
    x = sc.postproc_wav("lead_tail")
    s, e = opp.trim_bounds(x)
    assert torch.equal(ae.trim_silence(x), x[:, s:e])
    y, _ = opp.normalize_loudness(x, 44100, -19.0)
    torch.testing.assert_close(ae.normalize_loudness(x, 44100), y, rtol=WAV_RTOL, atol=1e-9)


def test_rejects_bad_arguments():
    pp = _post()
    x = torch.zeros(1, 1024, device="cuda")
    with pytest.raises(AssertionError):
        pp.run(x, [2048])
    from zonos_b200 import _lib
    import ctypes

    pp._ensure_tables(1)
    plan = pp._plan(False)
    z = ctypes.c_void_p(0)
    with pytest.raises(ValueError):
        _lib.check(_lib.lib().zb200_codes_postprocess(ctypes.byref(plan), z, ctypes.c_int64(1024), z, z, z, 1, 1024, z,
                                                      z, z, z, z, ctypes.c_int64(0), z), "codes_postprocess")


def test_save_codes_and_batched_tensor_form(tmp_path):
    """codes as one [B, 9, T] tensor (one batched decode), waveforms kept on the device, and save_codes' files."""
    from zonos_b200.autoencoder import DACAutoencoder

    ae = DACAutoencoder().to("cuda")
    ae.load_dac_state_dict(shapes.synth_dac_weights(seed=7, arch=shapes.DAC_TINY))
    codes = torch.randint(0, 1024, (3, 9, 25), generator=torch.Generator().manual_seed(2))
    on_dev = ae.codes_to_wavs(codes.cuda(), device_output=True)
    on_host = ae.codes_to_wavs(codes)
    assert len(on_dev) == 3 and all(w.is_cuda for w in on_dev)
    for d, h, c in zip(on_dev, on_host, codes):
        assert torch.equal(d.cpu(), h)
        ref = opp.post(ae.decode(c[None].cuda()).cpu().squeeze(0))
        torch.testing.assert_close(h, ref["wav"], rtol=WAV_RTOL, atol=1e-9)
    paths = [str(tmp_path / f"u{i}.wav") for i in range(3)]
    ae.save_codes(paths, codes)
    for p, h in zip(paths, on_host):
        raw = open(p, "rb").read()
        assert raw[:4] == b"RIFF" and np.array_equal(np.frombuffer(raw[44:], dtype="<f4"), h[0].numpy())
    with pytest.raises(AssertionError):
        ae.save_codes(paths[:2], codes)
    with pytest.raises(ValueError):
        ae.codes_to_wavs(torch.zeros(2, 2, 9, 4, dtype=torch.long))
