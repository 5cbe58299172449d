"""codes_to_wavs post-processing, CPU side: the oracle against the reference's golden vectors and BS.1770's published
numbers, and the host logic of zonos_b200/postproc.py (block tables, chunked-filter transition)."""
import numpy as np
import pytest
import scipy.signal
import torch

from oracle import postproc as opp
from oracle import scenarios as sc


@pytest.mark.parametrize("name", sc.POSTPROC_GOLDEN)
def test_trim_and_fades_bit_exact_against_reference(golden, name):
    g = golden("postproc.pt")[name]
    wav = sc.postproc_wav(name)
    assert float(wav.double().sum()) == g["in_sum"]  # the seeded input is the one the reference saw
    r = opp.post(wav, loudness=False)
    assert torch.equal(r["wav"], g["wav"])
    assert r["end"] - r["start"] == g["trimmed_len"]


def test_k_weighting_matches_bs1770_table():
    # ITU-R BS.1770-4 table 1 / table 2 (48 kHz).  pyloudnorm designs the filters from (G, Q, fc) instead of using the
    # table: the shelf agrees to 1e-4, its high-pass numerator is 0.995 * [1, -2, 1] (a known property of that library).
    (b1, a1), (b2, a2) = opp.k_weighting(48000)
    np.testing.assert_allclose(b1, [1.53512485958697, -2.69169618940638, 1.19839281085285], atol=2e-4)
    np.testing.assert_allclose(a1, [1.0, -1.69065929318241, 0.73248077421585], atol=2e-4)
    np.testing.assert_allclose(a2, [1.0, -1.99004745483398, 0.99007225036621], atol=5e-5)
    np.testing.assert_allclose(b2 / b2[0], [1.0, -2.0, 1.0], atol=1e-12)
    assert abs(b2[0] - 0.995) < 1e-3


@pytest.mark.parametrize("rate", [44100, 48000])
def test_known_answer_997hz_sine(rate):
    # BS.1770-4 annex: a 0 dBFS 997 Hz sine in one front channel reads -3.01 LKFS; pyloudnorm's high-pass pass-band
    # gain (0.995 -> -0.043 dB) puts its own reading at -3.05.
    t = np.arange(rate * 10) / rate
    x = np.sin(2 * np.pi * 997.0 * t).astype(np.float32)
    lufs = opp.integrated_loudness(x[:, None], rate)
    assert abs(lufs - (-3.01 + 20 * np.log10(0.995))) < 5e-3
    # linearity of the meter: -20 dB in, -20 LU out
    assert abs(opp.integrated_loudness(0.1 * x[:, None], rate) - (lufs - 20.0)) < 1e-4


def test_too_short_audio_is_returned_unchanged():
    wav = sc.postproc_wav("three_frames")
    out, lufs = opp.normalize_loudness(wav, 44100, -23.0)
    assert lufs is None and torch.equal(out, wav)


def test_host_block_tables_follow_the_meter():
    from zonos_b200 import postproc as hp

    for n in (4410, 4608, 88200, 88201, 92160, 122880, 44100 * 30 + 7):
        sel, nb = hp.loudness_blocks(n, 44100)
        bs = 0.400 if n > 2.0 * 44100 else 0.100
        assert sel == (0 if bs == 0.400 else 1)
        assert nb == int(np.round(((n / 44100 - bs) / (bs * 0.25))) + 1)
    assert hp.loudness_blocks(4409, 44100) == (-1, 0)
    lo, hi = hp.block_bounds(0.400, 44100, 300)
    assert lo[:3] == [0, 4410, 8820] and hi[0] == 17640
    assert all(h - l in (17639, 17640, 17641) for l, h in zip(lo, hi))


def test_host_chunk_transition_reproduces_the_unchunked_filter():
    """The kernels filter chunks independently and stitch them with M^chunk (csrc/postproc.cu); the same stitching
    in numpy must give scipy's answer."""
    from zonos_b200 import postproc as hp

    shelf, hpf = hp.k_weighting(44100)
    for (b, a), (bo, ao) in zip((shelf, hpf), opp.k_weighting(44100)):
        np.testing.assert_array_equal(b, bo)
        np.testing.assert_array_equal(a, ao)
    chunk = 256
    M = hp.chunk_transition(shelf, hpf, chunk)
    x = np.random.default_rng(0).standard_normal(chunk * 9 + 57)
    full = scipy.signal.lfilter(*hpf, scipy.signal.lfilter(*shelf, x))

    def run(seg, s):
        y = np.empty_like(seg)
        s = list(s)
        (b0a, b1a, b2a), (_, a1a, a2a) = shelf
        (b0b, b1b, b2b), (_, a1b, a2b) = hpf
        for i, v in enumerate(seg):
            ya = b0a * v + s[0]
            s[0] = b1a * v - a1a * ya + s[1]
            s[1] = b2a * v - a2a * ya
            yb = b0b * ya + s[2]
            s[2] = b1b * ya - a1b * yb + s[3]
            s[3] = b2b * ya - a2b * yb
            y[i] = yb
        return y, np.array(s)

    segs = [x[i : i + chunk] for i in range(0, len(x), chunk)]
    ends = [run(seg, np.zeros(4))[1] for seg in segs]
    s_in, out = np.zeros(4), []
    for seg, e in zip(segs, ends):
        out.append(run(seg, s_in)[0])
        s_in = M @ s_in + e
    np.testing.assert_allclose(np.concatenate(out), full, rtol=0, atol=1e-10)


def test_wav_writer_header(tmp_path):
    import wave  # noqa: F401  (the stdlib reader rejects format 3; parse the header by hand)
    import struct

    from zonos_b200.postproc import write_wav_f32

    w = torch.linspace(-1, 1, 1000).unsqueeze(0)
    p = tmp_path / "a.wav"
    write_wav_f32(str(p), w, 44100)
    raw = p.read_bytes()
    assert raw[:4] == b"RIFF" and raw[8:16] == b"WAVEfmt "
    fmt, ch, rate, _, _, bits = struct.unpack("<HHIIHH", raw[20:36])
    assert (fmt, ch, rate, bits) == (3, 1, 44100, 32)
    assert np.array_equal(np.frombuffer(raw[44:], dtype="<f4"), w[0].numpy())


def test_host_edge_tables_cover_every_block():
    from zonos_b200 import postproc as hp

    for bs in hp.WavPostprocessor.BLOCK_SIZES:
        lo, hi = hp.block_bounds(bs, 44100, 700)
        edges, e_lo, e_hi = hp.edge_tables(lo, hi)
        assert edges == sorted(set(edges)) and edges[0] == 0
        for j in (0, 1, 5, 333, 699):
            assert edges[e_lo[j]] == lo[j] and edges[e_hi[j]] == hi[j] and e_lo[j] < e_hi[j]
        for chunk in (512, 2048):
            assert hp.max_pieces(edges, chunk) <= hp.MAX_PIECES
