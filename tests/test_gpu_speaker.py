"""Speaker encoder on the device (zonos_b200/speaker_cloning.py over csrc/speaker.cu) against the golden
outputs of the reference's own modules (tests/golden/speaker.pt, made by oracle/gen_golden.py --speaker) and against the
CPU oracle on a fresh input.  fp32 on both sides: the differences are summation order (and, in the features, the
reference's FFT against a direct DFT with fp64 accumulators, which the log of small mel energies amplifies)."""
import pytest
import torch

from oracle import speaker as osp
from oracle.gen_golden import SPEAKER_CASES, speaker_wav

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", list(SPEAKER_CASES))
def test_speaker_encoder_vs_reference_golden(golden, name):
    from zonos_b200.speaker_cloning import SpeakerEmbeddingLDA

    ip, blocks, seed, seconds, sr, ch = SPEAKER_CASES[name]
    c = golden("speaker.pt")[name]
    sd, lda = osp.synth_speaker_weights(ip, blocks, seed)
    assert float(sum(float(v.double().abs().sum()) for v in sd.values())) == pytest.approx(c["weights_checksum"], rel=1e-9)
    enc = SpeakerEmbeddingLDA("cuda", state_dict=sd, lda_state_dict=lda)
    net = enc.model.model
    wav = speaker_wav(seed, seconds, sr, ch)
    # resampler (44.1 kHz stereo case) against torchaudio's output
    x = enc.model.prepare_input(wav, sr)
    assert x.shape == c["wav16k"].shape and float((x.cpu() - c["wav16k"]).abs().max()) < 1e-5
    # features on the golden 16 kHz signal
    feat = net.features(c["wav16k"].cuda())
    d = (feat.cpu() - c["feat"]).abs()
    print(f"{name}: log-mel max|d| {float(d.max()):.2e} mean|d| {float(d.mean()):.2e}")
    assert float(d.max()) < 5e-4 and float(d.mean()) < 1e-5
    # the network on the golden features: summation order only
    scale = float(c["emb"].abs().max())
    emb = net.embed(c["feat"].cuda())
    e1 = float((emb.cpu() - c["emb"]).abs().max()) / scale
    # whole path from the raw signal, through the public API
    emb2, out2 = enc(wav, sr)
    e2 = float((emb2.cpu() - c["emb"]).abs().max()) / scale
    e3 = float((out2.cpu() - c["lda"]).abs().max()) / float(c["lda"].abs().max())
    print(f"{name}: embedding rel. max error {e1:.2e} (golden features), {e2:.2e} / LDA {e3:.2e} (from the raw signal)")
    assert emb2.shape == (1, 256) and out2.shape == (1, 128) and emb2.dtype == torch.float32
    assert e1 < 1e-4 and e2 < 1e-3 and e3 < 1e-3


def test_speaker_encoder_batch_vs_oracle():
    """A batch of two longer signals through features + network against the CPU oracle (narrow stack)."""
    from zonos_b200.speaker_cloning import ResNet293B200

    ip, blocks = 16, (2, 2, 3, 2)
    sd, _ = osp.synth_speaker_weights(ip, blocks, seed=21)
    net = ResNet293B200(sd, "cuda")
    wav = torch.cat([speaker_wav(31, 2.37, 16000, 1), speaker_wav(32, 2.37, 16000, 1)], 0)
    feat_ref = osp.log_fbank(wav)
    emb_ref = osp.embedding_from_features(feat_ref, sd, blocks)
    emb = net.embed(feat_ref.cuda())
    scale = float(emb_ref.abs().max())
    assert float((emb.cpu() - emb_ref).abs().max()) < 5e-4 * scale
    emb2 = net(wav.cuda())
    assert float((emb2.cpu() - emb_ref).abs().max()) < 1e-2 * scale
    assert float((emb_ref[0] - emb_ref[1]).abs().max()) > 1e-2 * scale  # the two signals do not collapse


def test_make_speaker_embedding_on_the_model(golden):
    """Zonos.make_speaker_embedding (zonos/model.py:90-95): bf16 [1, 1, 128] = the LDA output of the clip."""
    from oracle import shapes
    from zonos_b200.config import ZonosConfig
    from zonos_b200.model import Zonos
    from zonos_b200.speaker_cloning import SpeakerEmbeddingLDA

    ip, blocks, seed, seconds, sr, ch = SPEAKER_CASES["tiny_44k_stereo"]
    c = golden("speaker.pt")["tiny_44k_stereo"]
    sd, lda = osp.synth_speaker_weights(ip, blocks, seed)
    m = Zonos.from_state_dict(ZonosConfig.from_dict(shapes.reference_config_dict("tiny")), shapes.synth_transformer_weights("tiny"), "cuda")
    m.spk_clone_model = SpeakerEmbeddingLDA("cuda", state_dict=sd, lda_state_dict=lda)
    e = m.make_speaker_embedding(speaker_wav(seed, seconds, sr, ch), sr)
    assert e.shape == (1, 1, 128) and e.dtype == torch.bfloat16
    assert float((e.float().cpu()[0] - c["lda"]).abs().max()) < 2e-2 * float(c["lda"].abs().max())  # bf16 rounding of the output
