"""Speaker encoder on the device (csrc/speaker.cu): the reference's `zonos/speaker_cloning.py` API --
`SpeakerEmbedding(ckpt_path, device)`, `SpeakerEmbeddingLDA(device)`, `.forward(wav, sample_rate)` -> `(emb, lda(emb))`
(:355-411) -- over the C-ABI kernels: resample to 16 kHz, log mel filterbank, ResNet293 of SimAM blocks (conv + folded
BatchNorm + ReLU, SimAM + shortcut + ReLU), attentive statistics pooling, bottleneck and LDA Linear layers.  fp32 like the
reference; torch only owns the buffers and there is no CPU path.

Weights are the reference's state dicts (`ResNet293_SimAM_ASP_base.pt`, `ResNet293_SimAM_ASP_base_LDA-128.pt` of
Zyphra/Zonos-v0.1-speaker-embedding, :390-405): pass them (`state_dict=`, `lda_state_dict=`) or let the constructor fetch them
like the reference does; offline that raises with the repository and file names.
"""
import ctypes
import math

import torch

from ._lib import check, lib, stream_ptr

_p = lambda t: ctypes.c_void_p(0 if t is None else t.data_ptr())
BN_EPS = 1e-5
SPEAKER_REPO = "Zyphra/Zonos-v0.1-speaker-embedding"


def sinc_resample_kernel(orig_freq: int, new_freq: int, lowpass_filter_width: int = 6, rolloff: float = 0.99):
    """The FIR bank of torchaudio.transforms.Resample (sinc_interp_hann; the defaults SpeakerEmbedding._get_resampler uses,
    speaker_cloning.py:371-373), evaluated in fp64 on the host: -> (fp32 [new'][2*width + orig'], width, orig', new')."""
    gcd = math.gcd(int(orig_freq), int(new_freq))
    orig, new = int(orig_freq) // gcd, int(new_freq) // gcd
    base = min(orig, new) * rolloff
    width = math.ceil(lowpass_filter_width * orig / base)
    idx = torch.arange(-width, width + orig, dtype=torch.float64)[None, :] / orig
    t = (torch.arange(0, -new, -1, dtype=torch.float64)[:, None] / new + idx) * base
    t = t.clamp(-lowpass_filter_width, lowpass_filter_width)
    window = torch.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t = t * math.pi
    kern = torch.where(t == 0, torch.ones_like(t), t.sin() / t) * window * (base / orig)
    return kern.to(torch.float32).contiguous(), width, orig, new


def mel_filterbank(n_freqs: int = 257, sample_rate: int = 16000, n_mels: int = 80) -> torch.Tensor:
    """Triangular HTK mel filters without area normalisation (MelSpectrogram's defaults, speaker_cloning.py:22-28): [257][80]."""
    hz2mel = lambda f: 2595.0 * math.log10(1.0 + f / 700.0)
    all_freqs = torch.linspace(0, sample_rate // 2, n_freqs)
    m_pts = torch.linspace(hz2mel(0.0), hz2mel(sample_rate / 2.0), n_mels + 2)
    f_pts = 700.0 * (10.0 ** (m_pts / 2595.0) - 1.0)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts[None, :] - all_freqs[:, None]
    return torch.clamp(torch.min(-slopes[:, :-2] / f_diff[:-1], slopes[:, 2:] / f_diff[1:]), min=0.0).contiguous()


class ResNet293B200:
    """`ResNet293_based` (speaker_cloning.py:198-223) behind its feature front end: weights repacked once (conv filters to
    [C_in][k*k][C_out], eval-mode BatchNorm to scale / shift), every layer one C-ABI launch."""

    def __init__(self, sd: dict, device):
        self.device = torch.device(device)
        f = lambda k: sd[k].detach().to(self.device, torch.float32).contiguous()

        def conv(k):  # [C_out, C_in, kh, kw] -> [C_in][kh*kw][C_out]
            w = f(k)
            co, ci, kh, kw = w.shape
            assert kh == kw and kh in (1, 3)
            return w.reshape(co, ci, kh * kw).permute(1, 2, 0).contiguous(), kh

        def bn(k):
            scale = f(k + ".weight") / torch.sqrt(f(k + ".running_var") + BN_EPS)
            return scale.contiguous(), (f(k + ".bias") - f(k + ".running_mean") * scale).contiguous()

        self.stem = (conv("front.conv1.weight"), bn("front.bn1"))
        self.blocks = []
        for li in range(1, 5):
            bi = 0
            while f"front.layer{li}.{bi}.conv1.weight" in sd:
                p = f"front.layer{li}.{bi}"
                blk = dict(c1=conv(p + ".conv1.weight"), b1=bn(p + ".bn1"), c2=conv(p + ".conv2.weight"), b2=bn(p + ".bn2"),
                           stride=(1 if li == 1 else 2) if bi == 0 else 1, down=None)
                if p + ".downsample.0.weight" in sd:
                    blk["down"] = (conv(p + ".downsample.0.weight"), bn(p + ".downsample.1"))
                self.blocks.append(blk)
                bi += 1
        lin = lambda k: (f(k + ".weight").reshape(sd[k + ".weight"].shape[0], -1).t().contiguous()[:, None, :].contiguous(),
                         f(k + ".bias"))  # Linear / 1x1 Conv1d [C_out, C_in(,1)] -> [C_in][1][C_out]
        self.att0 = lin("pooling.attention.0")
        self.att_bn = bn("pooling.attention.2")
        self.att3 = lin("pooling.attention.3")
        self.bottleneck = lin("bottleneck")
        self.window = torch.hann_window(400, periodic=True).to(self.device)
        self.mel_fb = mel_filterbank().to(self.device)

    # ---- launches
    def _conv2d(self, x, cw, bnp, stride, relu):
        (w, k), (scale, shift) = cw, bnp
        B, C_in, H, W = x.shape
        C_out, pad = w.shape[2], k // 2
        Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
        out = torch.empty(B, C_out, Ho, Wo, dtype=torch.float32, device=self.device)
        check(lib().zb200_conv2d_f32(_p(x), _p(w), _p(scale), _p(shift), _p(out), B, C_in, H, W, C_out, k, stride, int(relu),
                                     stream_ptr()), "conv2d_f32")
        return out

    def _conv1d(self, x, wb):
        w, b = wb
        B, C_in, T = x.shape
        out = torch.empty(B, w.shape[2], T, dtype=torch.float32, device=self.device)
        check(lib().zb200_conv1d_f32(_p(x), _p(w), _p(b), _p(None), _p(out), B, C_in, w.shape[2], T, T, 1, 1, 1, 0, stream_ptr()),
              "conv1d_f32")
        return out

    def features(self, wav16k: torch.Tensor) -> torch.Tensor:
        """logFbankCal.forward (:30-34): fp32 [B, n] at 16 kHz -> [B, 80, 1 + n // 160]."""
        wav16k = wav16k.to(self.device, torch.float32).contiguous()
        B, n = wav16k.shape
        T = 1 + n // 160
        out = torch.empty(B, 80, T, dtype=torch.float32, device=self.device)
        check(lib().zb200_logfbank_f32(_p(wav16k), _p(self.window), _p(self.mel_fb), _p(out), B, n, T, stream_ptr()), "logfbank_f32")
        return out

    def embed(self, feat: torch.Tensor) -> torch.Tensor:
        """ResNet.forward + ASP + bottleneck (:185-191, 52-60, 219-222) on features [B, 80, T] -> [B, embd_dim]."""
        x = self._conv2d(feat[:, None].contiguous(), *self.stem, 1, True)
        for blk in self.blocks:
            h = self._conv2d(x, blk["c1"], blk["b1"], blk["stride"], True)
            h = self._conv2d(h, blk["c2"], blk["b2"], 1, False)
            sc = x if blk["down"] is None else self._conv2d(x, *blk["down"], blk["stride"], False)
            B, C, H, W = h.shape
            x = torch.empty_like(h)
            check(lib().zb200_simam_residual_f32(_p(h), _p(sc), _p(x), B * C, H * W, ctypes.c_float(1e-4), stream_ptr()),
                  "simam_residual_f32")
        B, C, H, W = x.shape
        x = x.view(B, C * H, W)
        a = self._conv1d(x, self.att0)
        h = torch.empty_like(a)
        check(lib().zb200_relu_affine_f32(_p(a), _p(self.att_bn[0]), _p(self.att_bn[1]), _p(h), B, a.shape[1], W, stream_ptr()),
              "relu_affine_f32")
        logits = self._conv1d(h, self.att3)
        pooled = torch.empty(B, 2 * C * H, dtype=torch.float32, device=self.device)
        check(lib().zb200_asp_pool_f32(_p(x), _p(logits), _p(pooled), B, C * H, W, stream_ptr()), "asp_pool_f32")
        return self._linear(pooled, self.bottleneck)

    def _linear(self, x, wb):
        w, b = wb  # [C_in][1][C_out]
        out = torch.empty(x.shape[0], w.shape[2], dtype=torch.float32, device=self.device)
        check(lib().zb200_linear_f32(_p(x.contiguous()), _p(w), _p(b), _p(out), x.shape[0], w.shape[0], w.shape[2], stream_ptr()),
              "linear_f32")
        return out

    def __call__(self, wav16k: torch.Tensor) -> torch.Tensor:
        return self.embed(self.features(wav16k))


def _hub_state_dict(filename: str):
    try:
        from huggingface_hub import hf_hub_download

        return torch.load(hf_hub_download(repo_id=SPEAKER_REPO, filename=filename), weights_only=True, map_location="cpu")
    except Exception as e:
        raise RuntimeError(f"speaker encoder weights {SPEAKER_REPO}/{filename} are not available (offline?); pass the state dict "
                           "to the constructor") from e


class SpeakerEmbedding(torch.nn.Module):
    """speaker_cloning.py:355-384.  `state_dict`: the contents of ResNet293_SimAM_ASP_base.pt (else `ckpt_path` is loaded)."""

    def __init__(self, ckpt_path: str = "ResNet293_SimAM_ASP_base.pt", device: str = "cuda", state_dict: dict | None = None):
        super().__init__()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("zonos_b200's speaker encoder runs on CUDA only (there is no CPU fallback)")
        if state_dict is None:
            state_dict = torch.load(ckpt_path, weights_only=True, mmap=True, map_location="cpu")
        self.model = ResNet293B200(state_dict, self.device)
        self._resamplers = {}

    dtype = torch.float32

    def prepare_input(self, wav: torch.Tensor, sample_rate: int) -> torch.Tensor:
        """Mono mix + resample to 16 kHz (:375-380) -> fp32 [1, n] on the device."""
        assert wav.ndim < 3
        wav = wav.to(self.device, torch.float32)
        wav = wav.mean(0, keepdim=True) if wav.ndim == 2 else wav[None]
        if sample_rate == 16000:
            return wav.contiguous()
        if sample_rate not in self._resamplers:
            kern, width, orig, new = sinc_resample_kernel(sample_rate, 16000)
            self._resamplers[sample_rate] = (kern.to(self.device), width, orig, new)
        kern, width, orig, new = self._resamplers[sample_rate]
        wav = wav.contiguous()
        B, n = wav.shape
        n_out = math.ceil(new * n / orig)
        out = torch.empty(B, n_out, dtype=torch.float32, device=self.device)
        check(lib().zb200_resample_f32(_p(wav), _p(kern), _p(out), B, n, n_out, orig, new, width, stream_ptr()), "resample_f32")
        return out

    @torch.no_grad()
    def forward(self, wav: torch.Tensor, sample_rate: int):
        return self.model(self.prepare_input(wav, sample_rate))


class SpeakerEmbeddingLDA(torch.nn.Module):
    """speaker_cloning.py:387-411: forward(wav, sample_rate) -> (emb fp32 [1, 256], lda(emb) fp32 [1, 128])."""

    def __init__(self, device: str = "cuda", state_dict: dict | None = None, lda_state_dict: dict | None = None):
        super().__init__()
        self.device = torch.device(device)
        if state_dict is None:
            state_dict = _hub_state_dict("ResNet293_SimAM_ASP_base.pt")
        if lda_state_dict is None:
            lda_state_dict = _hub_state_dict("ResNet293_SimAM_ASP_base_LDA-128.pt")
        self.model = SpeakerEmbedding(device=device, state_dict=state_dict)
        w = lda_state_dict["weight"].detach().to(self.device, torch.float32)
        self.lda = (w.t().contiguous()[:, None, :].contiguous(), lda_state_dict["bias"].detach().to(self.device, torch.float32).contiguous())

    @torch.no_grad()
    def forward(self, wav: torch.Tensor, sample_rate: int):
        emb = self.model(wav, sample_rate).to(torch.float32)
        return emb, self.model.model._linear(emb, self.lda)
