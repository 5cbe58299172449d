"""DACAutoencoder.decode / codes_to_wavs on B200 (zonos/autoencoder.py:12-19,44-47,172-267) -- see zonos_b200/dac.py
and zonos_b200/postproc.py for the kernels' host side."""
import logging

import torch

logger = logging.getLogger(__name__)


class DACAutoencoder:
    """Same attributes as the reference wrapper (num_codebooks / codebook_size / sampling_rate, decode, codes_to_wavs).

    Weights: like zonos/autoencoder.py:13-19 the constructor loads `descript/dac_44khz` through transformers -- local
    files first, then the hub.  Where neither is reachable (this build container, the GPU box) the reference would raise
    in __init__; here the error is kept and raised by the first decode(), so that tests and the benchmark can hand over a
    state dict with the HF key names instead (load_dac_state_dict: `decoder.*`, `quantizer.*` and, for encode(), `encoder.*`)."""

    PRETRAINED = "descript/dac_44khz"

    def __init__(self, load_pretrained: bool = True):
        self.codebook_size = 1024
        self.num_codebooks = 9
        self.sampling_rate = 44100
        self._decoder = None
        self._encoder = None
        self._sd = None  # HF state dict waiting for a CUDA device
        self._load_error = "no DAC weights given"
        self._postproc = {}
        self._device = torch.device("cpu")
        if load_pretrained:
            self._load_pretrained()

    def _load_pretrained(self):
        try:
            from transformers.models.dac import DacModel

            try:
                dac = DacModel.from_pretrained(self.PRETRAINED, local_files_only=True)
            except Exception:
                dac = DacModel.from_pretrained(self.PRETRAINED, local_files_only=False)
            self.codebook_size = dac.config.codebook_size
            self.num_codebooks = dac.quantizer.n_codebooks
            self.sampling_rate = dac.config.sampling_rate
            self._sd = {k: v.detach() for k, v in dac.state_dict().items() if k.startswith(("decoder.", "quantizer.", "encoder."))}
            self._load_error = None
        except Exception as e:  # offline and not cached: deferred to the first decode()
            self._load_error = f"{self.PRETRAINED} could not be loaded through transformers ({type(e).__name__}: {e})"

    # the reference moves the codec with `model.autoencoder.dac.to(device)` (zonos/model.py:80)
    @property
    def dac(self):
        return self

    @property
    def device(self) -> torch.device:
        return self._device

    def to(self, device):
        self._device = torch.device(device)
        if self._decoder is not None:
            self._decoder.to(self._device)
        return self

    def load_dac_state_dict(self, sd: dict):
        """HF DacModel key names (`decoder.*`, `quantizer.quantizers.{k}.*`); replaces whatever was loaded before."""
        self._sd, self._decoder, self._encoder, self._load_error = dict(sd), None, None, None
        if self._device.type == "cuda":
            self._build()
        return self

    def _build(self):
        from .dac import DacDecoderB200

        if self._sd is None:
            raise RuntimeError(f"DAC weights not loaded: {self._load_error}; call autoencoder.load_dac_state_dict(state_dict) "
                               "with the HF descript/dac_44khz tensors")
        if self._device.type != "cuda":
            raise RuntimeError("zonos_b200 runs on CUDA (B200) only; move the autoencoder with .to('cuda')")
        self._decoder = DacDecoderB200(self._sd, self._device)

    def preprocess(self, wav: torch.Tensor, sr: int) -> torch.Tensor:
        """zonos/autoencoder.py:21-25: resample to 44.1 kHz and zero-pad on the LEFT to a whole number of 512-sample frames."""
        import torchaudio

        wav = torchaudio.functional.resample(wav, sr, self.sampling_rate)
        pad = -wav.shape[-1] % 512
        return torch.nn.functional.pad(wav, (pad, 0))

    def encode(self, wav: torch.Tensor) -> torch.Tensor:
        """zonos/autoencoder.py:27-28: fp32 [B, 1, T] -> codes int64 [B, 9, T / 512] (HF DacModel.encode(...).audio_codes),
        fp32 kernels like the reference's un-autocast encode (zonos_b200/dac_encoder.py)."""
        if self._encoder is None:
            from .dac_encoder import DacEncoderB200

            if self._sd is None or "encoder.conv1.weight" not in self._sd:
                raise RuntimeError(f"DAC encoder weights not loaded ({self._load_error or 'the state dict has no encoder.* tensors'})")
            if self._device.type != "cuda":
                raise RuntimeError("zonos_b200 runs on CUDA (B200) only; move the autoencoder with .to('cuda')")
            self._encoder = DacEncoderB200(self._sd, self._device)
        return self._encoder.encode(wav)

    def load_prefix_audio(self, audio_path, device):
        """zonos/autoencoder.py:30-42."""
        import torchaudio

        wav, sr = torchaudio.load(audio_path)
        wav = self.preprocess(wav.mean(dim=0, keepdim=True), sr).to(device, dtype=torch.float32)
        return self.encode(wav.unsqueeze(0))

    def decode(self, codes: torch.Tensor) -> torch.Tensor:
        """codes int64 [B, 9, T] -> fp32 [B, 1, 512*T]."""
        assert codes.shape[1] == self.num_codebooks, f"Expected {self.num_codebooks} codebooks, got {codes.shape[1]}"
        if self._decoder is None:
            self._build()
        return self._decoder.decode(codes)

    # ---- post-processing (autoencoder.py:49-90,172-267), batched on the device -----------------------------------
    def _post(self, target_lufs: float):
        from .postproc import WavPostprocessor

        key = (str(self._device), float(target_lufs))
        if key not in self._postproc:
            self._postproc[key] = WavPostprocessor(self._device, self.sampling_rate, target_lufs)
        return self._postproc[key]

    def _padded(self, wav: torch.Tensor) -> torch.Tensor:
        """[1, n] (any device) -> contiguous fp32 [1, round_up(n, 4)] on the device."""
        n = wav.shape[1]
        buf = torch.zeros(1, (n + 3) // 4 * 4, dtype=torch.float32, device=self._device)
        buf[:, :n] = wav
        return buf

    def trim_silence(self, wav: torch.Tensor) -> torch.Tensor:
        """autoencoder.py:49-90 with its default threshold / frame size: [1, n] -> [1, end - start]."""
        assert wav.ndim == 2 and wav.shape[0] == 1, "Expected mono audio tensor"
        out, bounds, _, _ = self._post(-23.0).run(self._padded(wav), [wav.shape[1]], loudness=False, fades=False)
        start, end = bounds[0].tolist()
        return out[:, : end - start].to(wav.device)

    def normalize_loudness(self, audio: torch.Tensor, sr: int, target_lufs: float = -19.0) -> torch.Tensor:
        """autoencoder.py:172-186: audio * 10 ** ((target - integrated loudness) / 20); unchanged when too short."""
        assert sr == self.sampling_rate and audio.ndim == 2 and audio.shape[0] == 1
        _, _, _, gain = self._post(target_lufs).run(self._padded(audio), [audio.shape[1]], fades=False)
        return audio * gain.to(audio.device)

    def codes_to_wavs(self, codes, device_output: bool = False) -> list:
        """autoencoder.py:188-245: codes [B, 9, T] / [9, T] / a list of those -> list of fp32 [1, n_i] waveforms
        (loudness-normalised to -23 LUFS, silence-trimmed, faded).  Utterances of equal length are decoded as one batch
        (the decoder treats rows independently, so this equals the reference's one-by-one loop); the post-processing
        runs once over the ragged batch.  Returns CPU tensors like the reference unless device_output."""
        if isinstance(codes, (list, tuple)):
            code_list = [c.unsqueeze(0) if c.dim() == 2 else c for c in codes]
        elif codes.dim() == 2:
            code_list = [codes.unsqueeze(0)]
        elif codes.dim() == 3:
            code_list = [codes[i : i + 1] for i in range(codes.shape[0])]
        else:
            raise ValueError(f"Invalid shape for codes: {codes.shape}. Expected [num_codebooks, num_codes] or "
                             "[batch_size, num_codebooks, num_codes]")
        kept = []
        for c in code_list:
            if c.shape[2] == 0:
                logger.warning("Empty codebook detected, skipping decoding.")
                continue
            kept.extend(c[i : i + 1] for i in range(c.shape[0]))
        if not kept:
            return []
        by_len = {}
        for i, c in enumerate(kept):
            by_len.setdefault(c.shape[2], []).append(i)
        if len(by_len) == 1:
            wav = self.decode(torch.cat(kept, 0).to(self._device)).squeeze(1)
            n_samples = [wav.shape[1]] * len(kept)
        else:
            decoded = {}
            for idx in by_len.values():
                w = self.decode(torch.cat([kept[i] for i in idx], 0).to(self._device)).squeeze(1)
                for k, i in enumerate(idx):
                    decoded[i] = w[k]
            n_samples = [decoded[i].shape[0] for i in range(len(kept))]
            wav = torch.zeros(len(kept), (max(n_samples) + 3) // 4 * 4, dtype=torch.float32, device=self._device)
            for i in range(len(kept)):
                wav[i, : n_samples[i]] = decoded[i]
        out, bounds, _, _ = self._post(-23.0).run(wav.contiguous(), n_samples)
        if not device_output:
            host = torch.empty(out.shape, dtype=out.dtype, pin_memory=True)
            host.copy_(out, non_blocking=True)
            out = host
        b = bounds.tolist()  # synchronises
        return [out[i : i + 1, : e - s] for i, (s, e) in enumerate(b)]

    def save_codes(self, paths, codes) -> None:
        """autoencoder.py:247-268."""
        from .postproc import write_wav_f32

        if isinstance(paths, str):
            paths = [paths]
        wavs = self.codes_to_wavs(codes)
        assert len(paths) == len(wavs), f"Number of paths ({len(paths)}) must match number of codes ({len(wavs)})"
        for p, w in zip(paths, wavs):
            write_wav_f32(p, w, self.sampling_rate)
