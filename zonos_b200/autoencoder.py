"""DACAutoencoder.decode on B200 (zonos/autoencoder.py:12-19,44-47) -- see zonos_b200/dac.py for the kernels' host side."""
import torch


class DACAutoencoder:
    """Same attributes as the reference wrapper (num_codebooks / codebook_size / sampling_rate, decode).  Weights are
    the HF `descript/dac_44khz` decoder + quantizer tensors, given as a state dict (load_dac_state_dict) because this
    package does not download checkpoints."""

    def __init__(self):
        self.codebook_size = 1024
        self.num_codebooks = 9
        self.sampling_rate = 44100
        self._decoder = None
        self._device = torch.device("cpu")

    def to(self, device):
        self._device = torch.device(device)
        if self._decoder is not None:
            self._decoder.to(self._device)
        return self

    def load_dac_state_dict(self, sd: dict):
        from .dac import DacDecoderB200

        self._decoder = DacDecoderB200(sd, self._device)
        return self

    def decode(self, codes: torch.Tensor) -> torch.Tensor:
        """codes int64 [B, 9, T] -> fp32 [B, 1, 512*T]."""
        assert codes.shape[1] == self.num_codebooks, f"Expected {self.num_codebooks} codebooks, got {codes.shape[1]}"
        if self._decoder is None:
            raise RuntimeError("DAC weights not loaded: call autoencoder.load_dac_state_dict(state_dict) first")
        return self._decoder.decode(codes)
