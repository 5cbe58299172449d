"""Host side of the DAC encoder + residual vector quantiser (csrc/dac_encoder.cu).

DACAutoencoder.encode (zonos/autoencoder.py:27-28) -> HF DacModel.encode (modeling_dac.py:581-605): the reference runs it
in fp32, once per request, on the audio prefix.  Weights: the HF state dict (`encoder.*`, `quantizer.quantizers.{k}.*`),
conv filters permuted once to [C_in][K][C_out]; every launch goes through the C ABI (zb200_conv1d_f32, zb200_snake_f32,
zb200_rvq_encode); torch only owns the buffers.
"""
import ctypes
import math

import torch

from ._lib import check, lib, stream_ptr

_p = lambda t: ctypes.c_void_p(0 if t is None else t.data_ptr())


class DacEncoderB200:
    def __init__(self, sd: dict, device):
        self.device = torch.device(device)
        f = lambda k: sd[k].detach().to(self.device, torch.float32).contiguous()
        conv = lambda k: (f(k + ".weight").permute(1, 2, 0).contiguous(), f(k + ".bias"))  # [C_out, C_in, K] -> [C_in][K][C_out]
        alpha = lambda k: f(k).reshape(-1).contiguous()
        self.conv1 = conv("encoder.conv1")
        self.blocks = []
        j = 0
        while f"encoder.block.{j}.conv1.weight" in sd:
            p = f"encoder.block.{j}."
            units = [dict(a1=alpha(p + f"res_unit{r}.snake1.alpha"), c1=conv(p + f"res_unit{r}.conv1"),
                          a2=alpha(p + f"res_unit{r}.snake2.alpha"), c2=conv(p + f"res_unit{r}.conv2"), dil=d)
                     for r, d in ((1, 1), (2, 3), (3, 9))]
            k = sd[p + "conv1.weight"].shape[2]
            self.blocks.append(dict(units=units, a=alpha(p + "snake1.alpha"), down=conv(p + "conv1"), stride=k // 2))
            j += 1
        self.snake_f = alpha("encoder.snake1.alpha")
        self.conv2 = conv("encoder.conv2")
        self.hop = math.prod(b["stride"] for b in self.blocks)
        # quantiser, stacked over the codebooks
        Q = 0
        while f"quantizer.quantizers.{Q}.codebook.weight" in sd:
            Q += 1
        qk = lambda name: torch.stack([f(f"quantizer.quantizers.{k}.{name}") for k in range(Q)])
        self.n_q = Q
        self.in_w = qk("in_proj.weight").squeeze(-1).contiguous()    # [Q][D][H]
        self.in_b = qk("in_proj.bias").contiguous()                  # [Q][D]
        self.codebook = qk("codebook.weight").contiguous()           # [Q][N][D]
        self.out_w = qk("out_proj.weight").squeeze(-1).contiguous()  # [Q][H][D]
        self.out_b = qk("out_proj.bias").contiguous()                # [Q][H]
        # F.normalize(codebook) and its squared norm, with the reference's own torch expressions (modeling_dac.py:158-163)
        cbn = torch.nn.functional.normalize(self.codebook, dim=-1)
        self.cb_norm = torch.cat([cbn, cbn.pow(2).sum(-1, keepdim=True)], dim=-1).contiguous()  # [Q][N][D+1]

    # ---- launches
    def _conv(self, x, wb, stride=1, dil=1, pad=0, resid=None):
        w, b = wb
        C_in, K, C_out = w.shape
        B, _, T = x.shape
        T_out = (T + 2 * pad - dil * (K - 1) - 1) // stride + 1
        out = torch.empty(B, C_out, T_out, dtype=torch.float32, device=self.device)
        check(lib().zb200_conv1d_f32(_p(x), _p(w), _p(b), _p(resid), _p(out), B, C_in, C_out, T, T_out, K, stride, dil, pad,
                                     stream_ptr()), "conv1d_f32")
        return out

    def _snake(self, x, alpha):
        out = torch.empty_like(x)
        check(lib().zb200_snake_f32(_p(x), _p(alpha), _p(out), x.shape[0], x.shape[1], x.shape[2], stream_ptr()), "snake_f32")
        return out

    def latent(self, wav: torch.Tensor) -> torch.Tensor:
        """DacEncoder.forward (modeling_dac.py:462-472): fp32 [B, 1, T] -> [B, hidden, T / hop]."""
        x = self._conv(wav, self.conv1, pad=3)
        for blk in self.blocks:
            for u in blk["units"]:  # DacResidualUnit (:187-207): x + conv2(snake2(conv1(snake1(x))))
                y = self._conv(self._snake(x, u["a1"]), u["c1"], dil=u["dil"], pad=3 * u["dil"])
                x = self._conv(self._snake(y, u["a2"]), u["c2"], resid=x)
            s = blk["stride"]
            x = self._conv(self._snake(x, blk["a"]), blk["down"], stride=s, pad=math.ceil(s / 2))
        return self._conv(self._snake(x, self.snake_f), self.conv2, pad=1)

    def encode(self, wav: torch.Tensor) -> torch.Tensor:
        """fp32 [B, 1, T] -> codes int64 [B, n_q, T / hop] (DacModel.encode(...).audio_codes)."""
        if not wav.is_cuda:
            raise ValueError("zonos_b200 kernels need CUDA tensors (there is no CPU fallback)")
        assert wav.ndim == 3 and wav.shape[1] == 1, "expected mono audio [B, 1, T]"
        z = self.latent(wav.to(self.device, torch.float32).contiguous())
        B, H, T = z.shape
        codes = torch.empty(B, self.n_q, T, dtype=torch.int64, device=self.device)
        D, N = self.codebook.shape[2], self.codebook.shape[1]
        check(lib().zb200_rvq_encode(_p(z), _p(self.in_w), _p(self.in_b), _p(self.codebook), _p(self.cb_norm), _p(self.out_w),
                                     _p(self.out_b), _p(codes), B, self.n_q, H, D, N, T, stream_ptr()), "rvq_encode")
        return codes
