"""DAC 44.1 kHz decoder on CPU, fp32 (test infrastructure; see oracle/__init__.py).

Restates zonos/autoencoder.py:44-47 (DACAutoencoder.decode; fp32 on CPU) and the third-party arithmetic it
calls -- transformers (pinned 4.48.3 in the reference's uv.lock:2063-2065; 5.5.0 installed here)
models/dac/modeling_dac.py: DacResidualVectorQuantizer.from_codes (:345-369), DacDecoder.forward (:429-439),
DacDecoderBlock.forward (:255-262), DacResidualUnit.forward (:187-207), Snake1d.forward (:94-99),
DacModel.decode (:608-638).  Weights: plain dict with the HF key names.
"""
import math

import torch
import torch.nn.functional as F


def snake(x, alpha):
    """x + (alpha + 1e-9)^-1 * sin(alpha * x)^2, alpha [1,C,1]."""
    return x + (alpha + 1e-9).reciprocal() * torch.sin(alpha * x).pow(2)


def from_codes(W, codes):
    """codes int64 [B,Q,T] -> z fp32 [B,latent,T] = sum_k (out_proj_k(codebook_k[codes_k]) ), k ascending."""
    z = 0.0
    for k in range(codes.shape[1]):
        p = f"quantizer.quantizers.{k}."
        lat = F.embedding(codes[:, k, :], W[p + "codebook.weight"]).transpose(1, 2)
        z = z + F.conv1d(lat, W[p + "out_proj.weight"], W[p + "out_proj.bias"])
    return z


def _res_unit(W, p, x, dilation):
    y = F.conv1d(snake(x, W[p + "snake1.alpha"]), W[p + "conv1.weight"], W[p + "conv1.bias"],
                 dilation=dilation, padding=3 * dilation)
    y = F.conv1d(snake(y, W[p + "snake2.alpha"]), W[p + "conv2.weight"], W[p + "conv2.bias"])
    return x + y  # same length: the crop branch (modeling_dac.py:202-205) is never taken with pad = 3*dil


def decoder(W, z, strides=(8, 8, 4, 2), taps=None):
    """z [B,latent,T] -> [B,1,prod(strides)*T].  `taps` (dict) collects intermediate activations if given."""
    x = F.conv1d(z, W["decoder.conv1.weight"], W["decoder.conv1.bias"], padding=3)
    if taps is not None:
        taps["conv1"] = x
    for j, s in enumerate(strides):
        p = f"decoder.block.{j}."
        x = F.conv_transpose1d(snake(x, W[p + "snake1.alpha"]), W[p + "conv_t1.weight"], W[p + "conv_t1.bias"],
                               stride=s, padding=math.ceil(s / 2))
        if taps is not None:
            taps[f"block{j}.convt"] = x
        for r, dil in ((1, 1), (2, 3), (3, 9)):
            x = _res_unit(W, p + f"res_unit{r}.", x, dil)
        if taps is not None:
            taps[f"block{j}"] = x
    x = F.conv1d(snake(x, W["decoder.snake1.alpha"]), W["decoder.conv2.weight"], W["decoder.conv2.bias"], padding=3)
    return torch.tanh(x)


def decode(W, codes, strides=(8, 8, 4, 2), taps=None):
    """codes int64 [B,9,T] -> fp32 [B,1,512*T]."""
    assert codes.shape[1] == 9, f"Expected 9 codebooks, got {codes.shape[1]}"
    return decoder(W, from_codes(W, codes), strides, taps).float()


# ----------------------------------------------------------------------------------------------------------------
# fp16-storage model of the same decoder: what the reference computes on CUDA under torch.autocast(fp16)
# (zonos/autoencoder.py:46): conv weights and conv inputs are fp16, conv outputs are rounded to fp16, Snake runs in
# fp32 on the fp16 conv output and is rounded to fp16 when the next conv consumes it, residual adds are fp16 + fp16.
# Accumulation stays fp32 here (CPU), so this isolates the rounding POINTS; used to separate "fp16 storage error"
# (this model vs fp32) from kernel error (CUDA vs this model).
def _h(x):
    return x.half().float()


def decode_fp16_storage(W, codes, strides=(8, 8, 4, 2)):
    Wh = {k: (_h(v) if k.endswith("weight") and "codebook" not in k and "quantizer" not in k else v) for k, v in W.items()}
    x = _h(from_codes(W, codes))
    x = _h(F.conv1d(x, Wh["decoder.conv1.weight"], W["decoder.conv1.bias"], padding=3))
    for j, s in enumerate(strides):
        p = f"decoder.block.{j}."
        x = _h(F.conv_transpose1d(_h(snake(x, W[p + "snake1.alpha"])), Wh[p + "conv_t1.weight"], W[p + "conv_t1.bias"],
                                  stride=s, padding=math.ceil(s / 2)))
        for r, dil in ((1, 1), (2, 3), (3, 9)):
            q = p + f"res_unit{r}."
            y = _h(F.conv1d(_h(snake(x, W[q + "snake1.alpha"])), Wh[q + "conv1.weight"], W[q + "conv1.bias"],
                            dilation=dil, padding=3 * dil))
            y = _h(F.conv1d(_h(snake(y, W[q + "snake2.alpha"])), Wh[q + "conv2.weight"], W[q + "conv2.bias"]))
            x = _h(x + y)
    x = _h(F.conv1d(_h(snake(x, W["decoder.snake1.alpha"])), _h(W["decoder.conv2.weight"]), W["decoder.conv2.bias"], padding=3))
    return torch.tanh(x).float()


# ----------------------------------------------------------------------------------------------- encoder + RVQ (fp32)
# Restates zonos/autoencoder.py:27-28 (DACAutoencoder.encode) -> DacModel.encode (modeling_dac.py:581-605): DacEncoder.forward
# (:462-472), DacEncoderBlock.forward (:226-232), DacResidualUnit.forward (:187-207), DacResidualVectorQuantizer.forward
# (:300-343, eval mode, all codebooks) and DacVectorQuantize.forward / decode_latents (:128-170).
def encoder(W, wav, strides=(2, 4, 8, 8)):
    """wav fp32 [B,1,T] -> latent [B,hidden,T / prod(strides)]."""
    x = F.conv1d(wav, W["encoder.conv1.weight"], W["encoder.conv1.bias"], padding=3)
    for j, s in enumerate(strides):
        p = f"encoder.block.{j}."
        for r, dil in ((1, 1), (2, 3), (3, 9)):
            x = _res_unit(W, p + f"res_unit{r}.", x, dil)
        x = F.conv1d(snake(x, W[p + "snake1.alpha"]), W[p + "conv1.weight"], W[p + "conv1.bias"], stride=s, padding=math.ceil(s / 2))
    return F.conv1d(snake(x, W["encoder.snake1.alpha"]), W["encoder.conv2.weight"], W["encoder.conv2.bias"], padding=1)


def rvq_encode(W, latent, n_q=9, return_margins=False):
    """latent [B,H,T] -> codes int64 [B,n_q,T] (and, optionally, the gap between the best and the second-best distance of
    every pick, which is what decides whether a differing code is a tie)."""
    residual, codes, margins = latent, [], []
    B, H, T = latent.shape
    for k in range(n_q):
        p = f"quantizer.quantizers.{k}."
        proj = F.conv1d(residual, W[p + "in_proj.weight"], W[p + "in_proj.bias"])
        enc = F.normalize(proj.permute(0, 2, 1).reshape(B * T, -1))
        cb = F.normalize(W[p + "codebook.weight"])
        dist = -(enc.pow(2).sum(1, keepdim=True) - 2 * enc @ cb.t()) + cb.pow(2).sum(1, keepdim=True).t()
        top2 = dist.topk(2, dim=1)
        idx = dist.max(1)[1].reshape(B, T)
        q = F.embedding(idx, W[p + "codebook.weight"]).transpose(1, 2)
        residual = residual - F.conv1d(q, W[p + "out_proj.weight"], W[p + "out_proj.bias"])
        codes.append(idx)
        margins.append((top2.values[:, 0] - top2.values[:, 1]).reshape(B, T))
    codes = torch.stack(codes, dim=1)
    return (codes, torch.stack(margins, dim=1)) if return_margins else codes


def encode(W, wav, strides=(2, 4, 8, 8), return_margins=False):
    """wav fp32 [B,1,T] (T a multiple of the hop) -> codes int64 [B,9,T/hop]."""
    return rvq_encode(W, encoder(W, wav.float(), strides), return_margins=return_margins)
