"""Model shapes and seeded synthetic weights (test infrastructure; see oracle/__init__.py).

Shapes: SURVEY.md Appendix A (Zonos-v0.1-transformer) and a tiny variant with the same head_dim (128).
State-dict key names follow the reference module tree (zonos/model.py:36-37, zonos/backbone/_torch.py:61-62,
88-91,114-115,147-148) so the same dict loads into the reference and into zonos_b200.
"""
import math

import torch

EOS_ID = 1024
MASK_ID = 1025
N_CODEBOOKS = 9
V_HEAD = 1025  # rows of each head before the load hook pads by 1025 % 8 == 1 (zonos/utils.py:22-37)
V_PAD = 1026

SHAPES = {
    "tiny": dict(d_model=512, n_layer=2, n_heads=4, n_kv_heads=2, d_ff=1024),
    "small": dict(d_model=1024, n_layer=3, n_heads=8, n_kv_heads=2, d_ff=2048),
    "real": dict(d_model=2048, n_layer=26, n_heads=16, n_kv_heads=4, d_ff=8192),
}


def reference_config_dict(shape: str) -> dict:
    """config.json-style dict accepted by the reference's ZonosConfig.from_dict (zonos/config.py:56-62)."""
    s = SHAPES[shape]
    return dict(
        backbone=dict(
            d_model=s["d_model"],
            n_layer=s["n_layer"],
            attn_mlp_d_intermediate=s["d_ff"],
            attn_layer_idx=list(range(s["n_layer"])),
            ssm_cfg={},
            attn_cfg=dict(num_heads=s["n_heads"], num_heads_kv=s["n_kv_heads"]),
            norm_epsilon=1e-5,
        ),
        prefix_conditioner=dict(conditioners=[], projection="none"),
        eos_token_id=EOS_ID,
        masked_token_id=MASK_ID,
        pad_vocab_to_multiple_of=8,
    )


def _uniform(gen, shape, bound):
    return (torch.rand(shape, generator=gen, dtype=torch.float32) * 2 - 1) * bound


def synth_transformer_weights(shape: str, seed: int = 0, head_gain: float = 4.0, padded_heads: bool = True,
                              zero_eos_row: bool = True) -> dict:
    """bf16 weights. Linear ~ U(+-1/sqrt(fan_in)) (nn.Linear's default bound), embeddings ~ N(0,1),
    LayerNorm affine perturbed away from (1,0) so the affine path is exercised.  heads scaled by head_gain
    for a realistic entropy; codebook-0 EOS row zeroed so random weights do not stop early."""
    s = SHAPES[shape]
    d, L, H, Hkv, F = s["d_model"], s["n_layer"], s["n_heads"], s["n_kv_heads"], s["d_ff"]
    D = d // H
    g = torch.Generator().manual_seed(seed)
    W = {}
    for i in range(L):
        p = f"backbone.layers.{i}."
        W[p + "norm.weight"] = 1 + 0.1 * torch.randn(d, generator=g)
        W[p + "norm.bias"] = 0.05 * torch.randn(d, generator=g)
        W[p + "mixer.in_proj.weight"] = _uniform(g, (H * D + 2 * Hkv * D, d), 1 / math.sqrt(d))
        W[p + "mixer.out_proj.weight"] = _uniform(g, (d, H * D), 1 / math.sqrt(H * D))
        W[p + "norm2.weight"] = 1 + 0.1 * torch.randn(d, generator=g)
        W[p + "norm2.bias"] = 0.05 * torch.randn(d, generator=g)
        W[p + "mlp.fc1.weight"] = _uniform(g, (2 * F, d), 1 / math.sqrt(d))
        W[p + "mlp.fc2.weight"] = _uniform(g, (d, F), 1 / math.sqrt(F))
    W["backbone.norm_f.weight"] = 1 + 0.1 * torch.randn(d, generator=g)
    W["backbone.norm_f.bias"] = 0.05 * torch.randn(d, generator=g)
    for k in range(N_CODEBOOKS):
        W[f"embeddings.{k}.weight"] = torch.randn(V_PAD, d, generator=g)
    for k in range(N_CODEBOOKS):
        h = _uniform(g, (V_HEAD, d), 1 / math.sqrt(d)) * head_gain
        if k == 0 and zero_eos_row:
            h[EOS_ID] = 0
        if padded_heads:
            h = torch.cat([h, torch.zeros(V_PAD - V_HEAD, d)], 0)
        W[f"heads.{k}.weight"] = h
    return {k: v.to(torch.bfloat16) for k, v in W.items()}


def synth_conditioning(batch: int, l_cond: int, d_model: int, seed: int = 1) -> torch.Tensor:
    """prefix_conditioning [2B, L_cond, d] bf16: LayerNorm'd noise."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(2 * batch, l_cond, d_model, generator=g)
    return torch.nn.functional.layer_norm(x, (d_model,)).to(torch.bfloat16)


# ----------------------------------------------------------------------------- DAC 44.1 kHz decoder
DAC = dict(n_codebooks=9, codebook_size=1024, codebook_dim=8, latent=1024, dec_ch=1536, strides=(8, 8, 4, 2))
DAC_TINY = dict(n_codebooks=9, codebook_size=1024, codebook_dim=8, latent=64, dec_ch=128, strides=(8, 8, 4, 2))


def synth_dac_weights(seed: int = 7, arch: dict = DAC) -> dict:
    """fp32 weights with HF DacModel key names, variance-preserving init:
    conv sigma^2 = 1/(C_in*k); convT sigma^2 = s/(C_in*k); residual 1x1 scaled by 0.5; alpha ~ U(0.5,2)."""
    g = torch.Generator().manual_seed(seed)
    W = {}
    nq, cs, cd, lat = arch["n_codebooks"], arch["codebook_size"], arch["codebook_dim"], arch["latent"]
    for k in range(nq):
        p = f"quantizer.quantizers.{k}."
        W[p + "codebook.weight"] = torch.randn(cs, cd, generator=g)
        W[p + "out_proj.weight"] = torch.randn(lat, cd, 1, generator=g) / math.sqrt(cd * nq)
        W[p + "out_proj.bias"] = 0.01 * torch.randn(lat, generator=g)
    C = arch["dec_ch"]
    W["decoder.conv1.weight"] = torch.randn(C, lat, 7, generator=g) / math.sqrt(lat * 7)
    W["decoder.conv1.bias"] = 0.01 * torch.randn(C, generator=g)
    for j, s in enumerate(arch["strides"]):
        p = f"decoder.block.{j}."
        Co = C // 2
        W[p + "snake1.alpha"] = 0.5 + 1.5 * torch.rand(1, C, 1, generator=g)
        W[p + "conv_t1.weight"] = torch.randn(C, Co, 2 * s, generator=g) * math.sqrt(s / (C * 2 * s))
        W[p + "conv_t1.bias"] = 0.01 * torch.randn(Co, generator=g)
        for r in (1, 2, 3):
            q = p + f"res_unit{r}."
            W[q + "snake1.alpha"] = 0.5 + 1.5 * torch.rand(1, Co, 1, generator=g)
            W[q + "conv1.weight"] = torch.randn(Co, Co, 7, generator=g) / math.sqrt(Co * 7)
            W[q + "conv1.bias"] = 0.01 * torch.randn(Co, generator=g)
            W[q + "snake2.alpha"] = 0.5 + 1.5 * torch.rand(1, Co, 1, generator=g)
            W[q + "conv2.weight"] = 0.5 * torch.randn(Co, Co, 1, generator=g) / math.sqrt(Co)
            W[q + "conv2.bias"] = 0.01 * torch.randn(Co, generator=g)
        C = Co
    W["decoder.snake1.alpha"] = 0.5 + 1.5 * torch.rand(1, C, 1, generator=g)
    W["decoder.conv2.weight"] = 0.15 * torch.randn(1, C, 7, generator=g) / math.sqrt(C * 7)  # keep tanh unsaturated
    W["decoder.conv2.bias"] = 0.01 * torch.randn(1, generator=g)
    return W


DAC_ENC = dict(enc_ch=64, strides=(2, 4, 8, 8), latent=1024, n_codebooks=9, codebook_size=1024, codebook_dim=8)
DAC_ENC_TINY = dict(enc_ch=8, strides=(2, 4, 8, 8), latent=64, n_codebooks=9, codebook_size=1024, codebook_dim=8)


def synth_dac_encoder_weights(seed: int = 8, arch: dict = DAC_ENC) -> dict:
    """fp32 encoder + quantizer-input weights with the HF DacModel key names; variance-preserving like synth_dac_weights
    (conv sigma^2 = 1/(C_in*k), residual 1x1 at half gain, alpha ~ U(0.5,2)); in_proj rows ~ N(0, 1/H) so the projected
    latents are O(1) before F.normalize."""
    g = torch.Generator().manual_seed(seed)
    W = {}
    C = arch["enc_ch"]
    W["encoder.conv1.weight"] = torch.randn(C, 1, 7, generator=g) / math.sqrt(7)
    W["encoder.conv1.bias"] = 0.01 * torch.randn(C, generator=g)
    for j, s in enumerate(arch["strides"]):
        p = f"encoder.block.{j}."
        for r in (1, 2, 3):
            q = p + f"res_unit{r}."
            W[q + "snake1.alpha"] = 0.5 + 1.5 * torch.rand(1, C, 1, generator=g)
            W[q + "conv1.weight"] = torch.randn(C, C, 7, generator=g) / math.sqrt(C * 7)
            W[q + "conv1.bias"] = 0.01 * torch.randn(C, generator=g)
            W[q + "snake2.alpha"] = 0.5 + 1.5 * torch.rand(1, C, 1, generator=g)
            W[q + "conv2.weight"] = 0.5 * torch.randn(C, C, 1, generator=g) / math.sqrt(C)
            W[q + "conv2.bias"] = 0.01 * torch.randn(C, generator=g)
        W[p + "snake1.alpha"] = 0.5 + 1.5 * torch.rand(1, C, 1, generator=g)
        W[p + "conv1.weight"] = torch.randn(2 * C, C, 2 * s, generator=g) / math.sqrt(C * 2 * s)
        W[p + "conv1.bias"] = 0.01 * torch.randn(2 * C, generator=g)
        C *= 2
    H = arch["latent"]
    W["encoder.snake1.alpha"] = 0.5 + 1.5 * torch.rand(1, C, 1, generator=g)
    W["encoder.conv2.weight"] = torch.randn(H, C, 3, generator=g) / math.sqrt(C * 3)
    W["encoder.conv2.bias"] = 0.01 * torch.randn(H, generator=g)
    for k in range(arch["n_codebooks"]):
        p = f"quantizer.quantizers.{k}."
        W[p + "in_proj.weight"] = torch.randn(arch["codebook_dim"], H, 1, generator=g) / math.sqrt(H)
        W[p + "in_proj.bias"] = 0.01 * torch.randn(arch["codebook_dim"], generator=g)
    return W
