"""Random-init weights of a Zonos architecture, generated on the device (bench / smoke; there is no network for
checkpoints).  Same distribution family as the test weights of oracle/shapes.py (nn.Linear-style uniform bounds,
N(0,1) embeddings, perturbed LayerNorm affine, heads x4, codebook-0 EOS row zeroed so random weights do not stop
early: SURVEY 8(c)), but drawn with the CUDA generator, so NOT element-identical to the CPU test weights."""
import math

import torch

REAL = dict(d_model=2048, n_layer=26, n_heads=16, n_kv_heads=4, d_ff=8192)
TINY = dict(d_model=512, n_layer=2, n_heads=4, n_kv_heads=2, d_ff=1024)


def config_dict(shape: dict) -> dict:
    return dict(
        backbone=dict(d_model=shape["d_model"], n_layer=shape["n_layer"], attn_mlp_d_intermediate=shape["d_ff"],
                      attn_layer_idx=list(range(shape["n_layer"])), ssm_cfg={},
                      attn_cfg=dict(num_heads=shape["n_heads"], num_heads_kv=shape["n_kv_heads"]), norm_epsilon=1e-5),
        prefix_conditioner=dict(conditioners=[], projection="none"),
        eos_token_id=1024, masked_token_id=1025, pad_vocab_to_multiple_of=8)


def transformer_state_dict(shape: dict, device="cuda", seed: int = 0, head_gain: float = 4.0) -> dict:
    g = torch.Generator(device=device).manual_seed(seed)
    d, L, H, Hkv, F = shape["d_model"], shape["n_layer"], shape["n_heads"], shape["n_kv_heads"], shape["d_ff"]
    D = d // H
    bf = torch.bfloat16

    def uni(shape_, bound):
        return ((torch.rand(shape_, generator=g, device=device) * 2 - 1) * bound).to(bf)

    def nrm(shape_, mean, std):
        return (mean + std * torch.randn(shape_, generator=g, device=device)).to(bf)

    W = {}
    for i in range(L):
        p = f"backbone.layers.{i}."
        W[p + "norm.weight"], W[p + "norm.bias"] = nrm((d,), 1, 0.1), nrm((d,), 0, 0.05)
        W[p + "mixer.in_proj.weight"] = uni((H * D + 2 * Hkv * D, d), 1 / math.sqrt(d))
        W[p + "mixer.out_proj.weight"] = uni((d, H * D), 1 / math.sqrt(H * D))
        W[p + "norm2.weight"], W[p + "norm2.bias"] = nrm((d,), 1, 0.1), nrm((d,), 0, 0.05)
        W[p + "mlp.fc1.weight"] = uni((2 * F, d), 1 / math.sqrt(d))
        W[p + "mlp.fc2.weight"] = uni((d, F), 1 / math.sqrt(F))
    W["backbone.norm_f.weight"], W["backbone.norm_f.bias"] = nrm((d,), 1, 0.1), nrm((d,), 0, 0.05)
    for k in range(9):
        W[f"embeddings.{k}.weight"] = nrm((1026, d), 0, 1)
    for k in range(9):
        h = uni((1025, d), head_gain / math.sqrt(d))
        if k == 0:
            h[1024] = 0
        W[f"heads.{k}.weight"] = h
    return W


def conditioning(batch: int, l_cond: int, d_model: int, device="cuda", seed: int = 1) -> torch.Tensor:
    g = torch.Generator(device=device).manual_seed(seed)
    x = torch.randn(2 * batch, l_cond, d_model, generator=g, device=device)
    return torch.nn.functional.layer_norm(x, (d_model,)).to(torch.bfloat16)


HYBRID = dict(d_model=2048, n_layer=46, attn_every=9, n_heads=16, n_kv_heads=4, d_ff=8192, expand=2, head_dim=64, d_state=128,
              n_groups=1, d_conv=4)  # recalled Zonos-v0.1-hybrid shape (the checkpoint's config.json is not available offline)


def hybrid_model(shape: dict = HYBRID, device="cuda", seed: int = 0, head_gain: float = 4.0):
    """Random-init hybrid model (Mamba2 blocks, attention + GatedMLP at every `attn_every`-th block) drawn on the device:
    returns zonos_b200.hybrid.ZonosHybridB200.  Mamba2 parameters in the ranges of its own initialisation (dt in
    [1e-3, 0.1], A in [1, 16], D = 1)."""
    from .hybrid import HybridBackboneB200, ZonosHybridB200

    g = torch.Generator(device=device).manual_seed(seed)
    bf = torch.bfloat16
    d, H, Hkv, F = shape["d_model"], shape["n_heads"], shape["n_kv_heads"], shape["d_ff"]
    Di = shape["expand"] * d
    Hm, N, G = Di // shape["head_dim"], shape["d_state"], shape["n_groups"]
    uni = lambda sh, bound: ((torch.rand(sh, generator=g, device=device) * 2 - 1) * bound).to(bf)
    nrm = lambda sh, mean, std: (mean + std * torch.randn(sh, generator=g, device=device)).to(bf)
    layers = []
    for l in range(shape["n_layer"]):
        if l % shape["attn_every"] == shape["attn_every"] - 1:
            p = {"kind": "attn", "in_proj.weight": uni(((H + 2 * Hkv) * 128, d), 1 / math.sqrt(d)),
                 "out_proj.weight": uni((d, H * 128), 1 / math.sqrt(H * 128)), "norm2.weight": nrm((d,), 1, 0.1),
                 "fc1.weight": uni((2 * F, d), 1 / math.sqrt(d)), "fc2.weight": uni((d, F), 1 / math.sqrt(F))}
        else:
            dt = torch.exp(torch.rand(Hm, generator=g, device=device) * (math.log(0.1) - math.log(1e-3)) + math.log(1e-3))
            p = {"kind": "mamba2", "in_proj.weight": uni((2 * Di + 2 * G * N + Hm, d), 1 / math.sqrt(d)),
                 "conv1d.weight": uni((Di + 2 * G * N, shape["d_conv"]), 1 / math.sqrt(shape["d_conv"])),
                 "conv1d.bias": nrm((Di + 2 * G * N,), 0, 0.1), "dt_bias": (dt + torch.log(-torch.expm1(-dt))).float(),
                 "A_log": torch.log(1 + 15 * torch.rand(Hm, generator=g, device=device)).float(),
                 "D": torch.ones(Hm, device=device), "mixer_norm.weight": nrm((Di,), 1, 0.1),
                 "out_proj.weight": uni((d, Di), 1 / math.sqrt(Di))}
        p["norm.weight"] = nrm((d,), 1, 0.1)
        layers.append(p)
    sd = {}
    for k in range(9):
        sd[f"embeddings.{k}.weight"] = nrm((1026, d), 0, 1)
        h = uni((1026, d), head_gain / math.sqrt(d))
        h[1024:] = 0  # no early EOS with random weights; slot 1025 is the pad row
        sd[f"heads.{k}.weight"] = h
    mshape = dict(d_model=d, d_inner=Di, n_heads=Hm, head_dim=shape["head_dim"], d_state=N, n_groups=G)
    net = HybridBackboneB200(layers, dict(weight=nrm((d,), 1, 0.1)), mshape, device, H, Hkv)
    return ZonosHybridB200(sd, net)


# ----------------------------------------------------------------------------------------------- DAC 44.1 kHz decoder
DAC = dict(n_codebooks=9, codebook_size=1024, codebook_dim=8, latent=1024, dec_ch=1536, strides=(8, 8, 4, 2))
DAC_TINY = dict(n_codebooks=9, codebook_size=1024, codebook_dim=8, latent=64, dec_ch=128, strides=(8, 8, 4, 2))


def dac_state_dict(arch: dict = DAC, device="cuda", seed: int = 7) -> dict:
    """Random decoder + quantizer tensors with the HF DacModel key names, drawn on the device.  HF's own init (std 0.02,
    alpha = 1) makes the waveform vanish, so the init is variance-preserving: conv sigma^2 = 1 / (C_in k),
    transposed conv sigma^2 = stride / (C_in k), the residual branch's 1x1 conv at half gain, Snake alpha ~ U(0.5, 2), last
    conv small enough to keep tanh unsaturated."""
    g = torch.Generator(device=device).manual_seed(seed)
    rn = lambda *sh: torch.randn(*sh, generator=g, device=device)
    alpha = lambda c: 0.5 + 1.5 * torch.rand(1, c, 1, generator=g, device=device)
    nq, cs, cd, lat = arch["n_codebooks"], arch["codebook_size"], arch["codebook_dim"], arch["latent"]
    W = {}
    for k in range(nq):
        p = f"quantizer.quantizers.{k}."
        W[p + "codebook.weight"] = rn(cs, cd)
        W[p + "out_proj.weight"] = rn(lat, cd, 1) / math.sqrt(cd * nq)
        W[p + "out_proj.bias"] = 0.01 * rn(lat)
    C = arch["dec_ch"]
    W["decoder.conv1.weight"] = rn(C, lat, 7) / math.sqrt(lat * 7)
    W["decoder.conv1.bias"] = 0.01 * rn(C)
    for j, s in enumerate(arch["strides"]):
        p, Co = f"decoder.block.{j}.", C // 2
        W[p + "snake1.alpha"] = alpha(C)
        W[p + "conv_t1.weight"] = rn(C, Co, 2 * s) * math.sqrt(s / (C * 2 * s))
        W[p + "conv_t1.bias"] = 0.01 * rn(Co)
        for r in (1, 2, 3):
            q = p + f"res_unit{r}."
            W[q + "snake1.alpha"] = alpha(Co)
            W[q + "conv1.weight"] = rn(Co, Co, 7) / math.sqrt(Co * 7)
            W[q + "conv1.bias"] = 0.01 * rn(Co)
            W[q + "snake2.alpha"] = alpha(Co)
            W[q + "conv2.weight"] = 0.5 * rn(Co, Co, 1) / math.sqrt(Co)
            W[q + "conv2.bias"] = 0.01 * rn(Co)
        C = Co
    W["decoder.snake1.alpha"] = alpha(C)
    W["decoder.conv2.weight"] = 0.15 * rn(1, C, 7) / math.sqrt(C * 7)
    W["decoder.conv2.bias"] = 0.01 * rn(1)
    return W


# ----------------------------------------------------------------------------------------------- Mamba2 mixer (hybrid)
def mamba2_mixer_params(d_model: int = 2048, expand: int = 2, head_dim: int = 64, d_state: int = 128, n_groups: int = 1,
                        d_conv: int = 4, device="cuda", seed: int = 0, dtype=torch.bfloat16) -> dict:
    """One Mamba2 mixer's parameters in the ranges of its own initialisation (dt in [1e-3, 0.1], A in [1, 16], D ~ 1),
    drawn on the device; keys as mamba_ssm's Mamba2 module plus a `shape` dict."""
    g = torch.Generator(device=device).manual_seed(seed)
    rn = lambda *sh: torch.randn(*sh, generator=g, device=device)
    ru = lambda *sh: torch.rand(*sh, generator=g, device=device)
    Di = expand * d_model
    H, cd = Di // head_dim, Di + 2 * n_groups * d_state
    dt = torch.exp(ru(H) * (math.log(0.1) - math.log(1e-3)) + math.log(1e-3)).clamp(min=1e-4)
    return {
        "in_proj.weight": (rn(2 * Di + 2 * n_groups * d_state + H, d_model) / d_model ** 0.5).to(dtype),
        "conv1d.weight": (rn(cd, d_conv) / d_conv ** 0.5).to(dtype),
        "conv1d.bias": (0.1 * rn(cd)).to(dtype),
        "dt_bias": (dt + torch.log(-torch.expm1(-dt))).float(),
        "A_log": torch.log(1 + 15 * ru(H)).float(),
        "D": torch.ones(H, device=device) + 0.1 * rn(H),
        "norm.weight": (1 + 0.1 * rn(Di)).to(dtype),
        "out_proj.weight": (rn(d_model, Di) / Di ** 0.5).to(dtype),
        "shape": dict(d_inner=Di, n_heads=H, head_dim=head_dim, d_state=d_state, n_groups=n_groups, eps=1e-5, d_model=d_model,
                      d_conv=d_conv),
    }


RESNET293_BLOCKS = (10, 20, 64, 3)


def speaker_state_dicts(in_planes: int = 64, num_blocks=RESNET293_BLOCKS, device="cuda", seed: int = 0, embd_dim: int = 256,
                        lda_dim: int = 128):
    """Random-init state dicts of the speaker encoder (`ResNet293_based` + the LDA Linear, zonos/speaker_cloning.py:198-223,
    404-405) under the reference's key names, drawn on the device: He-scaled convolutions, perturbed BatchNorm affine /
    running statistics, a small second BatchNorm weight per block so the residual stream stays O(1) over 97 blocks."""
    g = torch.Generator(device=device).manual_seed(seed)
    sd = {}
    rn = lambda *sh: torch.randn(*sh, generator=g, device=device)

    def bn(name, c, w_mean=1.0):
        sd[name + ".weight"] = w_mean * (1 + 0.1 * rn(c))
        sd[name + ".bias"] = 0.1 * rn(c)
        sd[name + ".running_mean"] = 0.1 * rn(c)
        sd[name + ".running_var"] = 1 + 0.2 * torch.rand(c, generator=g, device=device)

    sd["front.conv1.weight"] = rn(in_planes, 1, 3, 3) * math.sqrt(2.0 / 9)
    bn("front.bn1", in_planes)
    cin = in_planes
    for li, (mult, n) in enumerate(zip((1, 2, 4, 8), num_blocks)):
        planes = in_planes * mult
        for bi in range(n):
            stride = (1 if li == 0 else 2) if bi == 0 else 1
            p = f"front.layer{li + 1}.{bi}"
            sd[p + ".conv1.weight"] = rn(planes, cin, 3, 3) * math.sqrt(2.0 / (cin * 9))
            bn(p + ".bn1", planes)
            sd[p + ".conv2.weight"] = rn(planes, planes, 3, 3) * math.sqrt(2.0 / (planes * 9))
            bn(p + ".bn2", planes, w_mean=0.3)
            if stride != 1 or cin != planes:
                sd[p + ".downsample.0.weight"] = rn(planes, cin, 1, 1) * math.sqrt(2.0 / cin)
                bn(p + ".downsample.1", planes)
            cin = planes
    feat = in_planes * 8 * 10
    sd["pooling.attention.0.weight"] = rn(128, feat, 1) / math.sqrt(feat)
    sd["pooling.attention.0.bias"] = 0.1 * rn(128)
    bn("pooling.attention.2", 128)
    sd["pooling.attention.3.weight"] = rn(feat, 128, 1) / math.sqrt(128)
    sd["pooling.attention.3.bias"] = 0.1 * rn(feat)
    sd["bottleneck.weight"] = rn(embd_dim, 2 * feat) / math.sqrt(2 * feat)
    sd["bottleneck.bias"] = 0.1 * rn(embd_dim)
    lda = {"weight": rn(lda_dim, embd_dim) / math.sqrt(embd_dim), "bias": 0.1 * rn(lda_dim)}
    return sd, lda


def speaker_clip(seconds: float = 10.0, sample_rate: int = 44100, channels: int = 2, seed: int = 5) -> torch.Tensor:
    """A synthetic clip on the HOST (a few partials + noise), fp32 [channels, n] in [-1, 1]: what torchaudio.load returns."""
    g = torch.Generator().manual_seed(seed)
    n = int(seconds * sample_rate)
    t = torch.arange(n, dtype=torch.float64) / sample_rate
    wav = torch.zeros(channels, n, dtype=torch.float64)
    for c in range(channels):
        for f0 in (110.0 * (c + 1), 440.0, 1730.0, 5200.0):
            wav[c] += float(torch.rand(1, generator=g)) * torch.sin(2 * torch.pi * f0 * t + float(torch.rand(1, generator=g)))
        wav[c] += 0.1 * torch.randn(n, generator=g, dtype=torch.float64)
    return (0.25 * wav).float()
