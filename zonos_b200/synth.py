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
