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
