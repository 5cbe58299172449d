"""BACKBONES registry (the reference's plugin seam, zonos/backbone/__init__.py:1-12) with the B200 backbone.

B200ZonosBackbone keeps the parameter names of TorchZonosBackbone (zonos/backbone/_torch.py:52-152) so real
checkpoints load unchanged; its compute runs in libzonos_b200.so (tcgen05 GEMMs with fused LayerNorm-consumer /
RoPE / KV-page-write / SwiGLU / residual epilogues, paged GQA attention).
"""
import torch
import torch.nn as nn

from .config import BackboneConfig


class _Mixer(nn.Module):
    def __init__(self, d, n_heads, n_kv, head_dim):
        super().__init__()
        self.in_proj = nn.Linear(d, (n_heads + 2 * n_kv) * head_dim, bias=False)
        self.out_proj = nn.Linear(n_heads * head_dim, d, bias=False)


class _Mlp(nn.Module):
    def __init__(self, d, d_ff):
        super().__init__()
        self.fc1 = nn.Linear(d, 2 * d_ff, bias=False)
        self.fc2 = nn.Linear(d_ff, d, bias=False)


class _Block(nn.Module):
    def __init__(self, cfg: BackboneConfig):
        super().__init__()
        d, H, Hkv = cfg.d_model, cfg.attn_cfg["num_heads"], cfg.attn_cfg["num_heads_kv"]
        self.norm = nn.LayerNorm(d, eps=cfg.norm_epsilon)
        self.mixer = _Mixer(d, H, Hkv, d // H)
        self.norm2 = nn.LayerNorm(d, eps=cfg.norm_epsilon)
        self.mlp = _Mlp(d, cfg.attn_mlp_d_intermediate)


class B200ZonosBackbone(nn.Module):
    supported_architectures = ["transformer"]

    def __init__(self, config: BackboneConfig):
        assert not config.ssm_cfg, "This backbone implementation only supports the Transformer model."
        super().__init__()
        self.config = config
        self.layers = nn.ModuleList(_Block(config) for _ in range(config.n_layer))
        self.norm_f = nn.LayerNorm(config.d_model, eps=config.norm_epsilon)

    @property
    def n_heads(self):
        return self.config.attn_cfg["num_heads"]

    @property
    def n_kv_heads(self):
        return self.config.attn_cfg["num_heads_kv"]

    def allocate_inference_cache(self, batch_size: int, max_seqlen: int, dtype: torch.dtype = torch.bfloat16):
        """Paged pool + block table instead of one dense [B2, max_seq, 2, Hkv, D] tensor per layer (_torch.py:64-71)."""
        from .ops import PagedKV

        assert dtype == torch.bfloat16, "the B200 KV cache is bf16"
        dev = self.norm_f.weight.device
        return {"paged": PagedKV(self.config.n_layer, batch_size, max_seqlen, self.n_kv_heads, dev)}

    def forward(self, hidden_states, inference_params):
        raise NotImplementedError(
            "B200ZonosBackbone runs inside zonos_b200.Zonos (prefill / decode engine); a standalone forward is not exposed")


BACKBONES = {"b200": B200ZonosBackbone}
