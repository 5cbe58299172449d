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

    def _pack(self):
        """Kernel-layout copy of this module's own weights (built on first use, dropped when they are reloaded)."""
        from .engine import PackedWeights

        if getattr(self, "_packed", None) is None:
            dev = self.norm_f.weight.device
            if dev.type != "cuda":
                raise RuntimeError("zonos_b200 runs on CUDA (B200) only; move the backbone with .to('cuda')")
            self._packed = PackedWeights(self.state_dict(), self.config.n_layer, 0, dev, prefix="")
        return self._packed

    def _load_from_state_dict(self, *args, **kwargs):
        self._packed = None
        return super()._load_from_state_dict(*args, **kwargs)

    def forward(self, hidden_states: torch.Tensor, inference_params) -> torch.Tensor:
        """The plugin contract of zonos/backbone/_torch.py:73-80: hidden_states bf16 [B2, S, d] -> norm_f(blocks(...)),
        same shape.  Positions are seqlen_offset .. seqlen_offset + S - 1 (`lengths_per_sample` is uniform over the batch
        in every caller, zonos/model.py:319,421); the KV rows are written into inference_params.key_value_memory_dict
        (the paged pool of allocate_inference_cache); the caller advances seqlen_offset, as in the reference.  Runs the
        session engine's kernels (zb200_engine_forward): S > 1 is the causal prefill, S == 1 the decode step.  With this,
        the reference's own _prefill / _compute_logits / _decode_one_token run on BACKBONES["b200"] unchanged; the fast
        path (Zonos.generate of this package) does not go through here -- it keeps the loop on the device."""
        from .engine import Engine

        if not hidden_states.is_cuda or hidden_states.dtype != torch.bfloat16:
            raise ValueError("B200ZonosBackbone.forward needs CUDA bf16 hidden states (there is no CPU fallback)")
        B2, S, d = hidden_states.shape
        kv = inference_params.key_value_memory_dict["paged"]
        pos0 = int(inference_params.seqlen_offset)
        if B2 % 2 or kv.n_seq != B2:
            raise ValueError(f"hidden_states has {B2} rows; the cache was allocated for {kv.n_seq} (2 x batch) sequences")
        assert pos0 + S <= kv.max_seq, "KV cache overflow"  # _torch.py:44-45
        key = (id(kv), B2, S)
        eng = getattr(self, "_fwd_engine", None)
        if eng is None or self._fwd_key != key:
            if eng is not None:
                eng.close()
            eng = Engine(self._pack(), self.config, B2 // 2, S, kv.max_seq, 0, None, None, 0.0, kv=kv)
            self._fwd_engine, self._fwd_key = eng, key
        eng.x[: B2 * S].view(B2, S, d).copy_(hidden_states)
        out = torch.empty_like(hidden_states)
        eng.forward(S, pos0, out)
        return out


BACKBONES = {"b200": B200ZonosBackbone}
