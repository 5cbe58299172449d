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


# ---------------------------------------------------------------------------------------------------- hybrid
class _Mamba2Params(nn.Module):
    """Parameters of one mamba_ssm Mamba2 mixer, under mamba_ssm's names (in_proj, conv1d, dt_bias, A_log, D, norm,
    out_proj) so that a hybrid checkpoint's `backbone.layers.{i}.mixer.*` tensors load.  PARITY UNPINNED: mamba_ssm 2.2.4 is
    not installed here; names and shapes are restated from its published module."""

    def __init__(self, d, ssm_cfg):
        super().__init__()
        expand, hd = ssm_cfg.get("expand", 2), ssm_cfg.get("headdim", 64)
        N, G, K = ssm_cfg.get("d_state", 128), ssm_cfg.get("ngroups", 1), ssm_cfg.get("d_conv", 4)
        Di = expand * d
        H = Di // hd
        self.shape = dict(d_model=d, d_inner=Di, n_heads=H, head_dim=hd, d_state=N, n_groups=G, d_conv=K)
        self.in_proj = nn.Linear(d, 2 * Di + 2 * G * N + H, bias=False)
        self.conv1d = nn.Conv1d(Di + 2 * G * N, Di + 2 * G * N, K, groups=Di + 2 * G * N, padding=K - 1)
        self.dt_bias, self.A_log, self.D = nn.Parameter(torch.zeros(H)), nn.Parameter(torch.zeros(H)), nn.Parameter(torch.ones(H))
        self.norm = nn.Module()
        self.norm.weight = nn.Parameter(torch.ones(Di))
        self.out_proj = nn.Linear(Di, d, bias=False)


class _HybridBlock(nn.Module):
    def __init__(self, cfg: BackboneConfig, is_attn: bool):
        super().__init__()
        d = cfg.d_model
        Norm = (lambda: _RMSWeight(d)) if cfg.rms_norm else (lambda: nn.LayerNorm(d, eps=cfg.norm_epsilon))
        self.norm = Norm()
        if is_attn:
            H, Hkv = cfg.attn_cfg["num_heads"], cfg.attn_cfg["num_heads_kv"]
            self.mixer = _Mixer(d, H, Hkv, d // H)
            d_mlp = cfg.attn_mlp_d_intermediate
        else:
            self.mixer = _Mamba2Params(d, cfg.ssm_cfg)
            d_mlp = cfg.d_intermediate
        if d_mlp:
            self.norm2 = Norm()
            self.mlp = _Mlp(d, d_mlp)


class _RMSWeight(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(d))


class B200HybridZonosBackbone(nn.Module):
    """BACKBONES["b200_hybrid"]: the reference's MambaSSMZonosBackbone (zonos/backbone/_mamba_ssm.py:9-57) as a parameter
    container whose compute is zonos_b200.hybrid.HybridBackboneB200 (Mamba2 step / scan kernels + the transformer path's
    attention and GEMM kernels, host-sequenced).  Selected by from_local when `ssm_cfg` is non-empty, like the reference
    (zonos/model.py:70-77).  PARITY UNPINNED against mamba_ssm (absent here)."""

    supported_architectures = ["transformer", "hybrid"]

    def __init__(self, config: BackboneConfig):
        super().__init__()
        self.config = config
        attn = set(config.attn_layer_idx)
        self.layers = nn.ModuleList(_HybridBlock(config, i in attn) for i in range(config.n_layer))
        self.norm_f = _RMSWeight(config.d_model) if config.rms_norm else nn.LayerNorm(config.d_model, eps=config.norm_epsilon)
        self._net = None

    def _load_from_state_dict(self, *args, **kwargs):
        self._net = None
        return super()._load_from_state_dict(*args, **kwargs)

    def net(self):
        """The kernel-side backbone built from this module's tensors (once per load)."""
        from .hybrid import HybridBackboneB200

        if self._net is None:
            cfg, dev = self.config, self.norm_f.weight.device
            if dev.type != "cuda":
                raise RuntimeError("zonos_b200 runs on CUDA (B200) only; move the model with .to('cuda')")
            layers, shape = [], None
            for blk in self.layers:
                m = blk.mixer
                p = {"norm.weight": blk.norm.weight, "norm.bias": getattr(blk.norm, "bias", None),
                     "in_proj.weight": m.in_proj.weight, "out_proj.weight": m.out_proj.weight}
                if isinstance(m, _Mamba2Params):
                    shape = m.shape
                    p.update({"kind": "mamba2", "conv1d.weight": m.conv1d.weight.squeeze(1), "conv1d.bias": m.conv1d.bias,
                              "dt_bias": m.dt_bias, "A_log": m.A_log, "D": m.D, "mixer_norm.weight": m.norm.weight})
                else:
                    p["kind"] = "attn"
                if hasattr(blk, "mlp"):
                    p.update({"norm2.weight": blk.norm2.weight, "norm2.bias": getattr(blk.norm2, "bias", None),
                              "fc1.weight": blk.mlp.fc1.weight, "fc2.weight": blk.mlp.fc2.weight})
                layers.append({k: v for k, v in p.items() if v is not None})
            assert shape is not None, "a hybrid backbone needs at least one Mamba2 layer"
            nf = {"weight": self.norm_f.weight}
            if getattr(self.norm_f, "bias", None) is not None:
                nf["bias"] = self.norm_f.bias
            self._net = HybridBackboneB200(layers, nf, shape, dev, cfg.attn_cfg["num_heads"], cfg.attn_cfg["num_heads_kv"],
                                           rms_norm=cfg.rms_norm, eps=cfg.norm_epsilon)
        return self._net

    def allocate_inference_cache(self, batch_size: int, max_seqlen: int, dtype: torch.dtype = torch.bfloat16):
        assert dtype == torch.bfloat16, "the B200 caches are bf16"
        return self.net().allocate_inference_cache(batch_size, max_seqlen)

    def forward(self, hidden_states: torch.Tensor, inference_params) -> torch.Tensor:
        """[B2, S, d] -> [B2, S, d]; inference_params.key_value_memory_dict is allocate_inference_cache's dict."""
        B2, S, d = hidden_states.shape
        cache = inference_params.key_value_memory_dict
        cache["seqlen_offset"] = int(inference_params.seqlen_offset)
        return self.net().forward(hidden_states.reshape(B2 * S, d).contiguous(), cache, seq_len=S).view(B2, S, d)


BACKBONES["b200_hybrid"] = B200HybridZonosBackbone
