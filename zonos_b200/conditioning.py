"""Prefix conditioner container (zonos/conditioning.py:373-389).

The conditioner stack runs once per utterance before the hot path (out of scope for this
round; SURVEY 8(f) rank 3 lists it as a "next" row).  This module keeps the parameter names of the reference's
PrefixConditioner for the conditioner types that need no text front-end, so checkpoints load and
`prepare_conditioning` works for Fourier / Integer / Passthrough conditioners; the eSpeak phonemiser is not
re-implemented (the metric uses synthetic prefix_conditioning).
"""
import torch
import torch.nn as nn

from .config import PrefixConditionerConfig


class _Conditioner(nn.Module):
    def __init__(self, output_dim, name, cond_dim=None, projection="none", uncond_type="learned", **kwargs):
        super().__init__()
        self.name, self.output_dim = name, output_dim
        self.cond_dim = cond_dim = cond_dim or output_dim
        if projection == "linear":
            self.project = nn.Linear(cond_dim, output_dim)
        elif projection == "mlp":
            self.project = nn.Sequential(nn.Linear(cond_dim, output_dim), nn.SiLU(), nn.Linear(output_dim, output_dim))
        else:
            self.project = nn.Identity()
        self.uncond_vector = nn.Parameter(torch.zeros(output_dim)) if uncond_type == "learned" else None

    def apply_cond(self, *inputs):
        raise NotImplementedError

    def forward(self, inputs):
        if inputs is None:
            assert self.uncond_vector is not None
            return self.uncond_vector.data.view(1, 1, -1)
        cond = self.apply_cond(*inputs) if isinstance(inputs, tuple) else self.apply_cond(inputs)
        return self.project(cond.to(self.project_dtype()))

    def project_dtype(self):
        for p in self.project.parameters():
            return p.dtype
        return torch.bfloat16 if self.uncond_vector is None else self.uncond_vector.dtype


class FourierConditioner(_Conditioner):
    def __init__(self, output_dim, input_dim=1, std=1.0, min_val=0.0, max_val=1.0, **kwargs):
        assert output_dim % 2 == 0
        super().__init__(output_dim, **kwargs)
        self.register_buffer("weight", torch.randn([output_dim // 2, input_dim]) * std)
        self.input_dim, self.min_val, self.max_val = input_dim, min_val, max_val

    def apply_cond(self, x):
        assert x.shape[-1] == self.input_dim
        x = (x - self.min_val) / (self.max_val - self.min_val)
        f = 2 * torch.pi * x.to(self.weight.dtype) @ self.weight.T
        return torch.cat([f.cos(), f.sin()], dim=-1)


class IntegerConditioner(_Conditioner):
    def __init__(self, output_dim, min_val=0, max_val=512, **kwargs):
        super().__init__(output_dim, **kwargs)
        self.min_val, self.max_val = min_val, max_val
        self.int_embedder = nn.Embedding(max_val - min_val + 1, output_dim)

    def apply_cond(self, x):
        assert x.shape[-1] == 1
        return self.int_embedder(x.squeeze(-1) - self.min_val)


class PassthroughConditioner(_Conditioner):
    def __init__(self, output_dim, **kwargs):
        super().__init__(output_dim, **kwargs)

    def apply_cond(self, x):
        assert x.shape[-1] == self.cond_dim
        return x


class EspeakPhonemeConditioner(_Conditioner):
    """Holds the phoneme embedding table so checkpoints load; phonemisation itself (eSpeak) is outside this package."""

    N_SYMBOLS = 189  # len(symbols) of zonos/conditioning.py

    def __init__(self, output_dim, **kwargs):
        super().__init__(output_dim, **kwargs)
        self.phoneme_embedder = nn.Embedding(self.N_SYMBOLS, output_dim)

    def apply_cond(self, phoneme_ids, *rest):
        """Takes pre-computed phoneme ids [B, L] (int64)."""
        return self.phoneme_embedder(phoneme_ids)


_TYPES = {c.__name__: c for c in (PassthroughConditioner, EspeakPhonemeConditioner, FourierConditioner, IntegerConditioner)}


class PrefixConditioner(_Conditioner):
    def __init__(self, config: PrefixConditionerConfig, output_dim: int):
        super().__init__(output_dim, "prefix", projection=config.projection, uncond_type="none")
        self.conditioners = nn.ModuleList([_TYPES[c["type"]](output_dim, **c) for c in config.conditioners])
        self.norm = nn.LayerNorm(output_dim)
        self.required_keys = {c.name for c in self.conditioners if c.uncond_vector is None}

    def forward(self, cond_dict: dict) -> torch.Tensor:
        if not set(cond_dict).issuperset(self.required_keys):
            raise ValueError(f"Missing required keys: {self.required_keys - set(cond_dict)}")
        conds = [c(cond_dict.get(c.name)) for c in self.conditioners]
        max_bsz = max(map(len, conds))
        assert all(c.shape[0] in (max_bsz, 1) for c in conds)
        conds = [c.expand(max_bsz, -1, -1) for c in conds]
        return self.norm(self.project(torch.cat(conds, dim=-2)))
