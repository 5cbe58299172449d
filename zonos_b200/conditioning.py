"""Prefix conditioner on the device (zonos/conditioning.py:12-55, 295-389).

The modules below only DECLARE the reference's parameters, under the reference's names, so checkpoints load
(`prefix_conditioner.conditioners.{i}.{phoneme_embedder.weight | project.* | weight | int_embedder.weight |
uncond_vector}`, `prefix_conditioner.norm.*`).  They hold no arithmetic: PrefixConditioner.forward turns the `cond_dict`
into a table of segments and makes ONE call of zb200_prefix_conditioner (csrc/conditioning.cu: embedding rows / Fourier
features / projected pass-through vector / uncond vectors at their positions of the prefix, bf16 rounding where the eager
reference rounds, LayerNorm).  There is no CPU path.

Text -> phonemes (eSpeak) is host string processing and stays with the reference: the "espeak" entry of the cond_dict is
either an integer tensor of phoneme ids [B, L] (tokenize_phonemes' output, zonos/conditioning.py:181-191) or the reference's
(texts, languages) tuple, which needs the reference package with its eSpeak front-end installed.
"""
import ctypes

import torch
import torch.nn as nn

from .config import PrefixConditionerConfig

N_PHONEME_SYMBOLS = 189  # len(SPECIAL_TOKEN_IDS) + len(symbols) of zonos/conditioning.py:152-170
KIND_ROWS, KIND_VECTOR, KIND_FOURIER, KIND_LINEAR = range(4)


class CondSegment(ctypes.Structure):
    _fields_ = [("kind", ctypes.c_int32), ("start", ctypes.c_int32), ("len", ctypes.c_int32), ("in_dim", ctypes.c_int32),
                ("batch_stride", ctypes.c_int32), ("id_offset", ctypes.c_int32), ("min_val", ctypes.c_float),
                ("max_val", ctypes.c_float), ("table", ctypes.c_void_p), ("bias", ctypes.c_void_p), ("x", ctypes.c_void_p)]


class Conditioner(nn.Module):
    """Parameter holder for one conditioner.  Like the reference's base class (conditioning.py:12-40): `uncond_type`
    defaults to "none" (no uncond_vector: the conditioner is then a REQUIRED key), `projection` to "none"."""

    def __init__(self, output_dim: int, name: str, cond_dim: int | None = None, projection: str = "none",
                 uncond_type: str = "none", **kwargs):
        super().__init__()
        self.name, self.output_dim = name, output_dim
        self.cond_dim = cond_dim or output_dim
        if projection == "linear":
            self.project = nn.Linear(self.cond_dim, output_dim)
        elif projection == "none":
            self.project = nn.Identity()
        else:
            raise NotImplementedError(f"conditioner projection {projection!r} is not built on the device (the Zonos checkpoints "
                                      "use 'none' and 'linear')")
        self.uncond_vector = nn.Parameter(torch.zeros(output_dim)) if uncond_type == "learned" else None

    def segment(self, inputs, keep: list):
        """-> (CondSegment without start, batch size of the input).  `keep` collects tensors that must outlive the launch."""
        raise NotImplementedError

    def uncond_segment(self):
        assert self.uncond_vector is not None  # conditioning.py:46-48
        s = CondSegment()
        s.kind, s.len, s.table = KIND_VECTOR, 1, self.uncond_vector.data_ptr()
        return s, 1


class EspeakPhonemeConditioner(Conditioner):
    def __init__(self, output_dim: int, n_symbols: int = N_PHONEME_SYMBOLS, **kwargs):
        super().__init__(output_dim, **kwargs)
        self.phoneme_embedder = nn.Embedding(n_symbols, output_dim)

    def segment(self, inputs, keep):
        ids = inputs[0] if isinstance(inputs, (tuple, list)) and torch.is_tensor(inputs[0]) else inputs
        if not torch.is_tensor(ids):
            texts, languages = inputs
            try:  # the reference's own front-end (eSpeak + its symbol table), when that package is installed
                from zonos.conditioning import phonemize, tokenize_phonemes
            except Exception as e:
                raise RuntimeError("the 'espeak' conditioner was given (texts, languages) but the reference's phonemiser "
                                   "(zonos.conditioning, eSpeak) is not importable; pass tokenised phoneme ids [B, L]") from e
            ids, _ = tokenize_phonemes(phonemize(texts, languages))
        w = self.phoneme_embedder.weight
        ids = ids.to(device=w.device, dtype=torch.int32).contiguous()
        assert ids.ndim == 2 and int(ids.min()) >= 0 and int(ids.max()) < w.shape[0], "phoneme ids out of range"
        keep.append(ids)
        s = CondSegment()
        s.kind, s.len, s.table, s.x = KIND_ROWS, ids.shape[1], w.data_ptr(), ids.data_ptr()
        s.batch_stride = ids.shape[1] if ids.shape[0] > 1 else 0
        return s, ids.shape[0]


class FourierConditioner(Conditioner):
    def __init__(self, output_dim: int, input_dim: int = 1, std: float = 1.0, min_val: float = 0.0, max_val: float = 1.0, **kwargs):
        assert output_dim % 2 == 0
        super().__init__(output_dim, **kwargs)
        self.register_buffer("weight", torch.randn(output_dim // 2, input_dim) * std)
        self.input_dim, self.min_val, self.max_val = input_dim, float(min_val), float(max_val)

    def segment(self, inputs, keep):
        x = inputs[0] if isinstance(inputs, (tuple, list)) else inputs
        assert x.shape[-1] == self.input_dim and x.ndim == 3
        x = x.to(device=self.weight.device, dtype=torch.float32).contiguous()
        keep.append(x)
        s = CondSegment()
        s.kind, s.len, s.in_dim, s.min_val, s.max_val = KIND_FOURIER, x.shape[1], self.input_dim, self.min_val, self.max_val
        s.table, s.x = self.weight.data_ptr(), x.data_ptr()
        s.batch_stride = x.shape[1] * x.shape[2] if x.shape[0] > 1 else 0
        return s, x.shape[0]


class IntegerConditioner(Conditioner):
    def __init__(self, output_dim: int, min_val: int = 0, max_val: int = 512, **kwargs):
        super().__init__(output_dim, **kwargs)
        self.min_val, self.max_val = int(min_val), int(max_val)
        self.int_embedder = nn.Embedding(self.max_val - self.min_val + 1, output_dim)

    def segment(self, inputs, keep):
        x = inputs[0] if isinstance(inputs, (tuple, list)) else inputs
        assert x.shape[-1] == 1 and x.ndim == 3
        w = self.int_embedder.weight
        ids = x.squeeze(-1).to(device=w.device, dtype=torch.int32).contiguous()
        assert int(ids.min()) >= self.min_val and int(ids.max()) <= self.max_val, "integer conditioner input out of range"
        keep.append(ids)
        s = CondSegment()
        s.kind, s.len, s.id_offset, s.table, s.x = KIND_ROWS, ids.shape[1], -self.min_val, w.data_ptr(), ids.data_ptr()
        s.batch_stride = ids.shape[1] if ids.shape[0] > 1 else 0
        return s, ids.shape[0]


class PassthroughConditioner(Conditioner):
    def segment(self, inputs, keep):
        x = inputs[0] if isinstance(inputs, (tuple, list)) else inputs
        assert x.shape[-1] == self.cond_dim and x.ndim == 3
        if not isinstance(self.project, nn.Linear):
            raise NotImplementedError("a pass-through conditioner without a linear projection is not built on the device")
        w = self.project.weight
        x = x.to(device=w.device, dtype=torch.bfloat16).contiguous()
        keep.append(x)
        s = CondSegment()
        s.kind, s.len, s.in_dim, s.table, s.x = KIND_LINEAR, x.shape[1], self.cond_dim, w.data_ptr(), x.data_ptr()
        s.bias = None if self.project.bias is None else self.project.bias.data_ptr()
        s.batch_stride = x.shape[1] * x.shape[2] if x.shape[0] > 1 else 0
        return s, x.shape[0]


_TYPES = {c.__name__: c for c in (PassthroughConditioner, EspeakPhonemeConditioner, FourierConditioner, IntegerConditioner)}


class PrefixConditioner(nn.Module):
    """zonos/conditioning.py:373-389.  forward(cond_dict) -> bf16 [max batch, L_cond, d] on the device."""

    def __init__(self, config: PrefixConditionerConfig, output_dim: int):
        super().__init__()
        if config.projection != "none":
            raise NotImplementedError("a projection of the whole prefix is not built on the device (checkpoints use 'none')")
        self.name, self.output_dim = "prefix", output_dim
        self.conditioners = nn.ModuleList(_TYPES[c["type"]](output_dim, **c) for c in config.conditioners)
        self.norm = nn.LayerNorm(output_dim)
        self.required_keys = {c.name for c in self.conditioners if c.uncond_vector is None}

    def forward(self, cond_dict: dict) -> torch.Tensor:
        from ._lib import check, lib, stream_ptr

        if not set(cond_dict).issuperset(self.required_keys):
            raise ValueError(f"Missing required keys: {self.required_keys - set(cond_dict)}")
        w = self.norm.weight
        if not w.is_cuda or w.dtype != torch.bfloat16:
            raise RuntimeError("zonos_b200's prefix conditioner runs on CUDA in bf16 only (there is no CPU fallback)")
        keep, segs, sizes, pos = [], [], [], 0
        for c in self.conditioners:
            inputs = cond_dict.get(c.name)
            s, n = c.uncond_segment() if inputs is None else c.segment(inputs, keep)
            s.start = pos
            pos += s.len
            segs.append(s)
            sizes.append(n)
        if not segs:
            raise ValueError("the prefix conditioner has no conditioners")
        batch = max(sizes)
        assert all(n in (batch, 1) for n in sizes)  # conditioning.py:386-387
        arr = (CondSegment * len(segs))(*segs)
        out = torch.empty(batch, pos, self.output_dim, dtype=torch.bfloat16, device=w.device)
        check(lib().zb200_prefix_conditioner(arr, len(segs), batch, pos, self.output_dim, ctypes.c_void_p(w.data_ptr()),
                                             ctypes.c_void_p(self.norm.bias.data_ptr()), ctypes.c_float(self.norm.eps),
                                             ctypes.c_void_p(out.data_ptr()), stream_ptr()), "prefix_conditioner")
        return out
