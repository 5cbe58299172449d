"""Drop-in glue for an UNMODIFIED reference `zonos.model.Zonos` object (patch 2).

`generate(model, ...)` needs only three things of the reference object -- `model.config` (the dataclass tree of
zonos/config.py), `model.state_dict()` and `model.device` -- and runs the generation on a zonos_b200.Zonos built once from
them (cached on the object as `_zb200`; rebuilt when the parameters were reloaded).  `patch_reference(cls)` installs it as
`cls.generate`, keeping the reference's signature (zonos/model.py:224-245)."""
import dataclasses
import inspect

import torch

_GENERATE_PARAMS = ("prefix_conditioning", "audio_prefix_codes", "max_new_tokens", "cfg_scale", "batch_size",
                    "sampling_params", "progress_bar", "disable_torch_compile", "callback")


def _b200_model(model):
    from .config import ZonosConfig
    from .model import Zonos

    sd = model.state_dict()
    stamp = tuple((k, v.data_ptr(), v._version) for k, v in sd.items() if k.startswith("backbone.norm_f") or k.startswith("heads.0"))
    cached = getattr(model, "_zb200", None)
    if cached is not None and cached[0] == stamp:
        return cached[1]
    cfg = model.config
    cfg = ZonosConfig.from_dict(dataclasses.asdict(cfg)) if dataclasses.is_dataclass(cfg) else cfg
    skip = ("prefix_conditioner.", "spk_clone_model.")  # not on the hot path: the reference object keeps doing those
    from .autoencoder import DACAutoencoder

    mine = Zonos.from_state_dict(cfg, {k: v for k, v in sd.items() if not k.startswith(skip)}, str(model.device),
                                 autoencoder=DACAutoencoder(load_pretrained=False))  # generate() does not touch the codec
    object.__setattr__(model, "_zb200", (stamp, mine))
    return mine


@torch.inference_mode()
def generate(model, prefix_conditioning, audio_prefix_codes=None, max_new_tokens=86 * 30, cfg_scale=2.0, batch_size=1,
             sampling_params=None, progress_bar=True, disable_torch_compile=False, callback=None, **kw):
    """zonos/model.py:224-457 for a reference Zonos object, on the B200 engine.  Same arguments, defaults, return type
    (list of int64 [9, T_i] on the model device) and exceptions."""
    return _b200_model(model).generate(prefix_conditioning, audio_prefix_codes, max_new_tokens, cfg_scale, batch_size,
                                       sampling_params, progress_bar, disable_torch_compile, callback, **kw)


def patch_reference(zonos_cls):
    """Installs `generate` on the reference class (keeps the original as `generate_reference`).  Returns the class."""
    orig = zonos_cls.generate
    have = tuple(inspect.signature(orig).parameters)[1:]
    if have != _GENERATE_PARAMS:
        raise TypeError(f"unexpected reference signature for Zonos.generate: {have}")
    if not hasattr(zonos_cls, "generate_reference"):
        zonos_cls.generate_reference = orig

    def _generate(self, *args, **kwargs):
        return generate(self, *args, **kwargs)

    _generate.__doc__ = generate.__doc__
    _generate.__signature__ = inspect.signature(orig)
    zonos_cls.generate = _generate
    return zonos_cls
