"""zonos_b200: B200-native implementation of coezbek/Zonos's batched generation hot path."""


def generate(model, *args, **kwargs):
    """Free-function form of Zonos.generate for an unmodified reference model object (zonos_b200/integrate.py)."""
    from .integrate import generate as _g

    return _g(model, *args, **kwargs)


def patch_reference(zonos_cls):
    from .integrate import patch_reference as _p

    return _p(zonos_cls)
