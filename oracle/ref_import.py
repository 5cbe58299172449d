"""TEST INFRASTRUCTURE ONLY. Imports the *unmodified* reference (coezbek/Zonos) from /root/reference.

Used only by oracle/gen_golden.py and by the CPU tests that cross-check the oracle restatement
against the live reference (skipped when /root/reference is absent, e.g. on the GPU box).

The reference's text front-end needs inflect/kanjize/phonemizer/sudachipy (zonos/conditioning.py:64-69),
none of which are installed here; they are only touched by EspeakPhonemeConditioner.apply_cond, which the
hot path (synthetic prefix_conditioning) never calls, so they are stubbed.  DacModel.from_pretrained
(zonos/autoencoder.py:15) needs the HF hub; it is replaced by a random-init DacModel(DacConfig(44100)).
"""
import os
import sys
import types

REF_ROOT = os.environ.get("ZONOS_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "zonos"))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules.setdefault(name, m)
    return sys.modules[name]


def import_reference():
    """Returns the reference `zonos` package (model, sampling, codebook_pattern, ... importable)."""
    if not available():
        raise RuntimeError("reference not present at %s" % REF_ROOT)
    if "zonos.model" in sys.modules:
        return sys.modules["zonos"]

    class _Any:
        def __init__(self, *a, **k):
            pass

        def __call__(self, *a, **k):
            return self

        def __getattr__(self, k):
            return _Any()

    _stub("inflect", engine=_Any)
    _stub("kanjize", number2kanji=lambda n: str(n))
    ph = _stub("phonemizer")
    be = _stub("phonemizer.backend", EspeakBackend=_Any)
    ph.backend = be
    _stub("sudachipy", Dictionary=_Any, SplitMode=_Any())

    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    import torch
    import zonos.utils as zu

    zu.set_device("cpu")
    from transformers.models.dac import DacConfig, DacModel
    import zonos.autoencoder as zae

    def _fake_from_pretrained(*a, **k):
        torch.manual_seed(1234)
        return DacModel(DacConfig(sampling_rate=44100))

    zae.DacModel = types.SimpleNamespace(from_pretrained=_fake_from_pretrained)
    import zonos.model  # noqa: F401
    import zonos

    return zonos
