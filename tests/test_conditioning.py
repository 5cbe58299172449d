"""Prefix conditioner: parameter names / required keys against the reference (CPU), device output
against the reference's own PrefixConditioner.forward (tests/golden/conditioning.pt, `python -m oracle.gen_golden
--conditioning`).  Tolerance (bf16 path): every conditioner row is rounded to bf16 before the norm in both implementations;
the Fourier features go through cos / sin of a bf16-rounded phase, where one input ulp moves the output by up to |phase| ulps,
so the gate is on the NORMALISED output: max|d| <= 0.0625 (2 bf16 ulps at |y| ~ 4) and rel-RMS <= 5e-3."""
import pytest
import torch

from oracle.gen_golden import COND_D, COND_LIST, cond_inputs


def _ours():
    from zonos_b200.conditioning import PrefixConditioner
    from zonos_b200.config import PrefixConditionerConfig

    return PrefixConditioner(PrefixConditionerConfig(conditioners=[dict(c) for c in COND_LIST], projection="none"), COND_D)


def test_names_and_required_keys_match_the_reference(golden):
    g = golden("conditioning.pt")
    pc = _ours()
    assert sorted(pc.required_keys) == g["required_keys"] == ["espeak"]  # uncond_type defaults to "none" (conditioning.py:19)
    ours = {k: tuple(v.shape) for k, v in pc.state_dict().items()}
    ref = {k: tuple(v.shape) for k, v in g["state_dict"].items()}
    assert ours == ref
    pc.load_state_dict(g["state_dict"])
    with pytest.raises(ValueError, match="Missing required keys"):
        pc.to(torch.bfloat16)({"speaker": torch.zeros(1, 1, 128)})
    with pytest.raises(RuntimeError, match="CUDA"):  # no CPU fallback
        pc({"espeak": torch.zeros(1, 4, dtype=torch.long)})


def test_against_the_live_reference_class():
    from oracle import ref_import

    if not ref_import.available():
        pytest.skip("reference checkout not present (GPU box)")
    ref_import.import_reference()
    import zonos.conditioning as zcond
    import zonos.config as zc

    ref = zcond.PrefixConditioner(zc.PrefixConditionerConfig(conditioners=[dict(c) for c in COND_LIST], projection="none"), COND_D)
    pc = _ours()
    assert pc.required_keys == ref.required_keys
    assert {k: tuple(v.shape) for k, v in pc.state_dict().items()} == {k: tuple(v.shape) for k, v in ref.state_dict().items()}
    assert [type(c).__name__ for c in pc.conditioners] == [type(c).__name__ for c in ref.conditioners]


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["full", "uncond", "partial"])
def test_device_output_vs_reference(golden, which):
    g = golden("conditioning.pt")
    pc = _ours()
    pc.load_state_dict(g["state_dict"])
    pc = pc.to("cuda", torch.bfloat16)
    inp = cond_inputs()
    keys = {"full": list(inp), "uncond": ["espeak"], "partial": ["espeak", "speaker", "language_id"]}[which]
    out = pc({k: inp[k] for k in keys}).cpu().float()
    want = g[which].float()
    assert out.shape == want.shape
    d = out - want
    worst, rel = float(d.abs().max()), float(d.pow(2).mean().sqrt() / want.pow(2).mean().sqrt())
    per_pos = d.abs().amax(dim=(0, 2))
    print(f"{which}: max|d| {worst:.4f}, rel-RMS {rel:.2e}; per position max {[round(float(v), 3) for v in per_pos]}")
    assert worst <= 0.0625 and rel <= 5e-3


@pytest.mark.gpu
def test_prepare_conditioning_end_to_end(golden):
    """Zonos.prepare_conditioning (zonos/model.py:210-218): [cond ; uncond] prefixes of equal length from one cond_dict."""
    from oracle import shapes
    from zonos_b200.config import ZonosConfig
    from zonos_b200.model import Zonos

    g = golden("conditioning.pt")
    cfg = shapes.reference_config_dict("tiny")
    cfg["prefix_conditioner"] = dict(conditioners=[dict(c) for c in COND_LIST], projection="none")
    m = Zonos(ZonosConfig.from_dict(cfg))
    sd = m.state_dict()
    sd.update({"prefix_conditioner." + k: v for k, v in g["state_dict"].items()})
    m.load_state_dict(sd)
    m = m.to("cuda", torch.bfloat16)
    pre = m.prepare_conditioning(cond_inputs()).cpu().float()
    want = torch.cat([g["full"], g["uncond"]]).float()
    assert pre.shape == want.shape == (4, 17, 512) and float((pre - want).abs().max()) <= 0.0625
