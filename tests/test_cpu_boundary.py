"""CPU-only checks of the drop-in boundary: the C-ABI library loads and exports every declared symbol, the host-side
mirror keeps the reference's names/defaults/exceptions, and the multi-GPU sharding logic works (gloo, world size 2)."""
import ctypes
import inspect
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built_lib():
    from zonos_b200 import build

    return build.build()


def test_library_exports_every_declared_symbol(built_lib):
    from zonos_b200 import _lib

    names = _lib.declared_symbols()
    assert len(names) >= 20 and "zb200_engine_step" in names and "zb200_dac_decode" in names
    lib = ctypes.CDLL(built_lib)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.zb200_version() == 1
    lib.zb200_last_error.restype = ctypes.c_char_p
    assert isinstance(lib.zb200_last_error(), bytes)


def test_argument_validation_needs_no_gpu(built_lib):
    """Bad shapes are rejected before any CUDA call, with the reference's exception types in the shim."""
    from zonos_b200 import _lib

    lib = _lib.lib()
    assert lib.zb200_packed_weight_bytes(128, 100, 64, 0, 0, 0) == -1  # K not a multiple of 64
    assert lib.zb200_packed_weight_bytes(3072, 2048, 64, 0, 0, 0) == 3072 * 2048 * 2
    assert lib.zb200_packed_weight_bytes(9 * 1026, 2048, 64, 2, 1026, 1088) == 9 * 1088 * 2048 * 2
    with pytest.raises(ValueError):
        _lib.check(lib.zb200_layernorm(None, None, None, None, 4, 100, ctypes.c_float(1e-5), None), "layernorm")
    with pytest.raises(ValueError):
        _lib.check(lib.zb200_delay_apply(None, None, 1, 0, 4, ctypes.c_int64(1025), None), "delay")


def test_struct_layouts_match_the_header(built_lib):
    """ctypes mirrors of the C structs: sizes must equal what the compiler laid out (checked via a tiny C program)."""
    from zonos_b200.dac import DacLayer, DacPlan
    from zonos_b200.engine import EngineDesc
    from zonos_b200.sampler import SamplerArgs

    src = '#include <stdio.h>\n#include "zonos_b200.h"\nint main(){printf("%zu %zu %zu %zu\\n", sizeof(zb200_sampler_args), sizeof(zb200_engine_desc), sizeof(zb200_dac_layer), sizeof(zb200_dac_plan));return 0;}\n'
    exe = os.path.join(ROOT, "zonos_b200", "build", "sizes")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    with open(exe + ".c", "w") as f:
        f.write(src)
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), exe + ".c", "-o", exe], check=True)
    sizes = [int(x) for x in subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()]
    assert sizes == [ctypes.sizeof(SamplerArgs), ctypes.sizeof(EngineDesc), ctypes.sizeof(DacLayer), ctypes.sizeof(DacPlan)]


def test_generate_signature_matches_the_reference():
    from zonos_b200.model import Zonos

    sig = inspect.signature(Zonos.generate)
    names = list(sig.parameters)
    assert names[:10] == ["self", "prefix_conditioning", "audio_prefix_codes", "max_new_tokens", "cfg_scale", "batch_size",
                          "sampling_params", "progress_bar", "disable_torch_compile", "callback"]
    p = sig.parameters
    assert p["max_new_tokens"].default == 86 * 30 and p["cfg_scale"].default == 2.0 and p["batch_size"].default == 1
    assert p["progress_bar"].default is True and p["disable_torch_compile"].default is False
    for fn in ("from_pretrained", "from_local", "prepare_conditioning", "embed_codes", "setup_cache"):
        assert hasattr(Zonos, fn)


def test_sampling_param_defaults_follow_the_reference():
    from zonos_b200.sampler import resolve_sampling_params

    d = resolve_sampling_params(None)  # generate()'s default dict (zonos/model.py:232-241)
    assert (d["linear"], d["conf"], d["quad"], d["repetition_penalty"], d["repetition_penalty_window"]) == (0.55, 0.4, 0.0, 3.0, 2)
    u = resolve_sampling_params(dict(min_p=0.1, repetition_penalty=2.0))  # user dict replaces it wholesale
    assert u["linear"] == 0.0 and u["min_p"] == 0.1 and u["temperature"] == 1.0
    with pytest.raises(KeyError):
        resolve_sampling_params(dict(min_p=0.1))  # zonos/model.py:342 indexes repetition_penalty


def test_model_errors_and_state_dict_names():
    from oracle import shapes
    from zonos_b200.backbone import BACKBONES
    from zonos_b200.config import ZonosConfig
    from zonos_b200.model import Zonos

    cfg = ZonosConfig.from_dict(shapes.reference_config_dict("tiny"))
    m = Zonos(cfg)
    keys = set(m.state_dict())
    for k in ("backbone.layers.0.norm.weight", "backbone.layers.1.mixer.in_proj.weight", "backbone.layers.0.mixer.out_proj.weight",
              "backbone.layers.0.norm2.bias", "backbone.layers.0.mlp.fc1.weight", "backbone.layers.1.mlp.fc2.weight",
              "backbone.norm_f.weight", "embeddings.8.weight", "heads.0.weight", "prefix_conditioner.norm.weight"):
        assert k in keys, k
    m.load_state_dict(m.state_dict())
    assert m.heads[0].weight.shape[0] == 1026  # 1025 % 8 == 1 pad row (zonos/utils.py:22-37)
    with pytest.raises(KeyError):
        BACKBONES["no_such_backbone"]
    x = torch.zeros(2, 4, 512, dtype=torch.bfloat16)
    with pytest.raises(AssertionError):
        m.generate(x, cfg_scale=1.0, batch_size=1)
    with pytest.raises(ValueError):
        m.generate(x, batch_size=2)
    with pytest.raises(RuntimeError):  # CPU model: there is no CPU fallback
        m.generate(x, batch_size=1, max_new_tokens=4, progress_bar=False)


def test_shard_bounds_cover_everything():
    from zonos_b200.parallel import shard_bounds

    for n in (0, 1, 7, 64, 512, 513):
        for w in (1, 2, 3, 8):
            spans = [shard_bounds(n, w, r) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1


def _gloo_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist

    from zonos_b200.parallel import gather_codes, shard_bounds

    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 5
    lo, hi = shard_bounds(n, world, rank)
    local = [torch.full((9, 3 + i), 100 * i + 7, dtype=torch.int64) for i in range(lo, hi)]
    out = gather_codes(local, n, dst=0)
    if rank == 0:
        q.put([(tuple(o.shape), int(o[0, 0])) for o in out])
    dist.destroy_process_group()


def test_gather_codes_gloo_world2():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    got = q.get(timeout=120)
    [p.join(60) for p in procs]
    assert got == [((9, 3 + i), 100 * i + 7) for i in range(5)]


def test_patch_reference_on_the_imported_reference_class():
    """INTEGRATION.md patch 2, applied to the UNMODIFIED reference class (when /root/reference is present): the signature
    survives, generate() only needs config / state_dict() / device of the reference object, the reference's exceptions
    keep their types, and on a CPU model the call ends in the 'no CPU fallback' error instead of the reference's loop."""
    from oracle import ref_import, shapes

    if not ref_import.available():
        pytest.skip("reference checkout not present (GPU box)")
    import zonos_b200

    ref_import.import_reference()
    import zonos.config as zc
    import zonos.model as zm
    from zonos.backbone import BACKBONES as REF_BACKBONES

    before = inspect.signature(zm.Zonos.generate)
    cls = zonos_b200.patch_reference(zm.Zonos)
    try:
        assert cls is zm.Zonos and inspect.signature(zm.Zonos.generate) == before and hasattr(zm.Zonos, "generate_reference")
        cfg = zc.ZonosConfig.from_dict(shapes.reference_config_dict("tiny"))
        m = zm.Zonos(cfg, REF_BACKBONES["torch"]).to("cpu", torch.bfloat16)
        sd = m.state_dict()
        sd.update(shapes.synth_transformer_weights("tiny", seed=0, padded_heads=False))
        m.load_state_dict(sd)
        x = torch.zeros(2, 4, 512, dtype=torch.bfloat16)
        with pytest.raises(AssertionError):
            m.generate(x, cfg_scale=1.0, batch_size=1)
        with pytest.raises(ValueError):
            m.generate(x, batch_size=2)
        with pytest.raises(RuntimeError, match="CUDA"):
            m.generate(x, batch_size=1, max_new_tokens=4, progress_bar=False)
        stamp, mine = m._zb200  # built from the reference object's config + state dict, cached on it
        assert mine.heads[0].weight.shape[0] == 1026 and mine.config.backbone.n_layer == cfg.backbone.n_layer
        assert torch.equal(mine.state_dict()["backbone.layers.1.mlp.fc2.weight"], sd["backbone.layers.1.mlp.fc2.weight"])
        with pytest.raises(RuntimeError, match="CUDA"):
            m.generate(x, batch_size=1, max_new_tokens=4, progress_bar=False)
        assert m._zb200[1] is mine  # cached, not rebuilt
    finally:
        zm.Zonos.generate = zm.Zonos.generate_reference
        del zm.Zonos.generate_reference
