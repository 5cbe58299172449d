"""Diagnostic: is generate() deterministic across polling modes?  (tap => sync after every frame; poll N => N graph replays
queued back to back.)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import shapes
from zonos_b200.config import ZonosConfig
from zonos_b200.model import Zonos

cfg = ZonosConfig.from_dict(shapes.reference_config_dict("tiny"))
m = Zonos.from_state_dict(cfg, shapes.synth_transformer_weights("tiny", seed=0), "cuda")
B, Lc, N = 4, 6, 12
sp = dict(top_p=0, top_k=0, min_p=0.1, linear=0, conf=0, quad=0, repetition_penalty=3.0, repetition_penalty_window=2, temperature=1.0)
cond = shapes.synth_conditioning(B, Lc, 512, seed=3).cuda()
half = torch.cat([cond[0:2], cond[4:6]])

def run(c, b, **kw):
    out = m.generate(c, None, max_new_tokens=N, batch_size=b, progress_bar=False, sampling_params=dict(sp), seed=1234, row_base=0, **kw)
    return torch.stack([o.cpu() for o in out]), dict(m.last_generate_stats)

ref4, st = run(cond, 4, _tap=lambda *a: None)
print("stats", st)
for name, kw in [("tap", dict(_tap=lambda *a: None)), ("poll32", {}), ("poll32 again", {}), ("poll1", dict(_poll_every=1)), ("poll4", dict(_poll_every=4))]:
    o4, _ = run(cond, 4, **kw)
    o2, _ = run(half, 2, **kw)
    d4 = (o4 != ref4).nonzero()
    d2 = (o2 != ref4[:2]).nonzero()
    print(f"{name:14s} B=4 vs tapped B=4: {len(d4)} cells differ (first {d4[0].tolist() if len(d4) else None}); "
          f"B=2 vs tapped B=4[:2]: {len(d2)} differ (first {d2[0].tolist() if len(d2) else None})")
