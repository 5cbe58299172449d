"""Kernel timeline (CUPTI) of the graph-replayed decode step of a Mamba2 layer stack at the hybrid's mixer shape.
  python scripts/timeline_hybrid.py [rows] [layers]"""
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, ".")
from oracle import mamba2 as om
from zonos_b200.hybrid import Mamba2StackB200

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n_layers = int(sys.argv[2]) if len(sys.argv) > 2 else 8
d = 2048
base = om.synth_mixer_params(d_model=d, seed=0, dtype=torch.bfloat16)
layers = []
for l in range(n_layers):
    p = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in base.items()}
    p["mixer_norm.weight"] = p.pop("norm.weight")
    p["norm.weight"] = torch.ones(d, dtype=torch.bfloat16)
    layers.append(p)
stack = Mamba2StackB200(layers, dict(weight=torch.ones(d, dtype=torch.bfloat16)), base["shape"], "cuda")
states = stack.allocate_inference_cache(rows)
x = torch.randn(rows, d, device="cuda").bfloat16()
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    for _ in range(3):
        out = stack.step(x, states)
    side.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        out = stack.step(x, states)
    for _ in range(5):
        graph.replay()
    side.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(4):
            graph.replay()
        side.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "memcpy" not in e.name.lower()
       and "memset" not in e.name.lower()]
evs.sort(key=lambda e: e.time_range.start)
n_launch = len(evs) // 4
frame = evs[2 * n_launch:3 * n_launch]
t0 = frame[0].time_range.start
print(f"{rows} rows, {n_layers} layers: {n_launch} kernels per step, span {(frame[-1].time_range.end - t0):.1f} us "
      f"= {(frame[-1].time_range.end - t0) / n_layers:.1f} us per layer")
short = lambda n: n.replace("void zb::", "").replace("zb::", "").split("(")[0][:70]
prev_end = t0
for i, e in enumerate(frame[:20]):
    s, en = e.time_range.start - t0, e.time_range.end - t0
    print(f"{i:4d} {s:9.2f} -> {en:9.2f}  dur {en - s:7.2f}  after-prev-end {e.time_range.start - prev_end:7.2f}  {short(e.name)}")
    prev_end = max(prev_end, e.time_range.end)
agg, cursor = {}, t0
for e in frame:
    s, en = e.time_range.start, e.time_range.end
    exposed = max(0.0, en - max(s, cursor))
    cursor = max(cursor, en)
    k = agg.setdefault(short(e.name), [0, 0.0, 0.0])
    k[0] += 1
    k[1] += en - s
    k[2] += exposed
print(f"{'kernel':72s} {'n':>4s} {'dur us':>9s} {'exposed us':>11s}")
for k, (n, dsum, x) in sorted(agg.items(), key=lambda kv: -kv[1][2]):
    print(f"{k:72s} {n:4d} {dsum:9.1f} {x:11.1f}")
