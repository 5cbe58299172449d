"""Kernel timeline of graph-replayed decode steps from CUPTI (torch.profiler): start / end of every kernel of one
frame relative to the frame's first kernel, so gaps and overlaps of the PDL chain are visible.
  python scripts/timeline_step.py --batch 64 --ctx 1432 [--layers 2]"""
import argparse
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, ".")
import bench
from zonos_b200 import synth
from zonos_b200.sampler import ST_POS

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--ctx", type=int, default=1432)
ap.add_argument("--layers", type=int, default=2)
ap.add_argument("--shape", default="real")
a = ap.parse_args()
dev = torch.device("cuda", 0)


class A:
    shape = a.shape


model, shape = bench.build_model(A, dev)
B = a.batch
cond = synth.conditioning(B, 24, shape["d_model"], device=dev)
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    sampling = dict(bench.C3_SAMPLING) if B > 1 else dict(temperature=0.0, repetition_penalty=1.0)
    loop, eng, delayed = model._open_session(cond, None, 2580, 2.0, B, sampling, 1, 0, None)
    loop.sample_first(eng.logits)
    sargs = loop.args(eng.logits, True)
    eng.step(sargs)
    loop.state[ST_POS] = a.ctx
    eng.capture(sargs)
    eng.run(10)
    side.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        eng.run(4)
        side.synchronize()
    eng.close()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "memcpy" not in e.name.lower()]
evs.sort(key=lambda e: e.time_range.start)
n_launch = len(evs) // 4
frame = evs[2 * n_launch:3 * n_launch]  # third replay
t0 = frame[0].time_range.start
print(f"batch {B} ctx {a.ctx}: {n_launch} kernels per frame, frame span {(frame[-1].time_range.end - t0):.1f} us")
prev_end = t0
per_layer = 0
short = lambda n: n.replace("void zb::", "").replace("zb::", "").split("(")[0][:70]
# print the first kernels up to the requested number of layers (a layer starts at every in_proj)
count = 0
for i, e in enumerate(frame):
    s, en = e.time_range.start - t0, e.time_range.end - t0
    nm = short(e.name)
    if count < 12 * a.layers + 3:
        print(f"{i:4d} {s:9.2f} -> {en:9.2f}  dur {en - s:7.2f}  gap-after-prev-end {s - (prev_end - t0):7.2f}  {nm}")
    prev_end = max(prev_end, e.time_range.end)
    count += 1
# aggregate: per kernel name, total duration and the "exposed" time (portion not overlapped by the previous kernels)
agg = {}
cursor = t0
for e in frame:
    s, en = e.time_range.start, e.time_range.end
    exposed = max(0.0, en - max(s, cursor))
    cursor = max(cursor, en)
    k = short(e.name)
    d = agg.setdefault(k, [0, 0.0, 0.0])
    d[0] += 1
    d[1] += en - s
    d[2] += exposed
print(f"{'kernel':72s} {'n':>4s} {'dur us':>9s} {'exposed us':>11s}")
for k, (n, d, x) in sorted(agg.items(), key=lambda kv: -kv[1][2]):
    print(f"{k:72s} {n:4d} {d:9.1f} {x:11.1f}")
