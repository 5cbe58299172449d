"""Kernel breakdown of the C3 prefill (batch 64, S = 138 -> 17,664 rows) from CUPTI (torch.profiler): per kernel name
the launches, total and mean duration, plus the first layer's launches in order.
  python scripts/timeline_prefill.py [--batch 64]"""
import argparse
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, ".")
import bench
from zonos_b200 import synth

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
a = ap.parse_args()
dev = torch.device("cuda", 0)


class A:
    shape = "real"


model, shape = bench.build_model(A, dev)
B, Lc, P = a.batch, 128, 9
cond = synth.conditioning(B, Lc, shape["d_model"], device=dev, seed=1)
prefix = torch.randint(0, 1024, (B, 9, P), device=dev)
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    loop, eng, delayed = model._open_session(cond, prefix, 2580, 2.0, B, dict(bench.C3_SAMPLING), 0, 0, None)
    S = Lc + P + 1
    eng.prefill(S)
    side.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        eng.prefill(S)
        side.synchronize()
    eng.close()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "memcpy" not in e.name.lower()]
evs.sort(key=lambda e: e.time_range.start)
t0 = evs[0].time_range.start
print(f"prefill batch {B} S {S}: {len(evs)} kernels, span {(evs[-1].time_range.end - t0) / 1e3:.2f} ms")
short = lambda n: n.replace("void zb::", "").replace("zb::", "").split("(")[0][:80]
for i, e in enumerate(evs[:14]):
    s, en = e.time_range.start - t0, e.time_range.end - t0
    print(f"{i:4d} {s:9.1f} -> {en:9.1f}  dur {en - s:8.1f}  {short(e.name)}")
agg = {}
for e in evs:
    d = agg.setdefault(short(e.name), [0, 0.0])
    d[0] += 1
    d[1] += e.time_range.end - e.time_range.start
print(f"{'kernel':82s} {'n':>4s} {'total ms':>9s} {'us/launch':>10s}")
for k, (n, d) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:82s} {n:4d} {d / 1e3:9.2f} {d / n:10.1f}")
