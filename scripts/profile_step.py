"""Runs a few eager decode steps at a chosen batch / context for ncu (profile range = cudaProfilerStart/Stop)."""
import argparse
import sys

import torch

sys.path.insert(0, ".")
import bench
from zonos_b200 import synth
from zonos_b200.sampler import ST_POS

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--ctx", type=int, default=1432)
ap.add_argument("--frames", type=int, default=2)
ap.add_argument("--shape", default="real")
a = ap.parse_args()
args = bench.parse.__wrapped__() if hasattr(bench.parse, "__wrapped__") else None
dev = torch.device("cuda", 0)


class A:
    shape = a.shape


model, shape = bench.build_model(A, dev)
cond = synth.conditioning(a.batch, 24, shape["d_model"], device=dev)
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    loop, eng, delayed = model._open_session(cond, None, 2580, 2.0, a.batch, dict(bench.C3_SAMPLING), 1, 0, None)
    loop.sample_first(eng.logits)
    sargs = loop.args(eng.logits, True)
    eng.step(sargs)
    loop.state[ST_POS] = a.ctx
    eng.step(sargs)
    side.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    for _ in range(a.frames):
        eng.step(sargs)
    side.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    eng.close()
print("done", model.last_generate_stats)
