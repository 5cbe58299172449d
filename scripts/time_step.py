"""Times the graph-replayed decode step alone (CUDA events around N replays) at a chosen batch and starting context:
the precise A/B metric for kernel changes (bench.py's short runs include prefill / capture overheads)."""
import argparse
import sys

import torch

sys.path.insert(0, ".")
import bench
from zonos_b200 import synth
from zonos_b200.sampler import ST_POS

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, nargs="+", default=[64, 1])
ap.add_argument("--ctx", type=int, default=1432)
ap.add_argument("--frames", type=int, default=200)
ap.add_argument("--shape", default="real")
a = ap.parse_args()
dev = torch.device("cuda", 0)


class A:
    shape = a.shape


model, shape = bench.build_model(A, dev)
hbm_peak, _ = bench.peaks()
for B in a.batch:
    cond = synth.conditioning(B, 24, shape["d_model"], device=dev)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        sampling = dict(bench.C3_SAMPLING) if B > 1 else dict(temperature=0.0, repetition_penalty=1.0)
        loop, eng, delayed = model._open_session(cond, None, 2580, 2.0, B, sampling, 1, 0, None)
        loop.sample_first(eng.logits)
        sargs = loop.args(eng.logits, True)
        eng.step(sargs)
        start = max(30, a.ctx - a.frames // 2)
        loop.state[ST_POS] = start
        eng.capture(sargs)
        eng.run(20)
        side.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        eng.run(a.frames)
        e1.record(side)
        side.synchronize()
        ms = e0.elapsed_time(e1) / a.frames
        launches = eng.launches_per_step()
        eng.close()
    mean_ctx = start + 20 + a.frames / 2
    algo = bench.step_bytes(shape, B, mean_ctx)
    print(f"batch {B} ctx~{mean_ctx:.0f}: {ms:.4f} ms/frame, {launches} launches, hbm_frac {algo / (ms * 1e-3) / 1e9 / hbm_peak:.4f}, "
          f"{B / (ms * 1e-3) / bench.FRAME_RATE:.1f} audio-s/s", flush=True)
