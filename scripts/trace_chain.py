"""In-kernel timeline of the fused chain kernel (csrc/chain.cuh): runs a few eager decode steps with the trace hook on and
prints, per phase of one mid-stack launch, when (relative to the launch's first CTA start, medians / extremes over CTAs)
  W0   first weight stage of the phase landed        A-dep  the phase's A operand became available (grid barrier passed)
  MMA0 first MMA issued                              MMAn   last accumulator complete
  done CTA arrived on the phase's counter            fin-dep all CTAs arrived (reduction may start)    fin reduction done."""
import argparse
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import bench
from zonos_b200 import synth
from zonos_b200.sampler import ST_POS

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, nargs="+", default=[64, 1])
ap.add_argument("--ctx", type=int, default=1432)
ap.add_argument("--layer", type=int, default=12)
a = ap.parse_args()
dev = torch.device("cuda", 0)


class A:
    shape = "real"


model, shape = bench.build_model(A, dev)
sms = torch.cuda.get_device_properties(0).multi_processor_count
L = shape["n_layer"]
names = ["W0", "A-dep", "MMA0", "MMAn", "done", "fin-dep", "fin", "start"]
for B in a.batch:
    cond = synth.conditioning(B, 24, shape["d_model"], device=dev)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        sampling = dict(bench.C3_SAMPLING) if B > 1 else dict(temperature=0.0, repetition_penalty=1.0)
        loop, eng, delayed = model._open_session(cond, None, 2580, 2.0, B, sampling, 1, 0, None)
        loop.sample_first(eng.logits)
        sargs = loop.args(eng.logits, True)
        eng.step(sargs)
        loop.state[ST_POS] = a.ctx
        buf = torch.zeros((L + 1) * sms * 4 * 8, dtype=torch.int64, device=dev)
        eng.set_chain_trace(buf)
        for _ in range(3):
            eng.step(sargs)
        side.synchronize()
        t = buf.cpu().numpy().reshape(L + 1, sms, 4, 8).astype(np.float64)
        eng.close()
    t[t == 0] = np.nan
    j = a.layer + 1
    t0 = np.nanmin(t[j, :, 0, 7])
    prev_end = np.nanmax(t[j - 1])  # last stamp of the previous chain launch: the attention kernel sits in between
    print(f"=== batch {B}, chain launch after layer {a.layer}'s attention; previous chain launch ended {(t0 - prev_end) / 1e3:.1f} us "
          f"before this one's first CTA started; CTA starts spread over {(np.nanmax(t[j, :, 0, 7]) - t0) / 1e3:.1f} us")
    for p in range(4):
        row = []
        for ev in range(7):
            v = (t[j, :, p, ev] - t0) / 1e3
            if np.all(np.isnan(v)):
                row.append(f"{names[ev]:>7s}      -      ")
            else:
                row.append(f"{names[ev]:>7s} {np.nanmin(v):5.1f}/{np.nanmedian(v):5.1f}/{np.nanmax(v):5.1f}")
        print(f"phase {p}: " + " | ".join(row))
    print(f"launch span {(np.nanmax(t[j]) - t0) / 1e3:.1f} us; next launch's first CTA starts {(np.nanmin(t[j + 1, :, 0, 7]) - t0) / 1e3:.1f} us after this one's"
          if j + 1 <= L else "")
