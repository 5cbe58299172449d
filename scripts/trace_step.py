"""In-kernel timeline of the decode step's weight-streaming GEMMs as the ENGINE launches them (build with
ZB200_TRACE=1): per launch the min / median / max over CTAs of each event, in us since the first traced event.
  python scripts/trace_step.py --batch 64 --ctx 1432 --first 8 --count 10"""
import argparse
import ctypes
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import bench
from zonos_b200 import synth
from zonos_b200._lib import lib
from zonos_b200.sampler import ST_POS

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--ctx", type=int, default=1432)
ap.add_argument("--first", type=int, default=8, help="first traced launch to print")
ap.add_argument("--count", type=int, default=10)
a = ap.parse_args()
dev = torch.device("cuda", 0)


class A:
    shape = "real"


model, shape = bench.build_model(A, dev)
B = a.batch
cond = synth.conditioning(B, 24, shape["d_model"], device=dev)
tbuf = torch.zeros(64 * 160 * 12, dtype=torch.int64, device=dev)
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    sampling = dict(bench.C3_SAMPLING) if B > 1 else dict(temperature=0.0, repetition_penalty=1.0)
    loop, eng, delayed = model._open_session(cond, None, 2580, 2.0, B, sampling, 1, 0, None)
    loop.sample_first(eng.logits)
    sargs = loop.args(eng.logits, True)
    eng.step(sargs)
    loop.state[ST_POS] = a.ctx
    side.synchronize()
    lib().zb200_debug_trace(ctypes.c_void_p(tbuf.data_ptr()))
    eng.capture(sargs)      # the trace slots are baked into the captured launches
    eng.run(3)
    side.synchronize()
    tbuf.zero_()
    eng.run(1)
    side.synchronize()
    n = lib().zb200_debug_trace(None)
    eng.close()
buf = tbuf.cpu().numpy().reshape(64, 160, 12)
ev = ["entry", "prologue", "w_issue0", "w_land0", "pdl_wait", "a_land0", "mma_done", "tmem_full0", "epi_done", "exit", "w_issued_all"]
t0 = None
print("traced launches:", n)
for l in range(a.first, min(64, a.first + a.count)):
    raw = buf[l, :, :11].astype(np.int64)
    live = raw[:, 0] > 0
    if not live.any():
        continue
    raw = raw[live]
    if t0 is None:
        t0 = raw[:, 0].min()
    t = (raw - int(t0)) / 1e3
    print(f"--- traced launch {l} ({live.sum()} CTAs)")
    for e in range(11):
        col = t[:, e][raw[:, e] > 0]
        if len(col):
            print(f"  {ev[e]:13s} min {col.min():8.2f}  med {np.median(col):8.2f}  max {col.max():8.2f}")
