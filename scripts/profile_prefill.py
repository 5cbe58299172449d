"""Runs the C3 prefill (batch 64, L_cond 128, 9 prefix frames: 17,664 rows) once warm and once inside a
cudaProfilerStart/Stop range, for ncu (--profile-from-start off)."""
import sys

import torch

sys.path.insert(0, ".")
import bench
from zonos_b200 import synth


class A:
    shape = "real"


dev = torch.device("cuda", 0)
model, shape = bench.build_model(A, dev)
B, Lc, P = 64, 128, 9
cond = synth.conditioning(B, Lc, shape["d_model"], device=dev, seed=1)
prefix = torch.randint(0, 1024, (B, 9, P), device=dev)
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    loop, eng, delayed = model._open_session(cond, prefix, 16, 2.0, B, dict(bench.C3_SAMPLING), 1, 0, None)  # warm prefill
    side.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    eng.prefill(Lc + P + 1)
    side.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    print("prefill launches", eng.launches_per_step())
    eng.close()
