"""In-kernel timeline of the weight-streaming GEMMs inside the Mamba2 stack's decode step (graph replay).
Build first:  ZB200_TRACE=1 ZB200_LIB_SUFFIX=_trace python -m zonos_b200.build ; run with ZB200_LIB_SUFFIX=_trace.
  python scripts/trace_hybrid.py [rows] [layers]"""
import ctypes
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from oracle import mamba2 as om
from zonos_b200._lib import lib
from zonos_b200.hybrid import Mamba2StackB200

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n_layers = int(sys.argv[2]) if len(sys.argv) > 2 else 4
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
tbuf = torch.zeros(64 * 160 * 12, dtype=torch.int64, device="cuda")
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    for _ in range(3):
        stack.step(x, states)
    side.synchronize()
    lib().zb200_debug_trace(ctypes.c_void_p(tbuf.data_ptr()))
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=side):
        stack.step(x, states)
    for _ in range(3):
        gr.replay()
    side.synchronize()
n = lib().zb200_debug_trace(None)
buf = tbuf.cpu().numpy().reshape(64, 160, 12)
ev = ["entry", "prologue", "w_issue0", "w_land0", "pdl_wait", "a_land0", "mma_done", "tmem_full0", "epi_done", "exit", "w_issued_all"]
t0 = buf[0][buf[0][:, 0] > 0, 0].min()
print(f"{rows} rows, {n_layers} layers, {n} GEMM launches traced (in_proj, out_proj per layer); times in us from the first CTA entry")
for l in range(min(n, 2 * n_layers)):
    live = buf[l][:, 0] > 0
    t = (buf[l][live][:, :11].astype(np.int64) - int(t0)) / 1e3
    print(f"--- launch {l} ({'in_proj' if l % 2 == 0 else 'out_proj'}), {int(live.sum())} CTAs")
    for e in range(11):
        col = t[:, e]
        print(f"  {ev[e]:13s} min {col.min():8.2f}  med {np.median(col):8.2f}  max {col.max():8.2f}")

flat = tbuf.cpu().numpy().reshape(64, 160 * 12)
for l in (2, 3):
    u = (flat[32 + l][:6 * 64].reshape(64, 6).astype(np.int64) - int(t0)) / 1e3
    print(f"--- launch {l}, CTA 0, per unit: w issued | loop start | w ready | a ready | MMAs issued | commits issued  (us)")
    for i in range(64):
        if flat[32 + l][6 * i + 1] == 0:
            break
        print(f"  unit {i:2d}  " + "  ".join(f"{u[i, k]:8.2f}" for k in range(6)))
