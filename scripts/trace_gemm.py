"""In-kernel timeline of the weight-streaming GEMM (build with ZB200_TRACE=1 python -m zonos_b200.build --force).
Launches a chain of GEMMs (graph replay) and prints, per launch, the min/median/max over CTAs of each event time
relative to the first CTA start of the FIRST launch (us)."""
import ctypes
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from zonos_b200 import ops
from zonos_b200._lib import lib

M = int(sys.argv[1]) if len(sys.argv) > 1 else 2
names = sys.argv[2].split(",") if len(sys.argv) > 2 else ["qkv", "out", "fc1", "fc2"]
SHAPES = {"qkv": (3072, 2048), "out": (2048, 2048), "fc1": (16384, 2048), "fc2": (2048, 8192)}
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
L = 3
packed = {n: [ops.pack_weight((torch.randn(*SHAPES[n], device=dev, generator=g) / 45).bfloat16()) for _ in range(L)] for n in names}
acts = {K: torch.randn(M, K, device=dev, generator=g).bfloat16() for K in (2048, 8192)}
ws = torch.empty(160 * 32768, dtype=torch.float32, device=dev)
tickets = torch.zeros(1024, dtype=torch.int32, device=dev)
side = torch.cuda.Stream()
tbuf = torch.zeros(64 * 160 * 12, dtype=torch.int64, device=dev)


def chain():
    for l in range(L):
        for n in names:
            N, K = SHAPES[n]
            ops.gemm_skinny_bf16(acts[K], packed[n][l], N, grid=148, ws=ws, tickets=tickets)


with torch.cuda.stream(side):
    chain()
    side.synchronize()
    lib().zb200_debug_trace(ctypes.c_void_p(tbuf.data_ptr()))
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=side):
        chain()
    gr.replay()
    gr.replay()
    side.synchronize()
n = lib().zb200_debug_trace(None)
buf = tbuf.cpu().numpy().reshape(64, 160, 12)
print("launches traced", n)
ev = ["entry", "prologue", "w_issue0", "w_land0", "pdl_wait", "a_land0", "mma_done", "tmem_full0", "epi_done", "exit", "w_issued_all"]
t0 = buf[0, :148, 0].min()
order = [n_ for _ in range(L) for n_ in names]
for l in range(min(n, len(order))):
    t = (buf[l, :148, :11].astype(np.int64) - int(t0)) / 1e3
    print(f"--- launch {l} ({order[l]}, M={M})")
    for e in range(11):
        col = t[:, e]
        print(f"  {ev[e]:13s} min {col.min():8.2f}  med {np.median(col):8.2f}  max {col.max():8.2f}")
