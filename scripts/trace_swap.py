"""In-kernel timeline of one batch-1 decode layer's GEMM chain on the operand-swapped kernel (build with ZB200_TRACE=1):
in_proj -> out_proj -> LayerNorm -> fc1 -> fc2 -> LayerNorm, graph-replayed over different layers' weights."""
import ctypes
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from zonos_b200 import ops
from zonos_b200._lib import check, lib, ptr, stream_ptr

M, L, dev = 2, 3, "cuda"
g = torch.Generator(device=dev).manual_seed(0)
SHAPES = {"qkv": (3072, 2048, 192), "out": (2048, 2048, 128), "fc1": (16384, 2048, 256), "fc2": (2048, 8192, 128)}
packed = {n: [ops.pack_weight((torch.randn(s[0], s[1], device=dev, generator=g) / 45).bfloat16()) for _ in range(L)] for n, s in SHAPES.items()}
a2k = torch.randn(M, 2048, device=dev, generator=g).bfloat16()
a8k = torch.randn(M, 8192, device=dev, generator=g).bfloat16()
lw, lb = torch.ones(2048, device=dev).bfloat16(), torch.zeros(2048, device=dev).bfloat16()
tbuf = torch.zeros(64 * 160 * 12, dtype=torch.int64, device=dev)
ws = torch.empty(1024 * 2048, dtype=torch.float32, device=dev)
tickets = torch.zeros(1024, dtype=torch.int32, device=dev)
outs = {n: torch.empty(M, s[0], dtype=torch.bfloat16, device=dev) for n, s in SHAPES.items()}
side = torch.cuda.Stream()


def gemm(name, a, l):
    N, K, grid = SHAPES[name]
    check(lib().zb200_gemm_swap_bf16(ptr(a), M, K, ptr(packed[name][l]), N, K, ptr(outs[name]), N, grid, ptr(ws),
                                     ctypes.c_int64(ws.numel()), ptr(tickets), tickets.numel(), stream_ptr()), "swap")


def ln():
    check(lib().zb200_layernorm(ptr(a2k), ptr(lw), ptr(lb), ptr(outs["out"]), M, 2048, ctypes.c_float(1e-5), stream_ptr()), "ln")


def chain():
    for l in range(L):
        gemm("qkv", a2k, l)
        gemm("out", a2k, l)
        ln()
        gemm("fc1", a2k, l)
        gemm("fc2", a8k, l)
        ln()


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
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(side)
    for _ in range(10):
        gr.replay()
    e1.record(side)
    side.synchronize()
print("us per layer chain:", e0.elapsed_time(e1) / 10 / L * 1e3)
n = lib().zb200_debug_trace(None)
buf = tbuf.cpu().numpy().reshape(64, 160, 12)
ev = ["entry", "prologue", "-", "w_land0", "pdl_wait", "a_land0", "mma_done", "tmem_full0", "epi_done", "exit"]
names = ["qkv", "out", "fc1", "fc2"] * L
t0 = None
for l in range(min(n, len(names))):
    raw = buf[l, :, :10].astype(np.int64)
    raw = raw[raw[:, 0] > 0]
    if t0 is None:
        t0 = raw[:, 0].min()
    t = (raw - int(t0)) / 1e3
    print(f"--- launch {l} ({names[l]}, {len(raw)} CTAs traced)")
    for e in (0, 3, 4, 5, 6, 7, 8, 9):
        col = t[:, e]
        print(f"  {ev[e]:11s} min {col.min():8.2f}  med {np.median(col):8.2f}  max {col.max():8.2f}")
