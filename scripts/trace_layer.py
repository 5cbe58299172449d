"""In-kernel timeline of one decode layer's GEMM chain as the engine launches it (build with ZB200_TRACE=1):
in_proj (ticketed, 3 contributors) -> out_proj (deferred) + resid_ln -> fc1 -> fc2 (deferred) + resid_ln, graph-replayed.
Prints per launch the min/median/max over CTAs of each event (us since the first CTA of the first launch)."""
import ctypes
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from zonos_b200 import ops
from zonos_b200._lib import check, lib, ptr, stream_ptr

M = int(sys.argv[1]) if len(sys.argv) > 1 else 2
L = 3
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
SHAPES = {"qkv": (3072, 2048), "out": (2048, 2048), "fc1": (16384, 2048), "fc2": (2048, 8192)}
packed = {n: [ops.pack_weight((torch.randn(*SHAPES[n], device=dev, generator=g) / 45).bfloat16()) for _ in range(L)] for n in SHAPES}
a2k = torch.randn(M, 2048, device=dev, generator=g).bfloat16()
a8k = torch.randn(M, 8192, device=dev, generator=g).bfloat16()
x = torch.randn(M, 2048, device=dev, generator=g).bfloat16()
xn = torch.empty_like(x)
lw = torch.ones(2048, device=dev).bfloat16()
lb = torch.zeros(2048, device=dev).bfloat16()
ws = torch.empty(160 * 32768, dtype=torch.float32, device=dev)
tickets = torch.zeros(1024, dtype=torch.int32, device=dev)
tbuf = torch.zeros(64 * 160 * 12, dtype=torch.int64, device=dev)
small = M <= 32
cap = 296 if small else 148
GRID = {"qkv": 144, "fc1": 256 if small else 128}
SPLITS = 8
side = torch.cuda.Stream()


def resid(a, name, l):
    N, K = SHAPES[name]
    check(lib().zb200_gemm_resid_ln(ptr(a), M, K, ptr(packed[name][l]), N, K, SPLITS, ptr(ws), ctypes.c_int64(ws.numel()),
                                    ptr(x), ptr(lw), ptr(lb), ptr(xn), ctypes.c_float(1e-5), stream_ptr()), "resid")


def chain():
    for l in range(L):
        ops.gemm_skinny_bf16(a2k, packed["qkv"][l], 3072, grid=GRID["qkv"], ws=ws, tickets=tickets)
        resid(a2k, "out", l)
        ops.gemm_skinny_bf16(a2k, packed["fc1"][l], 16384, grid=GRID["fc1"], ws=ws, tickets=tickets)
        resid(a8k, "fc2", l)


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
ev = ["entry", "prologue", "w_issue0", "w_land0", "pdl_wait", "a_land0", "mma_done", "tmem_full0", "epi_done", "exit", "w_issued_all"]
names = ["qkv", "out", "rln", "fc1", "fc2", "rln"] * L
t0 = None
for l in range(min(n, len(names))):
    raw = buf[l, :, :11].astype(np.int64)
    live = raw[:, 0] > 0
    raw = raw[live]
    if t0 is None:
        t0 = raw[:, 0].min()
    t = (raw - int(t0)) / 1e3
    print(f"--- launch {l} ({names[l]}, M={M}, {live.sum()} CTAs traced)")
    for e in range(11):
        col = t[:, e]
        print(f"  {ev[e]:13s} min {col.min():8.2f}  med {np.median(col):8.2f}  max {col.max():8.2f}")
