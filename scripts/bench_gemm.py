"""Microbenchmark of the decode-step GEMMs (real Zonos-v0.1 layer shapes) on one B200: the weight-streaming stream-K
kernel (csrc/gemm_skinny.cuh) at several grid sizes vs the tiled split-K kernel, as chains of launches over DIFFERENT
layers' weights (working set >> L2) captured in a CUDA graph, timed with CUDA events.  Prints GB/s of weight bytes."""
import argparse
import json
import sys

import torch

sys.path.insert(0, ".")
from zonos_b200 import ops

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=10)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--grids", default="74,111,148")
ap.add_argument("--rows", default="2,128")
a = ap.parse_args()
dev = "cuda"
SHAPES = {"qkv": (3072, 2048), "out": (2048, 2048), "fc1": (16384, 2048), "fc2": (2048, 8192)}
g = torch.Generator(device=dev).manual_seed(0)
packed = {}
for name, (N, K) in SHAPES.items():
    packed[name] = [ops.pack_weight((torch.randn(N, K, device=dev, generator=g) / K ** 0.5).bfloat16()) for _ in range(a.layers)]
ws = torch.empty(296 * 32768, dtype=torch.float32, device=dev)
tickets = torch.zeros(1024, dtype=torch.int32, device=dev)
side = torch.cuda.Stream()


def timed(fn):
    """fn() launches one chain; capture it in a graph, replay reps times."""
    with torch.cuda.stream(side):
        fn()
        side.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=side):
            fn()
        gr.replay()
        side.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        for _ in range(a.reps):
            gr.replay()
        e1.record(side)
        side.synchronize()
    return e0.elapsed_time(e1) / a.reps


res = []
for M in [int(r) for r in a.rows.split(",")]:
    acts = {K: torch.randn(M, K, device=dev, generator=g).bfloat16() for K in (2048, 8192)}
    for name, (N, K) in SHAPES.items():
        wbytes = N * K * 2
        x = acts[K]
        # tiled kernel with the engine's split plan
        n_tiles = N // 64
        s = 1
        if n_tiles < 148:
            s = 148 // n_tiles
            while s > 1 and (K // 64) // s < 4:
                s -= 1
        wst = torch.empty(n_tiles * s * 64 * 128, dtype=torch.float32, device=dev)
        tk = torch.zeros(n_tiles, dtype=torch.int32, device=dev)

        def old():
            for l in range(a.layers):
                ops.gemm_bf16(x, packed[name][l], N, splits=s)

        ms = timed(old) / a.layers
        res.append(dict(M=M, gemm=name, kernel="tiled", splits=s, us=round(ms * 1e3, 2), gbs=round(wbytes / ms / 1e6, 1)))
        print(res[-1], flush=True)
        for grid in [int(v) for v in a.grids.split(",")]:
            def new():
                for l in range(a.layers):
                    ops.gemm_skinny_bf16(x, packed[name][l], N, grid=grid, ws=ws, tickets=tickets)

            ms = timed(new) / a.layers
            res.append(dict(M=M, gemm=name, kernel="skinny", grid=grid, us=round(ms * 1e3, 2), gbs=round(wbytes / ms / 1e6, 1)))
            print(res[-1], flush=True)
    # a whole layer's four GEMMs back to back (what the decode step does, minus LN and attention)
    for grid in [int(v) for v in a.grids.split(",")]:
        def layer_chain():
            for l in range(a.layers):
                for name, (N, K) in SHAPES.items():
                    ops.gemm_skinny_bf16(acts[K], packed[name][l], N, grid=grid, ws=ws, tickets=tickets)

        ms = timed(layer_chain) / a.layers
        wb = sum(N * K * 2 for N, K in SHAPES.values())
        res.append(dict(M=M, gemm="layer(4)", kernel="skinny", grid=grid, us=round(ms * 1e3, 2), gbs=round(wb / ms / 1e6, 1)))
        print(res[-1], flush=True)
json.dump(res, open("gpurun_out/bench_gemm.json", "w"), indent=1)
