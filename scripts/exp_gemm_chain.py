"""Back-to-back (PDL-chained, graph-replayed) launches of ONE decode GEMM shape over distinct weights: us per launch and
effective weight GB/s for several M / grid settings.  Separates the streaming rate of the kernel from step latencies."""
import sys

import torch

sys.path.insert(0, ".")
from zonos_b200 import ops

dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
SHAPES = {"qkv": (3072, 2048), "out": (2048, 2048), "fc1": (16384, 2048), "fc2": (2048, 8192)}
NW = int(__import__("os").environ.get("NW", "10"))
cases = [a.split(":") for a in sys.argv[1:]] or [["fc1", "128", "128"], ["fc1", "128", "148"], ["fc1", "32", "256"], ["fc1", "2", "256"]]
packed = {}
side = torch.cuda.Stream()
for name, M, grid in cases:
    M, grid = int(M), int(grid)
    N, K = SHAPES[name]
    if name not in packed:
        packed[name] = [ops.pack_weight((torch.randn(N, K, device=dev, generator=g) / 45).bfloat16()) for _ in range(NW)]
    a = torch.randn(M, K, device=dev, generator=g).bfloat16()
    ws = torch.empty(max(grid, 160) * 2 * 2 * 32768 // 4 * 4, dtype=torch.float32, device=dev)
    tickets = torch.zeros(1024, dtype=torch.int32, device=dev)

    def chain():
        for rep in range(max(1, 10 // NW)):
            for w in packed[name]:
                ops.gemm_skinny_bf16(a, w, N, grid=grid, ws=ws, tickets=tickets)

    with torch.cuda.stream(side):
        chain()
        side.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=side):
            chain()
        gr.replay()
        side.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        for _ in range(5):
            gr.replay()
        e1.record(side)
        side.synchronize()
    us = e0.elapsed_time(e1) / 5 / (NW * max(1, 10 // NW)) * 1e3
    print(f"{name} M={M} grid={grid}: {us:.2f} us/launch, {N * K * 2 / us / 1e3:.0f} GB/s of weights", flush=True)
