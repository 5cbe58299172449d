"""Decode step of a Mamba2 layer stack at the hybrid's mixer shape (BASELINE configs[4]: batch 32 per GPU = 64 rows with
CFG; d_model 2048, d_inner 4096, 64 heads x 64, d_state 128), host-sequenced C-ABI kernels captured in a torch CUDA graph.
Algorithmic bytes per layer: in_proj + out_proj weights (51.6 MB) + SSM state read + write (2 MB per row).
usage: python scripts/bench_hybrid_stack.py [rows] [layers]"""
import sys

import torch

sys.path.insert(0, ".")
import bench
from oracle import mamba2 as om
from zonos_b200.hybrid import Mamba2StackB200

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n_layers = int(sys.argv[2]) if len(sys.argv) > 2 else 40
d = 2048
base = om.synth_mixer_params(d_model=d, seed=0, dtype=torch.bfloat16)
layers = []
for l in range(n_layers):  # distinct tensors per layer (HBM traffic must be real), same values
    p = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in base.items()}
    p["mixer_norm.weight"] = p.pop("norm.weight")
    p["norm.weight"] = torch.ones(d, dtype=torch.bfloat16)
    layers.append(p)
stack = Mamba2StackB200(layers, dict(weight=torch.ones(d, dtype=torch.bfloat16)), base["shape"], "cuda")
states = stack.allocate_inference_cache(rows)
x = torch.randn(rows, d, device="cuda").bfloat16()
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    for _ in range(3):
        out = stack.step(x, states)
    side.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        out = stack.step(x, states)
    for _ in range(5):
        graph.replay()
    side.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 50
    e0.record(side)
    for _ in range(n):
        graph.replay()
    e1.record(side)
    side.synchronize()
ms = e0.elapsed_time(e1) / n
s = base["shape"]
w_bytes = (stack.n_in_real * d + d * s["d_inner"]) * 2
st_bytes = 2 * rows * s["n_heads"] * s["head_dim"] * s["d_state"] * 2
algo = n_layers * (w_bytes + st_bytes)
hbm_peak, _ = bench.peaks()
print(f"Mamba2 stack decode: {n_layers} layers, {rows} rows: {ms:.3f} ms/step ({ms / n_layers * 1e3:.1f} us/layer, "
      f"{6 * n_layers + 1} launches), algorithmic {algo / 1e9:.2f} GB -> {algo / ms / 1e6:.0f} GB/s = "
      f"{algo / ms / 1e6 / hbm_peak:.3f} of the measured HBM peak; finite={bool(torch.isfinite(out).all())}")
