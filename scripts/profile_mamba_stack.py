import sys, torch
sys.path.insert(0, ".")
from oracle import mamba2 as om
from zonos_b200.hybrid import Mamba2StackB200
rows, n_layers, d = 64, 8, 2048
base = om.synth_mixer_params(d_model=d, seed=0, dtype=torch.bfloat16)
layers = []
for l in range(n_layers):
    p = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in base.items()}
    p["mixer_norm.weight"] = p.pop("norm.weight"); p["norm.weight"] = torch.ones(d, dtype=torch.bfloat16)
    layers.append(p)
stack = Mamba2StackB200(layers, dict(weight=torch.ones(d, dtype=torch.bfloat16)), base["shape"], "cuda")
states = stack.allocate_inference_cache(rows)
x = torch.randn(rows, d, device="cuda").bfloat16()
for _ in range(3): stack.step(x, states)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
stack.step(x, states); torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
