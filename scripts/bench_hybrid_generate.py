"""BASELINE configs[4]-like run (hybrid, batch 32 per GPU): Zonos.generate on a random-init hybrid model of the recalled
Zonos-v0.1-hybrid shape (d_model 2048; 46 blocks, attention + GatedMLP(8192) at every 9th, Mamba2 elsewhere: d_inner 4096,
64 heads x 64, d_state 128 -- the real config.json is not available offline) through zonos_b200.hybrid.ZonosHybridB200,
i.e. host-driven eager launches (no per-frame CUDA graph yet): the number is host-bound and says so.
usage: python scripts/bench_hybrid_generate.py [batch] [new_frames]"""
import sys
import time

import torch

sys.path.insert(0, ".")
from zonos_b200 import synth

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
N_new = int(sys.argv[2]) if len(sys.argv) > 2 else 256
n_layers, d = synth.HYBRID["n_layer"], synth.HYBRID["d_model"]
model = synth.hybrid_model()
g = torch.Generator().manual_seed(0)
r = lambda *sh: torch.randn(*sh, generator=g)
BF = torch.bfloat16
cond = torch.nn.functional.layer_norm(r(2 * B, 128, d), (d,)).to(BF).cuda()
sp = dict(top_p=0, top_k=0, min_p=0.1, linear=0, conf=0, quad=0, repetition_penalty=3.0, repetition_penalty_window=2, temperature=1.0)
model.generate(cond, None, max_new_tokens=16, batch_size=B, sampling_params=sp, seed=1)


def run(n_new):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = model.generate(cond, None, max_new_tokens=n_new, batch_size=B, sampling_params=sp, seed=1)
    torch.cuda.synchronize()
    return time.perf_counter() - t0, out


t_short, _ = run(N_new // 4)
dt, out = run(N_new)
steps = model.last_generate_stats["steps"]
frames = sum(o.shape[1] for o in out)
per_frame = (dt - t_short) / (N_new - N_new // 4) * 1e3  # marginal cost of a frame: prefill / capture cancel out
print(f"hybrid generate (decode step replayed from a CUDA graph): batch {B}, {N_new} new frames, {n_layers} blocks: {dt:.2f} s -> "
      f"{frames / 86.1328 / dt:.1f} audio-s/s end to end (prefill + graph capture included); decode {per_frame:.2f} ms per frame = "
      f"{B / (per_frame * 1e-3) / 86.1328:.1f} audio-s/s; codes in range: {all(int(o.max()) < 1024 and int(o.min()) >= 0 for o in out)}")
