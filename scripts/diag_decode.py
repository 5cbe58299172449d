import sys, torch
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
from oracle import shapes, scenarios as sc
from oracle.gen_golden import MODEL_CASES
from zonos_b200.config import ZonosConfig
from zonos_b200.model import Zonos
import test_gpu_model as T

name = sys.argv[1] if len(sys.argv) > 1 else "b1_greedy"
cfgm = ZonosConfig.from_dict(shapes.reference_config_dict("tiny"))
m = Zonos.from_state_dict(cfgm, shapes.synth_transformer_weights("tiny", seed=0), "cuda")
B, Lc, P, N, sp, cfg, cond, pre = T._case_inputs(name)
taps = []
m.generate(cond.cuda(), None if pre is None else pre.cuda(), max_new_tokens=N, cfg_scale=cfg, batch_size=B,
           sampling_params=dict(sp), progress_bar=False, _noise=T._noise(77, 2 * (N + 9) + 2, B),
           _tap=lambda s, l, f: taps.append((l.cpu(), f.cpu())))
first = taps[0][1]
frames = torch.cat([f for _, f in taps[1:]], -1)
gl = torch.stack([l for l, _ in taps])
rl = T._teacher_forced_oracle(name, first, frames, len(taps) - 1)
print("stats", m.last_generate_stats)
for s in range(gl.shape[0]):
    f = torch.isfinite(rl[s]); fg = torch.isfinite(gl[s])
    d = (gl[s][f & fg] - rl[s][f & fg]).abs()
    print(s, "nonfinite gpu/ref", int((~fg).sum()), int((~f).sum()), "nan", int(torch.isnan(gl[s]).sum()), "maxdiff %.4f" % d.max().item(), "rms %.4f" % d.pow(2).mean().sqrt().item())
