"""Diagnostic: does a batch of 4 consume more noise draws (whole-batch re-draws on a first EOS, model.py:380-393) than its
two half-batches?  Prints the draw counter and the first step at which utterance 0's frames differ."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import shapes
from zonos_b200.config import ZonosConfig
from zonos_b200.model import Zonos

cfg = ZonosConfig.from_dict(shapes.reference_config_dict("tiny"))
m = Zonos.from_state_dict(cfg, shapes.synth_transformer_weights("tiny", seed=0), "cuda")
B, Lc, N = 4, 6, 12
sp = dict(top_p=0, top_k=0, min_p=0.1, linear=0, conf=0, quad=0, repetition_penalty=3.0, repetition_penalty_window=2, temperature=1.0)
cond = shapes.synth_conditioning(B, Lc, 512, seed=3).cuda()

def run(c, b, rb):
    taps = []
    m.generate(c, None, max_new_tokens=N, batch_size=b, progress_bar=False, sampling_params=dict(sp), seed=1234, row_base=rb,
               _tap=lambda s, l, f: taps.append((l.cpu(), f.cpu())))
    return taps

w = run(cond, 4, 0)
h = run(torch.cat([cond[0:2], cond[4:6]]), 2, 0)
for s, ((lw, fw), (lh, fh)) in enumerate(zip(w, h)):
    fin = torch.isfinite(lh)
    dl = float((lw[:2][fin] - lh[fin]).abs().max())
    eos_w = (fw[:, 0, 0] == 1024).tolist()
    print(f"step {s}: max|dlogit| utt0-1 {dl:.4g}; frames equal {bool((fw[:2] == fh).all())}; cb0==EOS per row (B=4) {eos_w}; "
          f"cb0 EOS logit rows {[round(float(x), 2) for x in lw[:, 0, 1024]]} max logit {[round(float(x), 2) for x in lw[:, 0, :1024].max(-1).values]}")
