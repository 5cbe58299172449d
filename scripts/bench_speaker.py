"""Speaker encoder (ResNet293 architecture, seeded weights): one 10 s clip at 44.1 kHz through SpeakerEmbeddingLDA.forward on
the device (CUDA events, host signal -> embedding on the host) against the CPU oracle on this box's cores, and the kernel
breakdown.   python scripts/bench_speaker.py [seconds] [--no-cpu]"""
import sys
import time

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, ".")
from oracle import speaker as osp
from oracle.gen_golden import speaker_wav
from zonos_b200.speaker_cloning import SpeakerEmbeddingLDA

seconds = float(sys.argv[1]) if len(sys.argv) > 1 and not sys.argv[1].startswith("-") else 10.0
sd, lda = osp.synth_speaker_weights(64, osp.RESNET293_BLOCKS, 13)
enc = SpeakerEmbeddingLDA("cuda", state_dict=sd, lda_state_dict=lda)
wav = speaker_wav(5, seconds, 44100, 2)
for _ in range(2):
    emb, out = enc(wav, 44100)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 5
e0.record()
for _ in range(n):
    emb, out = enc(wav, 44100)
    host = out.cpu()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
flop = 0
T = 1 + int(seconds * 16000) // 160
h, w, cin = 80, T, 64
for li, (mult, nb) in enumerate(zip((1, 2, 4, 8), osp.RESNET293_BLOCKS)):
    for bi in range(nb):
        st = (1 if li == 0 else 2) if bi == 0 else 1
        h, w = (h + st - 1) // st, (w + st - 1) // st
        flop += 2 * 9 * h * w * 64 * mult * (cin + 64 * mult)
        cin = 64 * mult
print(f"speaker encoder, {seconds:.0f} s clip (44.1 kHz stereo -> 16 kHz, {T} frames): {ms:.1f} ms on the device "
      f"({flop / 1e12:.2f} TFLOP of 3x3 convolutions -> {flop / ms / 1e9:.1f} TFLOP/s fp32); finite={bool(torch.isfinite(emb).all())}")
if "--no-cpu" not in sys.argv:
    torch.set_num_threads(torch.get_num_threads())
    t0 = time.perf_counter()
    emb_ref, out_ref = osp.speaker_embedding(wav, 44100, sd, lda)
    dt = time.perf_counter() - t0
    err = float((emb.cpu() - emb_ref).abs().max()) / float(emb_ref.abs().max())
    print(f"CPU oracle (torch fp32, {torch.get_num_threads()} threads): {dt * 1e3:.0f} ms -> {dt * 1e3 / ms:.1f}x; "
          f"embedding rel. max error vs the oracle {err:.2e}")
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    enc(wav, 44100)
    torch.cuda.synchronize()
agg = {}
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        k = e.name.split("(")[0][:60]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += e.device_time if hasattr(e, "device_time") else e.cuda_time
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:8]:
    print(f"  {k:62s} {c:5d} launches {t / 1e3:9.2f} ms")
