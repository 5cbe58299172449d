"""Tensor-bound pieces on one B200: the C3 prefill (batch 64, S = 138 -> 17,664 rows through 26 layers + heads) and the
C4 DAC decode (64 x 2580 frames), timed with CUDA events; prints algorithmic TFLOP/s against the measured bf16 peak."""
import json
import sys

import torch

sys.path.insert(0, ".")
import bench
from oracle import shapes as oshapes
from zonos_b200 import synth
from zonos_b200.autoencoder import DACAutoencoder

dev = torch.device("cuda", 0)
peak = json.load(open("MEASURED_PEAKS.json"))
out = {}


class A:
    shape = "real"


what = sys.argv[1] if len(sys.argv) > 1 else "both"
if what in ("both", "prefill"):
    model, shape = bench.build_model(A, dev)
    B, Lc, P = 64, 128, 9
    cond = synth.conditioning(B, Lc, shape["d_model"], device=dev, seed=1)
    prefix = torch.randint(0, 1024, (B, 9, P), device=dev)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        loop, eng, delayed = model._open_session(cond, prefix, 2580, 2.0, B, dict(bench.C3_SAMPLING), 0, 0, None)
        S = Lc + P + 1
        side.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        for _ in range(3):
            eng.prefill(S)
        e1.record(side)
        side.synchronize()
        ms = e0.elapsed_time(e1) / 3
        eng.close()
    rows = 2 * B * S
    flops = 2 * rows * 1581252608 + 2 * B * 26 * 4 * 16 * 128 * S * S / 2 * 2  # GEMMs + causal attention (qk and pv)
    out["prefill"] = dict(ms=round(ms, 2), rows=rows, tflops=round(flops / ms / 1e9, 1),
                          frac_of_burst_peak=round(flops / ms / 1e9 / peak["bf16_tflops"], 4))
    print(out["prefill"], flush=True)
    del model, eng, loop
    torch.cuda.empty_cache()
if what in ("both", "dac"):
    ae = DACAutoencoder().to(dev)
    ae.load_dac_state_dict(oshapes.synth_dac_weights(seed=7))
    B, T = 64, 2580
    codes = torch.randint(0, 1024, (B, 9, T), device=dev)
    for _ in range(2):
        wav = ae.decode(codes)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    wav = ae.decode(codes)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    fl = 1608302592 * B * T
    out["dac"] = dict(ms=round(ms, 1), audio_s_per_s=round(B * T / 86.1328 / (ms * 1e-3), 1), tflops=round(fl / ms / 1e9, 1),
                      frac_of_sustained_peak=round(fl / ms / 1e9 / peak["bf16_tflops_sustained"], 4))
    print(out["dac"], flush=True)
json.dump(out, open("gpurun_out/bench_tensor.json", "w"))
