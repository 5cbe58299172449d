#!/usr/bin/env python
"""Benchmark of the Zonos batched speech-generation hot path on B200 (contract: see DESIGN.md, "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...        (one rank per GPU, batch-sharded)

A "step" is one Zonos.generate() call over one batch of synthetic utterances (BASELINE.json configs[2] / SURVEY 8(d)
C3: batch 64 per GPU, L_cond 128, 9 prefix frames, 2580 new frames = 30 s, min-p 0.1 + repetition penalty 3.0,
cfg 2.0, random-init Zonos-v0.1-transformer weights).  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
FRAME_RATE = 44100 / 512  # 86.1328 frames per audio second
# attn_decode_kernel DRAM traffic per launch measured by ncu: key (sequences incl. CFG, kv heads, context)
NCU_ATTN_TRAFFIC = {(128, 4, 1432, 1): (390_396_160, "profiles/r02d_ncu_full_decode_b64.txt"),
                    (128, 4, 1432, 2): (390_500_000, "profiles/r01b_ncu_full_decode_b64.txt")}
C3_SAMPLING = dict(top_p=0, top_k=0, min_p=0.1, linear=0, conf=0, quad=0, repetition_penalty=3.0,
                   repetition_penalty_window=2, temperature=1.0)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="utterances per GPU")
    ap.add_argument("--max-new-tokens", type=int, default=2580)
    ap.add_argument("--l-cond", type=int, default=128)
    ap.add_argument("--prefix-frames", type=int, default=9)
    ap.add_argument("--shape", default="real", choices=["real", "tiny"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--no-dac", action="store_true")
    ap.add_argument("--no-mamba", action="store_true", help="skip the Mamba2 state-update micro-section (hybrid, first slice)")
    ap.add_argument("--cpu-frames", type=int, default=32, help="decode frames in the CPU baseline sample")
    ap.add_argument("--no-eager", action="store_true", help="skip the eager-PyTorch-on-GPU reference baseline")
    ap.add_argument("--no-extras", action="store_true", help="headline numbers only (no groups / prefill / latency / DAC / hybrid / baselines)")
    return ap.parse_args()


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower().startswith("active") for r in self.rows)]
        busy = [s for s in sm if s > 0.5 * max(sm)] if sm else []
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def dist_setup(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        backend = "nccl" if (args.impl == "b200" and torch.cuda.is_available()) else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    elif torch.cuda.is_available():
        torch.cuda.set_device(0)
    return world, rank, local


def barrier(world):
    if world > 1:
        import torch.distributed as dist

        dist.barrier()


def max_over_ranks(x: float, world: int, device) -> float:
    if world == 1:
        return x
    import torch.distributed as dist

    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(x: float, world: int, device) -> float:
    if world == 1:
        return x
    import torch.distributed as dist

    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


# ---------------------------------------------------------------------------------------------------- B200 arm
def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def build_model(args, device):
    from zonos_b200 import synth
    from zonos_b200.config import ZonosConfig
    from zonos_b200.model import Zonos

    shape = synth.REAL if args.shape == "real" else synth.TINY
    cfg = ZonosConfig.from_dict(synth.config_dict(shape))
    sd = synth.transformer_state_dict(shape, device=device, seed=0)
    model = Zonos.from_state_dict(cfg, sd, device)
    del sd
    model.packed()
    return model, shape


def step_bytes(shape, batch, ctx):
    """Algorithmic HBM bytes of one decode step at context `ctx`."""
    d, L, H, Hkv, F = shape["d_model"], shape["n_layer"], shape["n_heads"], shape["n_kv_heads"], shape["d_ff"]
    per_layer = ((H + 2 * Hkv) * 128 * d + H * 128 * d + 2 * F * d + F * d + 4 * d) * 2
    weights = L * per_layer + 2 * d * 2 + 9 * 1026 * d * 2
    kv_per_pos = L * 2 * Hkv * 128 * 2  # bytes per cached position per sequence
    b2 = 2 * batch
    return weights + b2 * kv_per_pos * ctx + b2 * kv_per_pos + batch * 9 * d * 2 + batch * 9 * 1026 * 4


def run_generate(model, cond, prefix, args, rank, seed):
    return model.generate(cond, prefix, max_new_tokens=args.max_new_tokens, cfg_scale=2.0, batch_size=args.batch,
                          sampling_params=dict(C3_SAMPLING), progress_bar=False, seed=seed, row_base=rank * args.batch)


def attn_roofline(shape, batch, ctx, device, hbm_peak):
    """Dominant kernel of the batch-64 decode step: paged GQA decode attention (KV bytes are ~75 % of the step).
    Timed alone with CUDA events on the launching stream over a pool larger than L2, at the mean context."""
    from zonos_b200 import ops

    H, Hkv = shape["n_heads"], shape["n_kv_heads"]
    n_seq, layers = 2 * batch, 4
    kv = ops.PagedKV(layers, n_seq, ctx + 64, Hkv, device)
    kv.pool.normal_()
    q = torch.randn(n_seq, H * 128, device=device).bfloat16()
    import math

    nsplit = max(1, min(8, math.ceil(256 / (n_seq * Hkv)), kv.pages_per_seq))  # the engine's choice (engine.py)
    for l in range(layers):
        ops.attn_decode(q, kv, l, ctx, H, nsplit)
    torch.cuda.synchronize()
    reps = 6
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        for l in range(layers):
            ops.attn_decode(q, kv, l, ctx, H, nsplit)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / (reps * layers)
    algo = n_seq * Hkv * ctx * 128 * 2 * 2 + 2 * n_seq * H * 128 * 2  # K+V rows read once, q in, out
    achieved = algo / (ms * 1e-3) / 1e9
    # DRAM bytes of one launch from the committed `ncu --set full` capture (dram__bytes_read.sum + dram__bytes_write.sum);
    # only valid for the configuration that was captured, otherwise null.
    traffic, traffic_src = NCU_ATTN_TRAFFIC.get((n_seq, Hkv, ctx, nsplit), (None, None))
    return {"kernel": "attn_decode_kernel", "bound": "hbm", "achieved": round(achieved, 1), "peak": hbm_peak, "unit": "GB/s",
            "frac": round(achieved / hbm_peak, 4), "traffic": traffic,
            "traffic_source": traffic_src,
            "bytes_per_launch": algo, "ms_per_launch": round(ms, 5), "ctx": ctx, "nsplit": nsplit}


def layer_param_count(shape):
    d, H, Hkv, F = shape["d_model"], shape["n_heads"], shape["n_kv_heads"], shape["d_ff"]
    return (H + 2 * Hkv) * 128 * d + H * 128 * d + 2 * F * d + F * d


def step_groups(model, shape, B, ctx, device, hbm_peak, ms_per_frame):
    """Where one decode frame goes, by kernel group.  A CUPTI pass (torch.profiler) over four graph replays at the mean
    context gives every kernel's start / end; per group: launches, summed durations, and the EXPOSED time (the part not
    overlapped by earlier kernels of the PDL chain -- exposed times add up to the frame).  Numbers taken under a profiler
    are not bench values: only the SHARES are used, scaled to the event-timed frame (`ms_per_frame`); the committed ncu
    launch list (profiles/) is the cross-reference.  frac = the group's algorithmic HBM bytes / its exposed time / peak."""
    from torch.profiler import ProfilerActivity, profile
    from zonos_b200 import synth
    from zonos_b200.sampler import ST_POS

    L, d, H, Hkv, F = shape["n_layer"], shape["d_model"], shape["n_heads"], shape["n_kv_heads"], shape["d_ff"]
    rows = 2 * B
    cond = synth.conditioning(B, 24, d, device=device)
    side = torch.cuda.Stream(device=device)
    side.wait_stream(torch.cuda.current_stream(device))
    with torch.cuda.stream(side):
        sampling = dict(C3_SAMPLING) if B > 1 else dict(temperature=0.0, repetition_penalty=1.0)
        loop, eng, delayed = model._open_session(cond, None, 2580, 2.0, B, sampling, 1, 0, None)
        try:
            loop.sample_first(eng.logits)
            a = loop.args(eng.logits, True)
            eng.step(a)
            loop.state[ST_POS] = ctx
            eng.capture(a)
            eng.run(8)
            side.synchronize()
            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                eng.run(4)
                side.synchronize()
            n_launch = eng.launches_per_step()
        finally:
            eng.close()
    torch.cuda.current_stream(device).wait_stream(side)
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "memcpy" not in e.name.lower()
           and "memset" not in e.name.lower()]
    evs.sort(key=lambda e: e.time_range.start)
    if len(evs) < 4 * n_launch:
        return {"error": f"profiler saw {len(evs)} kernels, expected {4 * n_launch}"}
    frame = evs[2 * n_launch:3 * n_launch]
    span = frame[-1].time_range.end - frame[0].time_range.start

    def group_of(name):
        n = name.lower()
        if "attn" in n:
            return "attention"
        if "gemm" in n or "chain" in n:
            return "projections"
        if "sampler" in n:
            return "sampler"
        return "reductions"

    kv_bytes = rows * Hkv * ctx * 128 * 2 * 2 * L
    w_bytes = (layer_param_count(shape) * L + 9 * 1026 * d) * 2
    red_bytes = L * (2 * 3 * rows * d * 2 + rows * (H + 2 * Hkv) * 128 * 2) + B * 9 * d * 2
    smp_bytes = B * 9 * (1026 * 4 + 2 * 8 + 8)
    algo = {"attention": kv_bytes, "projections": w_bytes, "reductions": red_bytes, "sampler": smp_bytes}
    agg = {k: [0, 0.0, 0.0] for k in algo}
    cursor = frame[0].time_range.start
    for e in frame:
        s0, s1 = e.time_range.start, e.time_range.end
        g = agg[group_of(e.name)]
        g[0] += 1
        g[1] += s1 - s0
        g[2] += max(0.0, s1 - max(s0, cursor))
        cursor = max(cursor, s1)
    out = {}
    for k, (n, dur, exposed) in agg.items():
        us = exposed / span * ms_per_frame * 1e3
        out[k] = {"launches": n, "us_per_frame": round(us, 1), "share": round(exposed / span, 4),
                  "kernel_us_sum": round(dur, 1), "algorithmic_bytes": int(algo[k]),
                  "frac": round(algo[k] / (us * 1e-6) / 1e9 / hbm_peak, 4) if us > 0 else None}
    out["_note"] = ("shares from one CUPTI pass over graph replays at ctx %d (exposed = not overlapped by earlier kernels of the "
                    "PDL chain), scaled to the event-timed frame; reductions / sampler sit at the launch-latency floor" % ctx)
    return out


def prefill_section(model, shape, B, Lc, P, device):
    """SURVEY 8(d) 'Algorithmic flops, prefill': the C3 prefill (2B x (L_cond + P + 1) rows through 26 layers, causal
    attention, heads on the last row) timed alone with CUDA events on its stream; tensor-pipe bound."""
    from zonos_b200 import synth

    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        burst, sustained = float(pk["bf16_tflops"]), float(pk["bf16_tflops_sustained"])
    except Exception:
        burst = sustained = None
    d, L, H = shape["d_model"], shape["n_layer"], shape["n_heads"]
    cond = synth.conditioning(B, Lc, d, device=device, seed=1)
    g = torch.Generator(device=device).manual_seed(2)
    prefix = torch.randint(0, 1024, (B, 9, P), generator=g, device=device) if P > 0 else None
    S, rows = Lc + P + 1, 2 * B * (Lc + P + 1)
    side = torch.cuda.Stream(device=device)
    side.wait_stream(torch.cuda.current_stream(device))
    with torch.cuda.stream(side):
        loop, eng, delayed = model._open_session(cond, prefix, 16, 2.0, B, dict(C3_SAMPLING), 1, 0, None)
        try:
            launches = eng.launches_per_step()
            x0 = eng.x[:rows].clone()  # the layers update the residual stream in place: restore it before every repetition
            reps, ms = 5, []
            for _ in range(reps + 1):
                eng.x[:rows].copy_(x0)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(side)
                eng.prefill(S)
                e1.record(side)
                side.synchronize()
                ms.append(e0.elapsed_time(e1))
            finite = bool(torch.isfinite(eng.logits[..., :1025]).all())
        finally:
            eng.close()
    torch.cuda.current_stream(device).wait_stream(side)
    t = statistics.median(ms[1:])
    flops = 2 * rows * layer_param_count(shape) * L + 2 * B * 4 * H * 128 * S * S // 2 * L + 2 * 2 * B * 9 * 1026 * d
    tf = flops / (t * 1e-3) / 1e12
    return {"workload": f"C3 prefill: {rows} rows ({2 * B} sequences x {S} positions), {L} layers, causal attention, heads + CFG",
            "ms": round(t, 2), "tflops": round(tf, 1), "bound": "tensor", "launches": launches,
            "frac_of_burst_bf16_peak": None if burst is None else round(tf / burst, 4),
            "frac_of_sustained_bf16_peak": None if sustained is None else round(tf / sustained, 4), "finite": finite}


def eager_gpu_reference(model, shape_name, B, Lc, P, sampling, frames, device):
    """The like-for-like GPU baseline BASELINE.md promises: the reference's eager PyTorch path (oracle/ port: the same torch
    ops -- F.linear, SDPA over the dense cache, ~25 sampling kernels, host syncs -- bit-exact to the reference on CPU) run on
    THIS GPU with THIS model's weights.  Bounded sample: prefill + `frames` decode frames of the full-length workload (the
    dense worst-case KV cache is allocated for all 2580 frames, like the reference); per-frame times with a device sync
    after every frame (the reference syncs several times per frame anyway).  The port vectorises the reference's
    per-utterance Python loop with its B host syncs per frame (model.py:410-414); nothing else about its host side is tuned."""
    from oracle import generate as og

    from zonos_b200 import synth

    W = dict(model.state_dict())
    d = model.config.backbone.d_model

    cond = synth.conditioning(B, Lc, d, device=device, seed=1)
    stamps = []

    def cb(frame, step, max_steps):
        torch.cuda.synchronize(device)
        stamps.append(time.perf_counter())
        return step < frames

    prefix = None if P == 0 else torch.randint(0, 1024, (B, 9, P), device=device)
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    with torch.inference_mode():  # the oracle follows the device of its inputs (no device-context mode taxing every op)
        og.generate_with_model(W, shape_name, cond, prefix, 2580, 2.0, dict(sampling), callback=cb)
    torch.cuda.synchronize(device)
    dts = sorted(b - a for a, b in zip(stamps[2:], stamps[3:]))  # skip the first frames (cuBLAS / SDPA autotuning)
    p50 = dts[len(dts) // 2]
    return {"ms_per_frame_p50": round(p50 * 1e3, 3), "audio_s_per_s": round(B / p50 / FRAME_RATE, 2),
            "prefill_plus_first_frame_s": round(stamps[0] - t0, 2), "frames": len(dts), "ctx": f"{Lc + P + 1}..{Lc + P + 1 + frames}",
            "kind": "port", "sample": f"batch {B}, L_cond {Lc}, {P} prefix frames, prefill + {frames} eager decode frames on the GPU"}


def dac_section(device, batch, frames, world=1):
    """BASELINE configs[3] (C4): DACAutoencoder.decode of `batch` x `frames` synthetic codes -> fp32 waveforms, device
    resident and end to end (codes from pinned host memory, waveform back to the host).  world > 1: every rank decodes
    its own batch (utterances are independent: no collective), times are the max over ranks, audio is the sum."""
    from zonos_b200 import synth
    from zonos_b200.autoencoder import DACAutoencoder

    try:
        tf_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"])
    except Exception:
        tf_peak = None
    ae = DACAutoencoder(load_pretrained=False).to(device)
    ae.load_dac_state_dict(synth.dac_state_dict(device=device, seed=7))
    g = torch.Generator(device=device).manual_seed(5)
    codes = torch.randint(0, 1024, (batch, 9, frames), generator=g, device=device)
    codes_host = codes.cpu().pin_memory()
    for _ in range(2):
        wav = ae.decode(codes)
    wav_host = torch.empty(wav.shape, dtype=wav.dtype).pin_memory()
    torch.cuda.synchronize()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    reps = 3
    barrier(world)
    e0.record()
    for _ in range(reps):
        wav = ae.decode(codes)
    e1.record()
    for _ in range(reps):
        wav = ae.decode(codes_host.to(device, non_blocking=True))
        wav_host.copy_(wav, non_blocking=True)
    e2.record()
    torch.cuda.synchronize()
    ms = max_over_ranks(e0.elapsed_time(e1) / reps, world, device)
    ms_e2e = max_over_ranks(e1.elapsed_time(e2) / reps, world, device)
    flops = 1608302592 * batch * frames * world  # conv MACs x 2 per code frame (DAC layer table)
    audio_s = world * batch * frames / FRAME_RATE
    if world > 1:
        finite = bool(torch.isfinite(wav_host).all())
        del ae
        return {"workload": f"DAC 44.1 kHz decode, batch {batch} x {frames} frames per GPU, {world} GPUs", "ms": round(ms, 1),
                "audio_s_per_s": round(audio_s / (ms * 1e-3), 1),
                "e2e_audio_s_per_s": round(audio_s / (ms_e2e * 1e-3), 1), "tflops": round(flops / ms / 1e9, 1),
                "bound": "tensor", "frac_of_sustained_bf16_peak":
                    None if tf_peak is None else round(flops / ms / 1e9 / tf_peak / world, 4),
                "launches": 31 * reps * 2, "finite": finite}
    # the public call one level up: host codes -> decode -> loudness / trim / fades -> list of host waveforms
    ae.codes_to_wavs(codes_host)
    t0 = time.perf_counter()
    wavs = ae.codes_to_wavs(codes_host)
    torch.cuda.synchronize()
    c2w_s = time.perf_counter() - t0
    c2w_len = sum(w.shape[1] for w in wavs)
    del ae, wavs
    return {"workload": f"DAC 44.1 kHz decode, batch {batch} x {frames} frames", "ms": round(ms, 1),
            "audio_s_per_s": round(audio_s / (ms * 1e-3), 1), "e2e_audio_s_per_s": round(audio_s / (ms_e2e * 1e-3), 1),
            "tflops": round(flops / ms / 1e9, 1), "bound": "tensor",
            "frac_of_sustained_bf16_peak": None if tf_peak is None else round(flops / ms / 1e9 / tf_peak, 4),
            "launches": 31 * reps * 2, "finite": bool(torch.isfinite(wav_host).all()),
            "codes_to_wavs": {"ms": round(c2w_s * 1e3, 1), "audio_s_per_s": round(audio_s / c2w_s, 1),
                              "kept_samples": int(c2w_len),
                              "note": "DACAutoencoder.codes_to_wavs(host codes): decode + post-processing, "
                                      "waveforms copied to pinned host memory"}}


def speaker_section(device, seconds=10.0):
    """Speaker encoder, ResNet293 architecture with seeded weights: one clip, host signal -> embedding
    on the host through SpeakerEmbeddingLDA.forward, CUDA events; the CPU oracle on this box's cores beside it."""
    from zonos_b200 import synth
    from zonos_b200.speaker_cloning import SpeakerEmbeddingLDA

    sd, lda = synth.speaker_state_dicts(device=device, seed=13)
    enc = SpeakerEmbeddingLDA(str(device), state_dict=sd, lda_state_dict=lda)
    wav = synth.speaker_clip(seconds, 44100, 2)
    for _ in range(2):
        emb, out = enc(wav, 44100)
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 5
    e0.record()
    for _ in range(n):
        emb, out = enc(wav, 44100)
        host = out.cpu()
    e1.record()
    torch.cuda.synchronize(device)
    ms = e0.elapsed_time(e1) / n
    from oracle import speaker as osp  # the checker / CPU baseline of this section

    sd_cpu, lda_cpu = {k: v.cpu() for k, v in sd.items()}, {k: v.cpu() for k, v in lda.items()}
    t0 = time.perf_counter()
    emb_ref, _ = osp.speaker_embedding(wav, 44100, sd_cpu, lda_cpu)
    cpu_ms = (time.perf_counter() - t0) * 1e3
    del enc
    return {"workload": f"SpeakerEmbeddingLDA.forward, ResNet293 architecture (seeded weights), one {seconds:.0f} s stereo clip at 44.1 kHz",
            "ms": round(ms, 2), "launches": 3 + 1 + 3 * 97 + 3 + 6, "dtype": "f32",
            "cpu_oracle_ms": round(cpu_ms, 1), "cpu_threads": torch.get_num_threads(),
            "rel_max_error_vs_cpu_oracle": float((emb.cpu() - emb_ref).abs().max() / emb_ref.abs().max())}


def postproc_section(device, hbm_peak, batch, frames, cpu_utts=2, chunk=None):
    """codes_to_wavs' post-processing (loudness normalisation, trim, fades) over a batch of decoded
    waveforms already in HBM, against the reference's per-utterance CPU loop (oracle/postproc.py, bounded sample).
    Algorithmic bytes: the signal has to be read once for the loudness, once more for the scaled output, and written
    once = 12 B per sample; the kernels move ~6x4 B (chunked IIR in two passes + 75 %-overlapped gating blocks)."""
    from zonos_b200.postproc import WavPostprocessor

    n = 512 * frames
    g = torch.Generator(device=device).manual_seed(3)
    t = torch.arange(n, device=device, dtype=torch.float32) / 44100.0
    wav = 0.2 * torch.sin(2 * math.pi * 180.0 * t).unsqueeze(0) * (0.6 + 0.4 * torch.sin(2 * math.pi * 2.3 * t))
    wav = (wav + 0.03 * torch.randn(batch, n, device=device, generator=g)).contiguous()
    wav[:, :1536] *= 1e-4
    pp = WavPostprocessor(device) if chunk is None else WavPostprocessor(device, chunk=chunk)
    lens = [n] * batch
    for _ in range(2):
        out, bounds, lufs, gain = pp.run(wav, lens)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        out, bounds, lufs, gain = pp.run(wav, lens)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    audio_s = batch * n / 44100.0
    from oracle import postproc as opp  # CPU leg: the reference's loop restated

    host = wav[:cpu_utts].cpu()
    t0 = time.perf_counter()
    refs = [opp.post(host[i : i + 1]) for i in range(cpu_utts)]
    dt = time.perf_counter() - t0
    b0 = bounds[0].tolist()
    ok = b0 == [refs[0]["start"], refs[0]["end"]] and abs(float(lufs[0]) - refs[0]["lufs"]) < 1e-4
    algo = 12 * batch * n
    return {"workload": f"codes_to_wavs post-processing, batch {batch} x {n} samples ({n / 44100.0:.1f} s each)",
            "ms": round(ms, 3), "audio_s_per_s": round(audio_s / (ms * 1e-3), 1), "launches": 6 * reps,
            "algorithmic_bytes": algo, "bound": "hbm",
            "hbm_frac": None if hbm_peak is None else round(algo / (ms * 1e-3) / 1e9 / hbm_peak, 4),
            "lufs_in": round(float(lufs[0]), 3), "matches_cpu_leg": bool(ok),
            "cpu_baseline": {"value": round(cpu_utts * n / 44100.0 / dt, 1), "unit": "audio-s/s", "cores": 1,
                             "kind": "port", "sample": f"{cpu_utts} utterances of {n / 44100.0:.1f} s through "
                                                       f"oracle/postproc.py ({dt:.2f} s)"}}


def mamba2_section(device, hbm_peak, rows=64, n_layers=40):
    """First slice of BASELINE configs[4] (hybrid, batch 32 per GPU = 64 rows with CFG) at the hybrid's mixer shape
    (d_model 2048, d_inner 4096, 64 heads x 64, d_state 128): (1) the SSD state update alone over distinct layer states
    (nothing stays in L2); (2) the decode step of a stack of Mamba2 blocks (fused add + norm, in_proj, conv, SSD, gated
    norm, out_proj), host-sequenced C-ABI kernels replayed from a torch CUDA graph.  Attention layers / engine: not yet."""
    from zonos_b200 import synth
    from zonos_b200.hybrid import Mamba2StackB200

    d = 2048
    base = synth.mamba2_mixer_params(d_model=d, device=device, seed=0, dtype=torch.bfloat16)
    s = base["shape"]
    layers = []
    for _ in range(n_layers):  # distinct tensors per layer: the HBM traffic must be real
        p = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in base.items()}
        p["mixer_norm.weight"] = p.pop("norm.weight")
        p["norm.weight"] = torch.ones(d, dtype=torch.bfloat16, device=device)
        layers.append(p)
    stack = Mamba2StackB200(layers, dict(weight=torch.ones(d, dtype=torch.bfloat16, device=device)), s, device)
    states = stack.allocate_inference_cache(rows)
    x = torch.randn(rows, d, device=device).bfloat16()
    from zonos_b200 import ops

    Di, N, H = s["d_inner"], s["d_state"], s["n_heads"]
    xbc = torch.randn(rows, Di + 2 * N, device=device).bfloat16()
    zx = torch.randn(rows, 2 * Di + 2 * N + H, device=device).bfloat16()
    L0 = stack.layers[0]
    ssd = lambda l: ops.mamba2_ssd_step(states[l]["ssm"], xbc, zx, 2 * Di + 2 * N, L0["A"], L0["D"], L0["dtb"], 1)
    side = torch.cuda.Stream(device=device)
    with torch.cuda.stream(side):
        for l in range(n_layers):
            ssd(l)
        side.synchronize()
        e0, e1, e2, e3 = (torch.cuda.Event(enable_timing=True) for _ in range(4))
        reps = 5
        e0.record(side)
        for _ in range(reps):
            for l in range(n_layers):
                ssd(l)
        e1.record(side)
        for st in states:
            st["ssm"].zero_()
        for _ in range(3):
            out = stack.step(x, states)
        side.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            out = stack.step(x, states)
        for _ in range(5):
            graph.replay()
        n = 30
        e2.record(side)
        for _ in range(n):
            graph.replay()
        e3.record(side)
        side.synchronize()
    us = e0.elapsed_time(e1) / (reps * n_layers) * 1e3
    ms_step = e2.elapsed_time(e3) / n
    st_bytes = 2 * rows * H * s["head_dim"] * N * 2
    w_bytes = (stack.n_in_real * d + d * Di) * 2
    algo = n_layers * (w_bytes + st_bytes)
    res = {"workload": f"hybrid first slice: {n_layers} Mamba2 blocks, {rows} rows (batch 32 x CFG), d_model 2048, state 64 x 64 x 128 bf16",
           "ssd_step": {"kernel": "mamba2_ssd_step_p64_kernel", "us_per_layer": round(us, 2), "bytes_per_layer": st_bytes,
                        "bound": "hbm", "achieved_gbs": round(st_bytes / us / 1e3, 1), "hbm_frac": round(st_bytes / us / 1e3 / hbm_peak, 4)},
           "stack_decode_step": {"ms_per_step": round(ms_step, 3), "us_per_layer": round(ms_step / n_layers * 1e3, 1),
                                 "launches_per_step": 6 * n_layers + 1, "algorithmic_bytes": algo, "bound": "hbm",
                                 "hbm_frac": round(algo / (ms_step * 1e-3) / 1e9 / hbm_peak, 4),
                                 "note": "host-sequenced kernels in a torch CUDA graph; attention layers and the C-side engine are not built yet"},
           "finite": bool(torch.isfinite(out).all())}
    del stack, states, graph
    res["generate_c5_like"] = hybrid_generate_section(device)
    return res


def hybrid_generate_section(device, world=1):
    """BASELINE configs[4]-like: Zonos.generate on a random-init hybrid model of the recalled Zonos-v0.1-hybrid shape,
    batch 32 per GPU, through zonos_b200.hybrid.ZonosHybridB200 (decode step replayed from a torch CUDA graph); the
    marginal cost of a frame.  world > 1: every rank generates its own batch (no collective), times are the max over
    ranks, audio is the sum."""
    from zonos_b200 import synth

    model = synth.hybrid_model(device=device)
    cond = synth.conditioning(32, 128, 2048, device=device, seed=1)

    def gen(n_new):
        torch.cuda.synchronize(device)
        barrier(world)
        t0 = time.perf_counter()
        out = model.generate(cond, None, max_new_tokens=n_new, batch_size=32, sampling_params=C3_SAMPLING, seed=1234)
        torch.cuda.synchronize(device)
        return max_over_ranks(time.perf_counter() - t0, world, device), out

    gen(16)
    t_a, _ = gen(128)
    t_b, out = gen(512)
    per_frame = (t_b - t_a) / (512 - 128)
    frames = sum_over_ranks(float(sum(o.shape[1] for o in out)), world, device)
    del model
    return {"workload": "hybrid Zonos.generate, recalled v0.1-hybrid shape (46 blocks, attention + GatedMLP every 9th), "
                        f"random-init, batch 32 per GPU, {world} GPU(s), L_cond 128, min_p 0.1 + rep. penalty 3.0, cfg 2.0",
            "ms_per_frame": round(per_frame * 1e3, 3), "audio_s_per_s": round(world * 32 / per_frame / FRAME_RATE, 1),
            "e2e_audio_s_per_s_512_frames": round(frames / FRAME_RATE / t_b, 1),
            "note": "host-sequenced C-ABI kernels, decode step replayed from a torch CUDA graph; no C-side engine yet"}


def latency_batch1(model, shape, args, device, hbm_peak):
    """BASELINE configs[1]: batch 1, L_cond 23, greedy, 30 s: per-frame latency of the graph-replayed decode step."""
    from zonos_b200 import synth

    cond = synth.conditioning(1, 23, shape["d_model"], device=device, seed=1)
    n = min(args.max_new_tokens, 2580)
    side = torch.cuda.Stream(device=device)  # stream capture is not allowed on the legacy default stream
    side.wait_stream(torch.cuda.current_stream(device))
    with torch.cuda.stream(side):
        loop, eng, delayed = model._open_session(cond, None, n, 2.0, 1, dict(temperature=0.0, repetition_penalty=1.0), 0, 0, None)
        try:
            loop.sample_first(eng.logits)
            a = loop.args(eng.logits, True)
            eng.step(a)
            eng.capture(a)
            side.synchronize()
            steps = loop.max_steps - 1
            evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
            evs[0].record(side)
            for i in range(steps):
                eng.run(1)
                evs[i + 1].record(side)
            side.synchronize()
            t = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(3, steps))
            launches = eng.launches_per_step()
        finally:
            eng.close()
    torch.cuda.current_stream(device).wait_stream(side)
    q = lambda f: t[min(len(t) - 1, int(f * len(t)))]
    mean_ctx = 23 + 1 + steps / 2
    algo = step_bytes(shape, 1, mean_ctx)
    p50 = q(0.5)
    return {"p50_ms": round(p50, 4), "p90_ms": round(q(0.9), 4), "p99_ms": round(q(0.99), 4), "frames": len(t),
            "launches_per_frame": launches, "algorithmic_bytes_per_frame": int(algo),
            "hbm_frac_at_p50": round(algo / (p50 * 1e-3) / 1e9 / hbm_peak, 4)}


# ---------------------------------------------------------------------------------------------------- CPU arm
def cpu_reference_sample(frames: int, shape_name: str = "real", threads=None, W=None, cond=None):
    """The reference's CPU path restated in oracle/ (the reference itself is Python source that cannot travel to the
    GPU box; oracle/ is pinned to it bit-for-bit by tests/test_oracle_golden.py): BASELINE configs[0] shape -- batch 1,
    L_cond 23, greedy -- on a bounded sample: prefill + `frames` decode frames.  W / cond: use these weights / this
    conditioning (the parity check feeds the GPU model's own).  Returns (audio-s/s, seconds, cores, codes)."""
    from oracle import generate as og
    from oracle import shapes

    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    if W is None:
        W = shapes.synth_transformer_weights(shape_name, seed=0)
    if cond is None:
        cond = shapes.synth_conditioning(1, 23, shapes.SHAPES[shape_name]["d_model"], seed=1)
    t0 = time.perf_counter()
    with torch.inference_mode():
        out = og.generate_with_model(W, shape_name, cond, None, frames, 2.0, dict(temperature=0.0, repetition_penalty=1.0))
    dt = time.perf_counter() - t0
    audio_s = sum(o.shape[1] for o in out) / FRAME_RATE
    return audio_s / dt, dt, cores, out


def cpu_reference_c3(args, n_steps: int, frames_per_step: int, budget_s: float = 170.0):
    """The reference arm at the B200 arm's OWN configuration (C3: batch `args.batch`, L_cond, prefix frames, min-p +
    repetition penalty, cfg 2), oracle port on all host cores.  A full step of that workload (2580 frames at batch 64) is
    ~10 minutes of CPU time, so the sample is bounded like this: the session is set up and PREFILLED once, outside the
    timed steps (prefill is ~1 % of the full workload on either arm; its wall time is reported), and every timed step is
    `frames_per_step` consecutive decode frames of all utterances, at the low contexts right after the prefill (attention is
    cheaper there than at the workload's mean context of ~1430: the sample flatters the CPU).  Stops early when the time
    budget is used up.  Returns dict(value, step_seconds, prefill_s, steps_run, cores, ctx0)."""
    from oracle import generate as og
    from oracle import shapes

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    B, Lc, P = args.batch, args.l_cond, args.prefix_frames
    W = shapes.synth_transformer_weights(args.shape, seed=0)
    cond = shapes.synth_conditioning(B, Lc, shapes.SHAPES[args.shape]["d_model"], seed=1)
    prefix = torch.randint(0, 1024, (B, 9, P), generator=torch.Generator().manual_seed(2)) if P > 0 else None
    want = (n_steps + 1) * frames_per_step  # the first group of frames is the warm-up
    stamps, t_start = [], time.perf_counter()

    def cb(frame, step, max_steps):
        stamps.append(time.perf_counter())
        return len(stamps) < want + 1 and (time.perf_counter() - t_start) < budget_s

    with torch.inference_mode():
        og.generate_with_model(W, args.shape, cond, prefix, want + 8, 2.0, dict(C3_SAMPLING), callback=cb)
    prefill_s = stamps[0] - t_start
    groups = [(stamps[k + frames_per_step] - stamps[k]) for k in range(frames_per_step, len(stamps) - frames_per_step, frames_per_step)]
    if not groups:  # budget ran out during the warm-up group: use whatever frames exist
        groups = [(stamps[-1] - stamps[0]) / max(1, len(stamps) - 1) * frames_per_step]
    audio_per_step = B * frames_per_step / FRAME_RATE
    return dict(value=audio_per_step * len(groups) / sum(groups), step_seconds=groups, prefill_s=prefill_s, steps_run=len(groups),
                cores=cores, ctx0=Lc + P + 1)


def main_reference(args, world, rank):
    """--impl reference: the reference's own CPU implementation of the path on the host cores (rank 0 only), on the B200
    arm's configuration (see cpu_reference_c3 for how the sample is bounded)."""
    if rank != 0:
        return
    fps = 2
    r = cpu_reference_c3(args, max(1, args.steps), fps)
    B, Lc, P = args.batch, args.l_cond, args.prefix_frames
    sample = (f"C3 config (batch {B}, L_cond {Lc}, {P} prefix frames, min_p 0.1 + rep. penalty 3.0, cfg 2.0): prefilled once outside "
              f"the timed steps ({r['prefill_s']:.1f} s), then {r['steps_run']} timed steps of {fps} decode frames of all {B} utterances "
              f"at contexts from {r['ctx0']} (mean context of the full 2580-frame workload: ~1430, where a CPU frame costs more)")
    value = round(r["value"], 5)
    line = {
        "impl": "reference", "metric": "audio-sec/sec", "value": value, "unit": "audio-s/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "steps_run": r["steps_run"],
        "ms_per_step": round(1e3 * sum(r["step_seconds"]) / len(r["step_seconds"]), 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "Zonos.generate C3: Zonos-v0.1-transformer shape, random-init, batch %d/GPU, L_cond %d, "
                               "%d prefix frames, %d new frames, min_p 0.1 + rep. penalty 3.0, cfg 2.0" % (B, Lc, P, args.max_new_tokens),
                   "shape": args.shape, "global_batch": B, "impl": "CPU reference path (oracle port, torch CPU bf16)", "sample": sample},
        "cpu_baseline": {"value": value, "unit": "audio-s/s", "cores": r["cores"], "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def parity_check(model, shape_name, device, frames, cpu_out, cond_cpu):
    """The bench's own sanity line: the GPU engine and the CPU oracle on the SAME weights (the bench model's, copied to the
    host) and conditioning -- greedy batch-1 tokens of the CPU-baseline run it already paid for, plus the prefill logits."""
    from oracle import delay as od
    from oracle import generate as og

    out = model.generate(cond_cpu.to(device), None, max_new_tokens=frames, batch_size=1, progress_bar=False,
                         sampling_params=dict(temperature=0.0, repetition_penalty=1.0))
    a, b = out[0].cpu(), cpu_out[0]
    n = min(a.shape[1], b.shape[1])
    same = (a[:, :n] == b[:, :n]).all(dim=0).long().cumprod(0)
    gl = model._debug_prefill_logits(cond_cpu.to(device), None, 2.0).cpu()
    W = {k: v.cpu() for k, v in model.state_dict().items() if k.startswith(("backbone.", "embeddings.", "heads."))}
    src = og.ModelLogits(W, shape_name, cond_cpu, 64, 2.0)
    with torch.inference_mode():
        rl = src.prefill(od.apply_delay_pattern(torch.full((1, 9, 8), -1, dtype=torch.int64), 1025)[..., :1])
    fin = torch.isfinite(rl)
    dl = gl[fin] - rl[fin]
    return {"greedy_frames_identical_to_cpu_oracle": int(same.sum()), "frames_compared": int(n),
            "prefill_logits_max_abs_diff": round(float(dl.abs().max()), 4),
            "prefill_logits_rel_rms": round(float(dl.pow(2).mean().sqrt() / rl[fin].pow(2).mean().sqrt()), 5),
            "masks_equal": bool(torch.equal(fin, torch.isfinite(gl))),
            "gates": "rel-RMS <= 2.5e-2, max <= 0.40 at 26 layers (tests/test_gpu_model_real.py); greedy tokens may part at a near-tie"}


def main_b200(args, world, rank, local):
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (there is no CPU fallback)")
    device = torch.device("cuda", local if world > 1 else 0)
    hbm_peak, peak_src = peaks()
    model, shape = build_model(args, device)
    from zonos_b200 import synth

    B, Lc, P = args.batch, args.l_cond, args.prefix_frames
    cond = synth.conditioning(B, Lc, shape["d_model"], device=device, seed=1 + rank)
    g = torch.Generator(device=device).manual_seed(2 + rank)
    prefix = torch.randint(0, 1024, (B, 9, P), generator=g, device=device) if P > 0 else None
    cond_host = cond.cpu().pin_memory()
    prefix_host = None if prefix is None else prefix.cpu().pin_memory()

    for i in range(args.warmup):
        run_generate(model, cond, prefix, args, rank, 1234)
    torch.cuda.synchronize()

    clocks = ClockSampler(local if world > 1 else 0)
    if rank == 0:
        clocks.start()
    # ---- device-resident inputs: `value`
    barrier(world)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    frames_out, launches = 0, 0
    e0.record()
    for i in range(args.steps):
        out = run_generate(model, cond, prefix, args, rank, 1234)
        frames_out += sum(o.shape[1] for o in out)
        st = model.last_generate_stats
        # counted, not estimated: the engine reports the kernels of its prefill and of one frame; + delay-pattern apply,
        # prefix embedding, first-frame sampler, delay-pattern revert
        launches += st["prefill_launches"] + st["launches_per_step"] * st["steps"] + 4
    e1.record()
    torch.cuda.synchronize()
    barrier(world)
    ms_dev = max_over_ranks(e0.elapsed_time(e1), world, device)
    total_frames = sum_over_ranks(frames_out, world, device)
    steps_per_gen = model.last_generate_stats["steps"]

    # ---- end to end through the public API with HOST buffers: pinned conditioning in, codes out
    if world > 1:  # warm-up of the result collective (NCCL sets its point-to-point channels up on first use: ~1.5 s at 8 ranks)
        from zonos_b200.parallel import gather_codes

        gather_codes([torch.zeros(9, 2, dtype=torch.int64, device=device) for _ in range(B)], B * world, dst=0)
    barrier(world)
    torch.cuda.synchronize()
    h2d = cond_host.numel() * cond_host.element_size() + (0 if prefix_host is None else prefix_host.numel() * 8)
    d2h, frames_e2e = 0, 0
    e0.record()
    for i in range(args.steps):
        c = cond_host.to(device, non_blocking=True)
        pfx = None if prefix_host is None else prefix_host.to(device, non_blocking=True)
        out = run_generate(model, c, pfx, args, rank, 1234)
        frames_e2e += sum(o.shape[1] for o in out)
        if world > 1:  # the only collective of the path -- results to rank 0 over NCCL (int32 codes)
            from zonos_b200.parallel import gather_codes

            gathered = gather_codes(out, B * world, dst=0)
            d2h = sum(o.numel() * 4 for o in out)
            if rank == 0:
                assert len(gathered) == B * world
        else:
            host_codes = [o.cpu() for o in out]
            d2h = sum(o.numel() * 8 for o in host_codes)
    e1.record()
    torch.cuda.synchronize()
    barrier(world)
    ms_e2e = max_over_ranks(e0.elapsed_time(e1), world, device)
    total_frames_e2e = sum_over_ranks(frames_e2e, world, device)
    clk = clocks.stop() if rank == 0 else None

    if rank != 0:
        del model
        torch.cuda.empty_cache()
        if not args.no_extras and not args.no_dac:  # C4 sharded: every rank decodes its own batch
            dac_section(device, B, args.max_new_tokens, world)
        if not args.no_extras and not args.no_mamba:  # C5-like sharded the same way
            torch.cuda.empty_cache()
            hybrid_generate_section(device, world)
        return
    value = total_frames / FRAME_RATE / (ms_dev * 1e-3)
    e2e = total_frames_e2e / FRAME_RATE / (ms_e2e * 1e-3)
    mean_ctx = Lc + P + 1 + steps_per_gen / 2
    ms_per_frame = ms_dev / args.steps / max(1, steps_per_gen + 1)
    algo_step = step_bytes(shape, B, mean_ctx)
    roof = attn_roofline(shape, B, int(mean_ctx), device, hbm_peak)
    roof["peak_source"] = peak_src
    roof["note"] = ("read-only KV stream; `peak` is the measured copy (read+write) bandwidth, which a read-only kernel can "
                    "exceed slightly")
    line = {
        "metric": "audio-sec/sec", "value": round(value, 2), "unit": "audio-s/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_dev / args.steps, 2), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "Zonos.generate C3: Zonos-v0.1-transformer shape, random-init, batch %d/GPU, L_cond %d, "
                               "%d prefix frames, %d new frames, min_p 0.1 + rep. penalty 3.0, cfg 2.0" % (B, Lc, P, args.max_new_tokens),
                   "shape": args.shape, "global_batch": B * world, "parallelism": (f"batch-sharded x{world}, no collective on the data path; e2e: codes gathered to rank 0 over NCCL "
                                   "(all_gather of lengths + gather of padded int32 codes) inside the timed region") if world > 1
                   else "batch-sharded x1, no collectives",
                   "l2": "inputs larger than L2 (3.2 GB weights + paged KV per step)", "dac_decode": "not in the timed region"},
        "e2e": {"value": round(e2e, 2), "unit": "audio-s/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches,
        "clocks": clk,
        "roofline": roof,
        "decode_step": {"ms_per_frame": round(ms_per_frame, 4), "algorithmic_bytes": int(algo_step),
                        "hbm_frac": round(algo_step / (ms_per_frame * 1e-3) / 1e9 / hbm_peak, 4), "mean_ctx": mean_ctx,
                        "launches_per_frame": model.last_generate_stats["launches_per_step"]},
    }
    step_frac = round(algo_step / (ms_per_frame * 1e-3) / 1e9 / hbm_peak, 4)
    roof["step_frac"] = step_frac  # the WHOLE decode frame against the HBM roofline (the dominant kernel alone is `frac`)
    extras = not args.no_extras
    if extras:
        line["prefill"] = prefill_section(model, shape, B, Lc, P, device)
    if extras and not args.no_latency:
        line["latency_batch1"] = latency_batch1(model, shape, args, device, hbm_peak)
    if extras and world == 1 and not args.no_eager:
        try:
            eager = {"c3_batch%d" % B: eager_gpu_reference(model, args.shape, B, Lc, P, C3_SAMPLING, 24, device)}
            torch.cuda.empty_cache()
            eager["c2_batch1"] = eager_gpu_reference(model, args.shape, 1, 23, 0, dict(temperature=0.0, repetition_penalty=1.0), 48, device)
            torch.cuda.empty_cache()
            eager["vs_eager_gpu"] = {"c3_frame_time_ratio": round(eager["c3_batch%d" % B]["ms_per_frame_p50"] / ms_per_frame, 2)}
            if "latency_batch1" in line:
                eager["vs_eager_gpu"]["c2_frame_time_ratio"] = round(eager["c2_batch1"]["ms_per_frame_p50"] / line["latency_batch1"]["p50_ms"], 2)
            line["reference_gpu_eager"] = eager
        except Exception as e:
            line["reference_gpu_eager"] = {"error": f"{type(e).__name__}: {e}"}
            torch.cuda.empty_cache()
    if extras and world == 1 and not args.no_cpu_baseline:
        # BASELINE configs[0]-like bounded sample on the host cores, on the bench model's OWN weights so that the same run
        # doubles as a parity check of the real-shape engine (the reference arm, --impl reference, runs the C3 config)
        frames = args.cpu_frames
        W_cpu = {k: v.cpu() for k, v in model.state_dict().items() if k.startswith(("backbone.", "embeddings.", "heads."))}
        cond1 = synth.conditioning(1, 23, shape["d_model"], device=device, seed=7).cpu()
        v, dt, cores, cpu_out = cpu_reference_sample(frames, args.shape, W=W_cpu, cond=cond1)
        del W_cpu
        line["cpu_baseline"] = {"value": round(v, 5), "unit": "audio-s/s", "cores": cores, "kind": "port",
                                "sample": f"batch 1, L_cond 23, greedy, prefill + {frames + 8} decode frames on the host "
                                          f"cores ({dt:.1f} s); oracle/ port of the reference's CPU path; the C3-config CPU "
                                          f"number is the reference arm's (--impl reference)"}
        try:
            line["parity_check"] = parity_check(model, args.shape, device, frames, cpu_out, cond1)
        except Exception as e:
            line["parity_check"] = {"error": f"{type(e).__name__}: {e}"}
    if extras and world == 1 and not args.no_dac:
        # text-less pipeline to finished waveforms, MEASURED as one pipeline on the same codes inside one event pair:
        # pinned conditioning -> generate -> codes_to_wavs (DAC decode, loudness, trim, fades) -> waveforms on the host
        try:
            from zonos_b200.autoencoder import DACAutoencoder

            ae = DACAutoencoder(load_pretrained=False).to(device)
            ae.load_dac_state_dict(synth.dac_state_dict(device=device, seed=7))
            warm = run_generate(model, cond, prefix, args, rank, 1234)
            ae.codes_to_wavs(warm)
            torch.cuda.synchronize()
            e0.record()
            c = cond_host.to(device, non_blocking=True)
            pfx = None if prefix_host is None else prefix_host.to(device, non_blocking=True)
            codes = run_generate(model, c, pfx, args, rank, 1234)
            wavs = ae.codes_to_wavs(codes)
            e1.record()
            torch.cuda.synchronize()
            audio = sum(o.shape[1] for o in codes) / FRAME_RATE
            line["e2e_wavs"] = {"value": round(audio / (e0.elapsed_time(e1) * 1e-3), 2), "unit": "audio-s/s", "measured": True,
                                "ms": round(e0.elapsed_time(e1), 1), "kept_samples": int(sum(w.shape[1] for w in wavs)),
                                "note": "ONE pipeline, one event pair: pinned host conditioning -> generate -> codes_to_wavs "
                                        "(DAC decode + loudness + trim + fades) -> waveforms on the host; random-init DAC"}
            del ae, wavs, codes, warm
        except Exception as e:
            line["e2e_wavs"] = {"error": f"{type(e).__name__}: {e}"}
    del model
    torch.cuda.empty_cache()
    if extras and not args.no_dac:
        line["dac_decode"] = dac_section(device, B, args.max_new_tokens, world)
    if extras and world == 1 and not args.no_mamba:
        line["hybrid_mamba2"] = mamba2_section(device, hbm_peak)
        torch.cuda.empty_cache()
    elif extras and not args.no_mamba:
        torch.cuda.empty_cache()
        line["hybrid_mamba2"] = {"generate_c5_like": hybrid_generate_section(device, world)}
    if extras and world == 1 and not args.no_dac:
        line["postprocess"] = postproc_section(device, hbm_peak, B, args.max_new_tokens)
        torch.cuda.empty_cache()
    if extras and world == 1:
        try:
            line["speaker_encoder"] = speaker_section(device)
        except Exception as e:
            line["speaker_encoder"] = {"error": f"{type(e).__name__}: {e}"}
        torch.cuda.empty_cache()
    if extras:
        # LAST, on a fresh model: the CUPTI pass behind `roofline.groups`.  Once torch.profiler has attached CUPTI, every
        # later launch of the process pays for it (measured: the wall-clock-timed hybrid section read 4.46 instead of
        # 3.42 ms per frame when this ran before it), so nothing is measured after this point.
        try:
            model, _ = build_model(args, device)
            roof["groups"] = step_groups(model, shape, B, int(mean_ctx), device, hbm_peak, ms_per_frame)
            del model
        except Exception as e:  # CUPTI unavailable on the box: the headline numbers do not depend on it
            roof["groups"] = {"error": f"{type(e).__name__}: {e}"}
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    world, rank, local = dist_setup(args)
    if args.impl == "reference":
        main_reference(args, world, rank)
    else:
        main_b200(args, world, rank, local)
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
