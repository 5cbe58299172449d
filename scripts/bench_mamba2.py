"""Times the Mamba2 decode-step kernels at the hybrid's shape (BASELINE configs[4]: batch 32 per GPU,
64 rows with CFG): d_model 2048, d_inner 4096, 64 heads x 64, d_state 128.  The state update is HBM-bound: it reads and
writes rows x H x P x N bf16 once per token and layer.  usage: python scripts/bench_mamba2.py [rows] [layers]"""
import sys

import torch

sys.path.insert(0, ".")
import bench
from oracle import mamba2 as om
from zonos_b200 import ops

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 64
layers = int(sys.argv[2]) if len(sys.argv) > 2 else 40  # distinct states, so that nothing stays in the 126 MB L2
p = om.synth_mixer_params(d_model=2048, seed=0, dtype=torch.bfloat16)
s = p["shape"]
Di, H, P, N, G = s["d_inner"], s["n_heads"], s["head_dim"], s["d_state"], s["n_groups"]
dev = torch.device("cuda", 0)
ssm = [(0.1 * torch.randn(rows, H, P, N, device=dev)).bfloat16() for _ in range(layers)]
conv = [torch.randn(rows, Di + 2 * G * N, 4, device=dev).bfloat16() for _ in range(layers)]
zx = torch.randn(rows, 2 * Di + 2 * G * N + H, device=dev).bfloat16()
cw, cb = p["conv1d.weight"].to(dev), p["conv1d.bias"].to(dev)
A, D, dtb, nw = p["A_log"].to(dev), p["D"].to(dev), p["dt_bias"].to(dev), p["norm.weight"].to(dev)
xbc = ops.mamba2_conv_update(conv[0], zx, Di, cw, cb)
y = ops.mamba2_ssd_step(ssm[0], xbc, zx, 2 * Di + 2 * G * N, A, D, dtb, G)


def timed(fn, reps=5):
    for l in range(layers):
        fn(l)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        for l in range(layers):
            fn(l)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * layers) * 1e3  # us per launch


hbm_peak, _ = bench.peaks()
t_ssd = timed(lambda l: ops.mamba2_ssd_step(ssm[l], xbc, zx, 2 * Di + 2 * G * N, A, D, dtb, G))
t_conv = timed(lambda l: ops.mamba2_conv_update(conv[l], zx, Di, cw, cb))
t_norm = timed(lambda l: ops.mamba2_gated_rmsnorm(y, zx, nw, s["eps"], G))
ssd_bytes = 2 * rows * H * P * N * 2
print(f"rows {rows}: ssd_step {t_ssd:.2f} us/launch, {ssd_bytes / 1e6:.1f} MB state traffic -> {ssd_bytes / t_ssd / 1e3:.0f} GB/s = "
      f"{ssd_bytes / t_ssd / 1e3 / hbm_peak:.3f} of the measured HBM peak ({hbm_peak:.0f} GB/s); conv_update {t_conv:.2f} us, "
      f"gated_rmsnorm {t_norm:.2f} us (includes the wrappers' torch allocations)")

# prefill scan at the hybrid's C5-like shape: 64 sequences x 138 tokens
S = 138
zxs = torch.randn(rows * S, 2 * Di + 2 * G * N + H, device=dev).bfloat16()
xbcs = ops.mamba2_conv_seq(conv[0], zxs, S, Di, cw, cb)
t_scan = timed(lambda l: ops.mamba2_ssd_scan(ssm[l], xbcs, zxs, S, 2 * Di + 2 * G * N, A, D, dtb, G), reps=2)
t_cseq = timed(lambda l: ops.mamba2_conv_seq(conv[l], zxs, S, Di, cw, cb), reps=2)
print(f"prefill {rows} x {S} tokens: ssd_scan {t_scan:.1f} us/layer, conv_seq {t_cseq:.1f} us/layer")
