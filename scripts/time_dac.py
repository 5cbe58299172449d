"""Times DACAutoencoder.decode (CUDA events, after warm-up) at a given batch / frame count."""
import sys, time, torch
sys.path.insert(0, ".")
from oracle import shapes
from zonos_b200.autoencoder import DACAutoencoder
B, T = int(sys.argv[1]), int(sys.argv[2])
ae = DACAutoencoder().to("cuda"); ae.load_dac_state_dict(shapes.synth_dac_weights(seed=7))
codes = torch.randint(0, 1024, (B, 9, T), device="cuda")
for _ in range(2): w = ae.decode(codes)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 3
e0.record()
for _ in range(n): w = ae.decode(codes)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
print(f"dac decode B={B} T={T}: {ms:.2f} ms  -> {B*T*512/44100/(ms/1e3):.0f} audio-s/s, finite={bool(torch.isfinite(w).all())}, rms={float(w.pow(2).mean().sqrt()):.4f}")
