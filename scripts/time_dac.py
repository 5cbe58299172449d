import sys, time, torch
sys.path.insert(0, ".")
from oracle import shapes
from zonos_b200.autoencoder import DACAutoencoder
B, T = int(sys.argv[1]), int(sys.argv[2])
ae = DACAutoencoder().to("cuda"); ae.load_dac_state_dict(shapes.synth_dac_weights(seed=7))
codes = torch.randint(0, 1024, (B, 9, T), device="cuda")
for _ in range(2):
    wav = ae.decode(codes)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); wav = ae.decode(codes); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
fl = 1608302592 * B * T
print(f"B={B} T={T}: {ms:.1f} ms, {B*T/86.1328/(ms*1e-3):.0f} audio-s/s, {fl/ms/1e9:.1f} TFLOP/s algorithmic, mem {torch.cuda.max_memory_allocated()/1e9:.1f} GB")
