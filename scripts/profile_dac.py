import sys, torch
sys.path.insert(0, ".")
from oracle import shapes
from zonos_b200.autoencoder import DACAutoencoder
B, T = int(sys.argv[1]), int(sys.argv[2])
ae = DACAutoencoder().to("cuda"); ae.load_dac_state_dict(shapes.synth_dac_weights(seed=7))
codes = torch.randint(0, 1024, (B, 9, T), device="cuda")
ae.decode(codes); torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
ae.decode(codes); torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("done")
