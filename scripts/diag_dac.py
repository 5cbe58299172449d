import sys, torch
sys.path.insert(0, ".")
from oracle import dac as odac, shapes
from zonos_b200.autoencoder import DACAutoencoder
rr = lambda a, b: float((a - b).pow(2).mean().sqrt() / b.pow(2).mean().sqrt())
for name, arch, cases in (("tiny", shapes.DAC_TINY, [(1, 1), (3, 7), (2, 86), (2, 300)]), ("real", shapes.DAC, [(1, 1), (1, 40)])):
    W = shapes.synth_dac_weights(seed=7, arch=arch)
    ae = DACAutoencoder().to("cuda"); ae.load_dac_state_dict(W)
    g = torch.Generator().manual_seed(9)
    for B, T in cases:
        codes = torch.randint(0, 1024, (B, 9, T), generator=g)
        ref32 = odac.decode(W, codes); ref16 = odac.decode_fp16_storage(W, codes)
        got = ae.decode(codes.cuda()).cpu()
        d = (got - ref16).abs()[0, 0]
        n = d.numel()
        print(f"{name} B{B} T{T}: gpu-vs-fp16model {rr(got, ref16):.2e}  gpu-vs-fp32 {rr(got, ref32):.2e}  fp16model-vs-fp32 {rr(ref16, ref32):.2e}"
              f"  | err head/mid/tail {d[:n//8].mean():.1e} {d[n//4:3*n//4].mean():.1e} {d[-n//8:].mean():.1e} max {d.max():.1e}", flush=True)
