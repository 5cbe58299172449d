import sys, traceback, torch
sys.path.insert(0, ".")
import bench
args = bench.parse()
dev = torch.device("cuda", 0)
model, shape = bench.build_model(args, dev)
from zonos_b200 import synth
for B, n in [(64, 600), (64, 2580), (8, 2580), (64, 2580)]:
    args.batch, args.max_new_tokens = B, n
    cond = synth.conditioning(B, 128, shape["d_model"], device=dev, seed=1)
    try:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = bench.run_generate(model, cond, None, args, 0, 1)
        e1.record(); torch.cuda.synchronize()
        print("OK", B, n, model.last_generate_stats, "ms", e0.elapsed_time(e1), "mem GB", torch.cuda.max_memory_allocated() / 1e9, flush=True)
    except Exception:
        print("FAIL", B, n, flush=True)
        traceback.print_exc()
        torch.cuda.synchronize()
