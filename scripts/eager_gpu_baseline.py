import sys, json, torch
sys.path.insert(0, ".")
import bench
class A: shape="real"
dev=torch.device("cuda",0)
model, shape = bench.build_model(A, dev)
r = {}
r["c3"] = bench.eager_gpu_reference(model, "real", 64, 128, 9, bench.C3_SAMPLING, 24, dev)
torch.cuda.empty_cache()
r["c2"] = bench.eager_gpu_reference(model, "real", 1, 23, 0, dict(temperature=0.0, repetition_penalty=1.0), 48, dev)
print(json.dumps(r))
