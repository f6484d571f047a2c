"""Dev tool: latency of one 2,048-pt cloud (BASELINE config 1 shape: 32 patches) eager vs CUDA-graph replay."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tools import synth
from puflow_b200.engine import PUFlowEngine
sd = torch.load("tests/golden/puflow-x4-pu1k.pt", map_location="cpu")
eng = PUFlowEngine(sd, "cuda:0")
for Bp in (32, 78, 256, 4096):
    xyz = torch.from_numpy(synth.make_patches(min(Bp, 256), 256, seed=1, shape="surface")).cuda().repeat(max(1, Bp // 256), 1, 1).contiguous()
    for name, fn in (("eager", eng.forward), ("graph", eng.forward_graphed)):
        for _ in range(3):
            fn(xyz, 4)
        torch.cuda.synchronize()
        n = 20 if Bp <= 256 else 5
        t0 = time.perf_counter()
        for _ in range(n):
            fn(xyz, 4)
        torch.cuda.synchronize()
        print(f"Bp={Bp:5d} {name}: {(time.perf_counter() - t0) / n * 1e3:8.3f} ms per call", flush=True)
