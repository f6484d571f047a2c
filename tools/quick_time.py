"""Per-kernel CUDA-event timing of the hot path on synthetic config-2-shaped input (dev tool)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tools import synth
from puflow_b200.engine import PUFlowEngine
from puflow_b200 import ops

Bp = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
sd = torch.load("tests/golden/puflow-x4-pu1k.pt", map_location="cpu")
eng = PUFlowEngine(sd, "cuda:0")
xyz = torch.from_numpy(synth.make_patches(Bp, 256, seed=1, shape="surface")).cuda()
for _ in range(2):
    eng.forward(xyz, 4)
torch.cuda.synchronize()

def timed(name, fn, n=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        out = fn()
    e1.record(); torch.cuda.synchronize()
    print(f"{name:28s} {e0.elapsed_time(e1)/n:9.3f} ms")
    return out

idx = timed("knn16", lambda: eng.knn(xyz, 16))
c = xyz.reshape(-1, 3)
cs = []
from puflow_b200.packing import FEAT_CFG
for i in range(6):
    f = eng.w["feat"][i]; _, g, Co = FEAT_CFG[i]
    pq = timed(f"linear pq{i}", lambda: ops.linear(c, f["wpq"], f["bpq"]))
    c = timed(f"edgeconv{i}", lambda: ops.edgeconv(pq, idx, f["wd"], f["wo"], Bp, 256, g, Co, 0.05))
    cs.append(timed(f"merge{i}", lambda: eng.merge(i, c)).view(Bp, 256, -1))
cond = timed("conditioning", lambda: eng.conditioning(cs))
z, _ = timed("flow_fwd", lambda: eng.f(xyz, cond))
zup = timed("interp", lambda: eng.interp(z, xyz, 4, idx))
timed("flow_inv", lambda: eng.g(zup, cond, 4))
timed("forward total", lambda: eng.forward(xyz, 4))
print("pts/s", Bp * 1024 / 1.0)
