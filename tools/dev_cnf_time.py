"""Dev tool: wall time of the continuous (CNF) variant's forward pass (conditioner + 6 forward + 6 inverse CNF blocks)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200 import ops
from puflow_b200.engine import CNFEngine
from tools import synth
sd = torch.load("tests/golden/puflow-x4-cnf-pu1k.pt", map_location="cpu")
E = CNFEngine(sd, "cuda:0")
for Bp in (32, 1024):
    xyz = torch.from_numpy(synth.make_patches(Bp, 256, seed=1, shape="surface")).cuda()
    E.forward(xyz, 4); torch.cuda.synchronize()
    dt = 1e9
    for _ in range(3):
        t = time.time(); r = E.forward(xyz, 4, return_all=True); torch.cuda.synchronize(); dt = min(dt, time.time() - t)
    print(f"CNF forward {Bp} patches x 256 -> x4: {dt*1e3:.1f} ms  ({Bp*1024/dt/1e6:.2f} M generated points/s)")
ctxs = E.contexts(r["cs"])
for i in (0, 5):
    b = E.w["cnf"][i]
    _, nfe = ops.cnf_flow(xyz.reshape(-1, 3), ctxs[i], b["blob"], 0.0, b["T"], return_nfe=True)
    print(f"block {i} forward: evaluations per point min {int(nfe.min())} mean {float(nfe.float().mean()):.1f} max {int(nfe.max())}")
