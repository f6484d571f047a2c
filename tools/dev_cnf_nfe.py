"""Dev tool: function evaluations per point of the inverse CNF blocks, and what lock-step lanes cost (warp / CTA maxima)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200 import ops
from puflow_b200.engine import CNFEngine
from tools import synth
sd = torch.load("tests/golden/puflow-x4-cnf-pu1k.pt", map_location="cpu")
E = CNFEngine(sd, "cuda:0")
xyz = torch.from_numpy(synth.make_patches(256, 256, seed=1, shape="surface")).cuda()
r = E.forward(xyz, 4, return_all=True)
ctxs = E.contexts(r["cs"])
y = r["zup"].reshape(-1, 3).contiguous()
tot = totw = totc = 0.0
for i in reversed(range(len(E.w["cnf"]))):
    b = E.w["cnf"][i]
    y, nfe = ops.cnf_flow(y, ctxs[i], b["blob"], b["T"], 0.0, 4, E.atol, E.rtol, return_nfe=True)
    n = nfe.float()
    w = n.view(-1, 32).max(1).values
    c = n.view(-1, 128).max(1).values
    print(f"inverse block {i}: nfe mean {n.mean():.1f} max {n.max():.0f}; mean of warp maxima {w.mean():.1f}; of CTA maxima {c.mean():.1f}")
    tot += n.mean().item(); totw += w.mean().item(); totc += c.mean().item()
print(f"all blocks: lanes useful {tot / totw:.3f} of the warp time; warp-max / CTA-max {totw / totc:.3f}")
