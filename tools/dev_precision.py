"""Dev tool: measured error of precision="fp16" (single-pass) vs "fp32" (split) against the reference fixtures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from puflow_b200.engine import PUFlowEngine
G = "tests/golden"
def nw(a, b):
    a, b = a.reshape(a.shape[0], -1), b.reshape(b.shape[0], -1)
    return float((np.abs(a - b).max(1) / np.abs(b).max(1)).max())
for tag in ("pu1k", "pugan", "pugeo"):
    e = np.load(f"{G}/net_{tag}.npz")
    sd = torch.load(f"{G}/puflow-x4-{tag}.pt", map_location="cpu")
    xyz = torch.from_numpy(e["xyz"]).cuda()
    for prec in ("fp32", "fp16"):
        r = PUFlowEngine(sd, "cuda:0", precision=prec).forward(xyz, 4, return_all=True)
        print(tag, prec, "x %.2e  z %.2e  feat5 %.2e" % (nw(r["x"].cpu().numpy(), e["x"]), nw(r["z"].cpu().numpy(), e["z"]),
              nw(r["feats"][5].cpu().numpy(), e["feat5"])))
