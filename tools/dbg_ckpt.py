import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import puflow_oracle as po
from tests import parity
from puflow_b200.engine import PUFlowEngine
dev = "cuda:0"
for tag in ("pu1k", "pugan", "pugeo"):
    sd = torch.load(f"tests/golden/puflow-x4-{tag}.pt", map_location="cpu")
    g = dict(np.load(f"tests/golden/net_{tag}.npz"))
    xyz = torch.from_numpy(g["xyz"])
    o64 = po.PUFlowOracle(sd, dtype=torch.float64).forward(xyz, 4, return_all=True)
    def nerr(a, b):
        a = np.asarray(a, dtype=np.float64).reshape(3, -1); b = np.asarray(b, dtype=np.float64).reshape(3, -1)
        return float((np.abs(a - b).max(1) / np.abs(b).max(1)).max())
    print(tag, "reference fp32 vs oracle fp64:  z %.2e  x %.2e" % (nerr(g["z"], o64["z"].numpy()), nerr(g["x"], o64["x"].numpy())))
    for mode in ("tc", "ffma"):
        r = PUFlowEngine(sd, dev, kernels=mode).forward(xyz.to(dev), 4, return_all=True)
        fz = r["zup"].cpu().view(3, 256, 4, 3).permute(0, 1, 3, 2).numpy()
        print(f"  {mode:5s} vs reference fp32: feat5 %.2e cs5 %.2e z %.2e fz %.2e x %.2e | vs fp64: z %.2e x %.2e" % (
            nerr(r["feats"][5].cpu().numpy(), g["feat5"]), nerr(r["cs"][5].cpu().numpy(), g["cs5"]),
            nerr(r["z"].cpu().numpy(), g["z"]), nerr(fz, g["fz"]) if "fz" in g else -1, nerr(r["x"].cpu().numpy(), g["x"]),
            nerr(r["z"].cpu().numpy(), o64["z"].numpy()), nerr(r["x"].cpu().numpy(), o64["x"].numpy())))
