"""Dev tool: distribution of the end-to-end z / x differences between the kernel path and the float64 oracle path (2 patches)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import cnf_oracle as co
from puflow_b200.engine import CNFEngine
from tools import synth
sd = torch.load("tests/golden/puflow-x4-cnf-pu1k.pt", map_location="cpu")
E = CNFEngine(sd, "cuda:0")
for seed in (9, 3):
    xyz = torch.from_numpy(synth.make_patches(2, 256, seed=seed, shape="surface"))
    r = E.forward(xyz.cuda(), 4, return_all=True)
    ref = co.CNFOracle(sd).forward(xyz, 4)
    for k in ("z", "x"):
        e = np.abs(r[k].cpu().double().numpy() - ref[k]).max(-1).reshape(-1) / np.abs(ref[k]).max()
        print(f"seed {seed} {k}: median {np.median(e):.2e} q95 {np.quantile(e, 0.95):.2e} q99 {np.quantile(e, 0.99):.2e} max {e.max():.2e}")
