"""Dev tool: one forward pass of the continuous (CNF) variant on 64 surface patches (the ncu target for cnf_flow_kernel)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200.engine import CNFEngine
from tools import synth
sd = torch.load("tests/golden/puflow-x4-cnf-pu1k.pt", map_location="cpu")
E = CNFEngine(sd, "cuda:0")
xyz = torch.from_numpy(synth.make_patches(int(os.environ.get("PF_PATCHES", "256")), 256, seed=1, shape="surface")).cuda()
E.forward(xyz, 4)
torch.cuda.synchronize()
print("ok")
