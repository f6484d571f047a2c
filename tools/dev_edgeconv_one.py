"""Dev tool: run pf_edgeconv_tc for ONE block shape at config-2 size a few times (target of `ncu -k regex:edgeconv_tc`)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200 import ops
from puflow_b200.engine import PUFlowEngine
from puflow_b200.packing import FEAT_CFG
blk = int(sys.argv[1]) if len(sys.argv) > 1 else 3
K = int(sys.argv[2]) if len(sys.argv) > 2 else 16
dev = "cuda:0"
sd = torch.load("tests/golden/puflow-x4-pu1k.pt", map_location="cpu")
E = PUFlowEngine(sd, dev)
Bp, N = 4096, 256
xyz = torch.randn(Bp, N, 3, device=dev) * 0.3
_, idx = ops.knn_topk(xyz, xyz, K, return_dist=False)
C, g, Co = FEAT_CFG[blk]
f = E.w["feat"][blk]
pq = torch.randn(Bp * N, 2 * (4 * g + Co), device=dev)
split = ops.split_tiles(Bp * N, Co, dev)
for _ in range(4):
    ops.edgeconv_tc(pq, idx, f["wtc"], Bp, N, g, Co, split=split)
torch.cuda.synchronize()
print("ok")
