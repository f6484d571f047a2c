"""Which tensor-core stage after the features costs accuracy in z?  Features from the FFMA kernels throughout."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import puflow_oracle as po
from puflow_b200.engine import PUFlowEngine, NUM_BLOCKS
from puflow_b200 import ops
dev = "cuda:0"
def nerr(a, b):
    a = np.asarray(a.cpu() if torch.is_tensor(a) else a, dtype=np.float64).reshape(3, -1); b = np.asarray(b, dtype=np.float64).reshape(3, -1)
    return float((np.abs(a - b).max(1) / np.abs(b).max(1)).max())
for tag in ("pu1k", "pugan"):
    sd = torch.load(f"tests/golden/puflow-x4-{tag}.pt", map_location="cpu")
    g = dict(np.load(f"tests/golden/net_{tag}.npz"))
    xyz = torch.from_numpy(g["xyz"]).to(dev)
    o64 = po.PUFlowOracle(sd, dtype=torch.float64).forward(xyz.cpu(), 4, return_all=True)
    E = PUFlowEngine(sd, dev)
    Bp, N, _ = xyz.shape; M = Bp * N
    idx = E.knn(xyz)
    E.use_tc = False
    cs_f, feats = E.feat_extract(xyz, idx, return_feats=True)
    for merge_tc in (0, 1):
      for raw_tc in (0, 1):
        for cond_tc in (0, 1):
          for fwd_tc in (0, 1):
            E.use_tc = bool(merge_tc)
            cs = [E.merge(i, feats[i].reshape(M, -1)).view(Bp, N, -1) for i in range(NUM_BLOCKS)]
            raw = torch.empty((NUM_BLOCKS, M, 192), dtype=torch.float32, device=dev)
            for i in range(NUM_BLOCKS):
                if raw_tc and E.w["wc_tc"][i] is not None:
                    ops.linear_tc(cs[i].reshape(M, -1), E.w["wc_tc"][i], 192, None, 0, out=raw[i])
                else:
                    ops.linear(cs[i].reshape(M, -1), E.w["wc"][i], None, 0, out=raw[i])
            sb = ops.flow_cond(raw, E.w["cond_tc"] if cond_tc else E.w["cond"])
            z, ssum = ops.flow_fwd(xyz.reshape(M, 3), raw, sb, E.w["flow_tc"] if fwd_tc else E.w["flow"])
            print(tag, f"merge_tc={merge_tc} raw_tc={raw_tc} cond_tc={cond_tc} fwd_tc={fwd_tc}  cs5 %.2e raw5 %.2e z %.2e" % (
                nerr(cs[5], o64["cs"][5].numpy()), 0.0, nerr(z.view(Bp, N, 3), o64["z"].numpy())))
