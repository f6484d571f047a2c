import os, sys
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from oracle import cnf_oracle as co
from puflow_b200 import ops, packing
sd = torch.load("tests/golden/puflow-x4-cnf-pu1k.pt", map_location="cpu")
for blk, cdim in ((5,128),(3,128)):
    pk = packing.pack_cnf_block(sd, blk); layers, T = co.block_weights(sd, blk)
    rng = np.random.default_rng(10 + blk)
    c = rng.standard_normal((333, cdim)) * 0.5
    y0 = (rng.standard_normal((333, 3)) * 0.6).astype(np.float32)
    ctx = torch.from_numpy((c @ pk["wctx"].double().numpy() + pk["bctx"].double().numpy()).astype(np.float32)).cuda()
    out, nfe = ops.cnf_flow(torch.from_numpy(y0).cuda(), ctx, pk["blob"].float().cuda(), 0.0, T, return_nfe=True)
    out = out.cpu().double().numpy(); nfe = nfe.cpu().numpy()
    ref, nr = co.dopri5(layers, y0.astype(np.float64), c, 0.0, T)
    tru = co.rk4_reference(layers, y0.astype(np.float64), c, 0.0, T, steps=2000)
    e1 = np.abs(out - tru).max(1); e2 = np.abs(ref - tru).max(1); e3 = np.abs(out-ref).max(1)
    i = e3.argmax()
    print(blk, "kernel vs truth", e1.max(), "oracle vs truth", e2.max(), "kernel vs oracle", e3.max(), "at", i, "nfe k/o", nfe[i], nr[i], "e1,e2 there", e1[i], e2[i], "|dy|", np.abs(tru[i]-y0[i]).max())
    print("   nfe kernel", nfe.min(), nfe.max(), "oracle", nr.min(), nr.max(), " p99 kernel-vs-truth", np.quantile(e1, .99))
