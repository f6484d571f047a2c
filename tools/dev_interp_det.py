"""Dev tool: run-to-run bit equality of pf_interp_tc (a race in the tile pipeline shows up as differing low bits)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200.engine import PUFlowEngine
dev = "cuda:0"
sd = torch.load("tests/golden/puflow-x4-pu1k.pt", map_location="cpu")
E = PUFlowEngine(sd, dev)
for Bp, N in ((4096, 256), (96, 256), (7, 200)):
    torch.manual_seed(1)
    xyz = torch.randn(Bp, N, 3, device=dev) * 0.3
    z = torch.randn(Bp, N, 3, device=dev)
    idx = E.knn(xyz)
    ref = E.interp(z, xyz, 4, idx).clone()
    bad = 0
    for it in range(10):
        out = E.interp(z, xyz, 4, idx)
        d = (out != ref)
        if d.any():
            bad += 1
            if bad == 1:
                w = d.nonzero()
                rows = w[:, 0] * (N * 4) + w[:, 1]
                gt = torch.unique(rows // 64).tolist()          # global tile (16 points x 4 outputs per tile, N % 16 == 0)
                grid = min(148, Bp * (N // 16))
                print("   tiles (cta, k-th tile of the cta):", [(g % grid, g // grid) for g in gt][:40], "of", -(-Bp * (N // 16) // grid))
                print(f"  Bp={Bp}: run {it}: {int(d.sum())} values differ, max |diff| {float((out - ref).abs().max()):.3e}, first at {w[0].tolist()}, patches {sorted(set(w[:, 0].tolist()))[:8]}")
    o1 = E.interp(z, xyz, 4, idx).clone(); o2 = E.interp(z, xyz, 4, idx).clone()
    print(f"Bp={Bp} N={N}: {bad} of 10 repeats differ from the first run; two later runs equal: {bool(torch.equal(o1, o2))}")
