"""Dev tool: pf_chamfer_nn (grid-pruned from 256 candidates) against the brute-force kernel (pf_knn_topk, K = 1) at config 2's
shape: 128 clouds, 8,192 up-sampled points against the 2,048 input points."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from puflow_b200 import ops
from tools import synth
B = int(os.environ.get("PF_B", "128"))
b = torch.from_numpy(synth.make_clouds(B, 2048, seed=5)).cuda()
rng = np.random.default_rng(0)
a = b.repeat_interleave(4, dim=1) + torch.from_numpy(rng.normal(0, 0.01, (B, 8192, 3)).astype(np.float32)).cuda()
a = a[:, torch.randperm(8192, device="cuda")].contiguous()       # FPS order is spatially random
def t(f, n=10):
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
tg = t(lambda: ops.chamfer_nn(a, b, both=False))
tb = t(lambda: ops.knn_topk(a, b, 1))
d1, _, i1, _ = ops.chamfer_nn(a, b, both=False)
dk, ik = ops.knn_topk(a, b, 1)
print(f"chamfer a->b, {B} x 8192 vs 2048: grid {tg:.3f} ms, brute force {tb:.3f} ms; equal: {torch.equal(i1, ik[..., 0]) and torch.equal(d1, dk[..., 0])}")
