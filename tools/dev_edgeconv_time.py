"""Dev tool: CUDA-event timing of pf_edgeconv_tc for the three block shapes at config-2 size (Bp = 4096, N = 256)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200 import ops
from puflow_b200.engine import PUFlowEngine
from puflow_b200.packing import FEAT_CFG
dev = "cuda:0"
sd = torch.load("tests/golden/puflow-x4-pu1k.pt", map_location="cpu")
E = PUFlowEngine(sd, dev)
Bp, N = 4096, 256
xyz = torch.randn(Bp, N, 3, device=dev) * 0.3
for K in (16, 8, 32):
    _, idx = ops.knn_topk(xyz, xyz, K, return_dist=False)
    for blk in ((0, 1, 3) if K == 16 else (3,)):
        C, g, Co = FEAT_CFG[blk]
        f = E.w["feat"][blk]
        pq = torch.randn(Bp * N, 2 * (4 * g + Co), device=dev)
        split = ops.split_tiles(Bp * N, Co, dev)
        for sp in (None, split):
            for _ in range(2):
                ops.edgeconv_tc(pq, idx, f["wtc"], Bp, N, g, Co, split=sp)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                ops.edgeconv_tc(pq, idx, f["wtc"], Bp, N, g, Co, split=sp)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            fl = 2.0 * (6 * g * g + 4 * g * Co) * Bp * N * K
            print(f"edgeconv_tc g={g} co={Co} K={K} split={'y' if sp is not None else 'n'}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s")
# single-pass fp16 variant of the main block (precision switch)
_, idx = ops.knn_topk(xyz, xyz, 16, return_dist=False)
C, g, Co = FEAT_CFG[3]
f = E.w["feat"][3]
pq = torch.randn(Bp * N, 2 * (4 * g + Co), device=dev)
for prec in (0, 1):
    for _ in range(2):
        ops.edgeconv_tc(pq, idx, f["wtc"], Bp, N, g, Co, precision=prec)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        ops.edgeconv_tc(pq, idx, f["wtc"], Bp, N, g, Co, precision=prec)
    e1.record(); torch.cuda.synchronize()
    print(f"edgeconv_tc g=32 co=128 K=16 precision={prec}: {e0.elapsed_time(e1) / 5:.3f} ms")
