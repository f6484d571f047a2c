"""Dev tool: CUDA-event timing of pf_knn_topk / pf_chamfer_nn at config-2 sizes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200 import ops
dev = "cuda:0"
def timed(fn, n=5):
    for _ in range(2): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for (B, N, K) in [(4096, 256, 16), (4096, 256, 8), (4096, 256, 32), (1024, 1024, 16), (64, 4096, 16)]:
    x = torch.randn(B, N, 3, device=dev) * 0.3
    ms = timed(lambda: ops.knn_topk(x, x, K, return_dist=False))
    print(f"knn B={B} N={N} K={K}: {ms:.3f} ms  {B * N * N / ms / 1e9:.3f} T pairs/s")
a = torch.randn(128, 8216, 3, device=dev); b = torch.randn(128, 2048, 3, device=dev)
ms = timed(lambda: ops.chamfer_nn(a, b, both=False))
print(f"chamfer a->b 128 x 8216 x 2048: {ms:.3f} ms  {128 * 8216 * 2048 / ms / 1e9:.3f} T pairs/s")
