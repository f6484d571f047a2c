"""Dev tool: CUDA-event timing of the FPS kernels: a batch (config 2), few large clouds (config 3), one cloud."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from puflow_b200 import ops
from tools import synth
dev = "cuda:0"
def timed(fn, n=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for (B, n, npoint) in [(128, 40960, 8216), (8, 99840, 20024), (1, 99840, 20024), (1, 40960, 8216), (2, 40960, 8216), (16, 40960, 8216)]:
    x = torch.from_numpy(np.stack([synth.make_cloud(n, 31 + b, synth.SHAPES[b % 3]) for b in range(min(B, 8))])).to(dev)
    x = x.repeat((B + 7) // 8, 1, 1)[:B].contiguous()
    for mode in ("bucket", "cluster", "queue"):
        ms = timed(lambda: ops.fps(x, npoint, mode=mode))
        print(f"fps B={B} n={n} npoint={npoint} {mode:8s}: {ms:8.3f} ms  {ms * 1e3 / npoint:.3f} us/round")
