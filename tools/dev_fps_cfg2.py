"""Dev tool: the merge FPS of config 2 (128 clouds, 8,216 of 40,960 candidates) through pf_fps_queued, CUDA events, and a
check of the indices against the bucket kernel's."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from puflow_b200 import ops
from tools import synth
B, n, npoint = 128, 40960, 8216
x = torch.from_numpy(np.stack([synth.make_cloud(n, 31 + b, synth.SHAPES[b % 3]) for b in range(8)])).cuda().repeat(16, 1, 1).contiguous()
def timed(fn, k=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / k
ms = timed(lambda: ops.fps(x, npoint, mode="queue"))
same = torch.equal(ops.fps(x[:8], npoint, mode="queue"), ops.fps(x[:8], npoint, mode="bucket"))
print(f"fps queue B={B} n={n} npoint={npoint}: {ms:.3f} ms ({ms * 1e3 / npoint:.3f} us per sample); equal to the bucket kernel: {same}")
