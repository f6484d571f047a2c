"""Dev tool: the threshold-queue FPS kernel against the bucket kernel (indices must be identical) + CUDA-event timing."""
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
def clouds(B, n, kind):
    rng = np.random.default_rng(n + B)
    if kind == "surface":
        return np.stack([synth.make_cloud(n, 31 + b, synth.SHAPES[b % 3]) for b in range(B)])
    if kind == "uniform":
        return rng.uniform(-1, 1, (B, n, 3)).astype(np.float32)
    if kind == "dupfar":      # 400 exact copies of each of 6 far points: more ties than the queue holds
        p = rng.uniform(-1, 1, (B, n, 3)).astype(np.float32)
        for j in range(6):
            p[:, 400 * j:400 * (j + 1)] = rng.uniform(2, 3, (B, 1, 3)).astype(np.float32) * (1 if j % 2 else -1)
        return p[:, rng.permutation(n)]
    if kind == "lattice":
        return np.stack([synth.make_cloud(n, 5 + b, "lattice") for b in range(B)])
    if kind == "same":
        return np.full((B, n, 3), 0.25, np.float32)
    if kind == "dup":
        base = rng.uniform(-1, 1, (B, n // 3, 3)).astype(np.float32)
        return np.concatenate([base, base, base], 1)[:, rng.permutation(n // 3 * 3)]
cases = [(128, 40960, 8216, "surface"), (8, 99840, 20024, "surface"), (1, 99840, 20024, "surface"), (1, 40960, 8216, "surface"),
         (4, 2048, 40, "surface"), (4, 5000, 78, "surface"), (3, 20000, 900, "uniform"), (2, 9000, 9000, "dup"),
         (2, 8000, 500, "lattice"), (2, 6000, 3000, "dupfar"), (1, 6000, 100, "same"), (2, 100, 100, "uniform"), (2, 33, 5, "uniform")]
if len(sys.argv) > 1: cases = cases[:int(sys.argv[1])]
for (B, n, npoint, kind) in cases:
    p = clouds(min(B, 8), n, kind)
    x = torch.from_numpy(np.ascontiguousarray(p)).to(dev)
    x = x.repeat((B + 7) // 8, 1, 1)[:B].contiguous()
    n = x.shape[1]; npoint = min(npoint, n)
    ref = ops.fps(x, npoint, mode="bucket")
    out = ops.fps(x, npoint, mode="queue")
    torch.cuda.synchronize()
    same = bool((ref == out).all())
    first = int((ref != out).any(0).nonzero()[0]) if not same else -1
    tb = timed(lambda: ops.fps(x, npoint, mode="bucket")); tq = timed(lambda: ops.fps(x, npoint, mode="queue"))
    print(f"fps B={B} n={n} npoint={npoint} {kind:8s}: bucket {tb:8.3f} ms  queue {tq:8.3f} ms ({tq * 1e3 / npoint:.3f} us/sample)  identical={same} first_diff={first}", flush=True)
