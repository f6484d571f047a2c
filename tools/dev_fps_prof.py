"""Per-phase clock64() totals of fps_queue_kernel at config 2 (library built with EXTRA=-DPF_PROF): thread 0 (the chain warp)
and thread 32 (a sweep-only warp) of CTA 0."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from puflow_b200 import _native, ops
from tools import synth
B, n, npoint = 128, 40960, 8216
x = torch.from_numpy(np.stack([synth.make_cloud(n, 31 + b, synth.SHAPES[b % 3]) for b in range(8)])).cuda().repeat(16, 1, 1).contiguous()
for _ in range(2):
    ops.fps(x, npoint, mode="queue")
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 32)()
lib = _native.lib
lib.pf_debug_fps_prof.argtypes = [ctypes.c_void_p]
assert lib.pf_debug_fps_prof(buf) == 0
v = list(buf)
names = ["update sweep", "wait B1", "collect: write sweep (after the slot atomic)", "wait B2", "chain (warp 0) / idle", "wait B3",
         "collect: block max + theta", "collect: count sweep"]
for base, who in ((0, "thread 0 (chain warp)"), (16, "thread 32 (warp 1)")):
    tot = sum(v[base:base + 8])
    print(f"{who}: {tot} cycles, {v[base + 11]} passes, {v[base + 12]} collect attempts -> {tot / max(v[base + 11], 1):.0f} cycles per pass")
    for k, nm in enumerate(names):
        print(f"  {nm:46s} {v[base + k]:10d}  {100 * v[base + k] / tot:5.1f}%  {v[base + k] / max(v[base + 11], 1):8.0f} per pass")
    print(f"  this warp per pass: {v[base + 8] / v[base + 11]:.1f} buckets updated in {v[base + 9] / v[base + 11]:.1f} batches of <= 4")
