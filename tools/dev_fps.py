"""Dev tool: merge-FPS timing at config-2 size; with a -DPF_FPS_TIMING build also the per-phase cycle split of CTA 0."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from puflow_b200 import ops, _native
from tools import synth
B, n, npoint = 128, 40960, 8216
rng = np.random.default_rng(0)
clouds = synth.make_clouds(B, 2048, seed=3)
# candidates like the merge step: 32 patches x 1280 points around the cloud -> jittered copies of surface points
pts = np.stack([np.repeat(c, 20, 0) + rng.normal(0, 0.02, (n, 3)).astype(np.float32) for c in clouds]).astype(np.float32)
x = torch.from_numpy(pts).cuda()
for mode in ("bucket",):
    ops.fps(x, npoint, mode); torch.cuda.synchronize()
    if hasattr(_native.lib, "pf_fps_timing"):
        buf = (ctypes.c_longlong * 8)(); _native.lib.pf_fps_timing(buf)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ops.fps(x, npoint, mode); e1.record(); torch.cuda.synchronize()
    print(mode, f"{e0.elapsed_time(e1):.3f} ms", f"{e0.elapsed_time(e1) / npoint * 1e3:.3f} us/round")
    if hasattr(_native.lib, "pf_fps_timing"):
        buf = (ctypes.c_longlong * 8)(); _native.lib.pf_fps_timing(buf)
        names = ["loop/phase3", "phase1 (box+update)", "sync1", "phase2 (winner)", "sync2"]
        tot = sum(buf[:5])
        for nme, v in zip(names, buf[:5]):
            print(f"  {nme:22s} {v / npoint:8.1f} cycles/round  {100 * v / tot:5.1f}%")
