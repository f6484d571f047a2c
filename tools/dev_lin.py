"""Dev timing of pf_linear_tc vs pf_linear at config-2 sizes (run on the GPU box)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200 import ops
from puflow_b200.packing import pack_linear_tc
M = 1 << 20
for K, N in ((128, 512), (128, 192), (128, 64), (64, 128), (64, 512), (32, 256)):
    a = torch.randn(M, K, device="cuda"); w = torch.randn(K, N, device="cuda") / K ** 0.5; b = torch.randn(N, device="cuda")
    wtc = pack_linear_tc(w.cpu()).cuda()
    r = {}
    for name, fn in (("ffma", lambda: ops.linear(a, w, b)), ("tc", lambda: ops.linear_tc(a, wtc, N, b))):
        o = fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): o = fn()
        e1.record(); torch.cuda.synchronize()
        r[name] = (e0.elapsed_time(e1) / 5, o)
    err = (r["tc"][1] - r["ffma"][1]).abs().max().item()
    gb = M * 4 * (K + N) / 1e9
    print(f"K={K:3d} N={N:3d}  ffma {r['ffma'][0]:6.3f} ms  tc {r['tc'][0]:6.3f} ms ({gb / r['tc'][0] * 1e3:6.0f} GB/s)  max|diff| {err:.2e}")
