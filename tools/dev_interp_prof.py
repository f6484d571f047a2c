"""Per-stage clock64() totals of interp_tc (library built with EXTRA=-DPF_PROF)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200 import _native, ops, packing
from puflow_b200.engine import PUFlowEngine
dev = "cuda:0"
sd = torch.load("tests/golden/puflow-x4-pu1k.pt", map_location="cpu")
E = PUFlowEngine(sd, dev)
Bp, N = 4096, 256
xyz = torch.randn(Bp, N, 3, device=dev) * 0.3
z = torch.randn(Bp, N, 3, device=dev)
idx = E.knn(xyz)
for _ in range(2):
    out = E.interp(z, xyz, 4, idx)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 32)()
lib = _native.lib if hasattr(_native, "lib") else _native._lib
lib.pf_debug_interp_prof.argtypes = [ctypes.c_void_p]
assert lib.pf_debug_interp_prof(buf) == 0
v = list(buf)
names = ["geo+publish", "chain: wait acc", "chain: act+slice+publish", "(side work / chain end)", "wait t0", "epi t0",
         "wait t1", "epi t1", "wait logits", "softmax+store"]
tiles = -(-Bp * N // 16) // 148 + 1
for base, who in ((0, "warpgroup 0 (dense chain), thread 0"), (20, "warpgroup 1 (side work), thread 128")):
    tot = sum(v[base:base + 10])
    print(f"{who} of CTA 0: {tot} cycles total, ~{tiles} tiles -> {tot / tiles:.0f} cycles per tile")
    for n, c in zip(names, v[base:base + 10]):
        print(f"  {n:28s} {c / tiles:8.0f} cyc/tile  {100 * c / tot:5.1f}%")
print("warpgroup 1 softmax split: ld logits", v[30] / tiles, " wait z", v[31] / tiles)
print(f"MMA thread: wait {v[16] / tiles:.0f} cyc/tile, issue {v[17] / tiles:.0f} cyc/tile")
