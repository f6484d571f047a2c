"""Per-stage clock64() totals of edgeconv_tc<32,128> (library built with EXTRA=-DPF_PROF)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200 import _native, ops
from puflow_b200.engine import PUFlowEngine
dev = "cuda:0"
sd = torch.load("tests/golden/puflow-x4-pu1k.pt", map_location="cpu")
E = PUFlowEngine(sd, dev)
Bp, N = 4096, 256
xyz = torch.randn(Bp, N, 3, device=dev) * 0.3
idx = E.knn(xyz)
c = torch.randn(Bp * N, 128, device=dev)
for _ in range(2):
    out = E.edgeconv_block(3, c, idx, Bp, N)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 32)()
lib = _native.lib
lib.pf_debug_edgeconv_prof.argtypes = [ctypes.c_void_p]
assert lib.pf_debug_edgeconv_prof(buf) == 0
v = list(buf)
names = ["l0 wait ring", "l0 read+lrelu+split+store+release", "l1-3 wait ring", "l1-3 read P+Q", "l1-3 wait MMA",
         "l1-3 ld+lrelu+split+store+release", "out wait ring", "out wait MMA", "out ld+fma+max+store", "-"]
tiles = -(-Bp * N * 16 // 128) // (148 * 2) + 1
tot = sum(v[:10])
print(f"epilogue thread 0 of CTA 0 (warpgroup 0): {tot} cycles, ~{tiles} tiles -> {tot / tiles:.0f} cycles per tile")
for n, c_ in zip(names, v[:10]):
    print(f"  {n:36s} {c_ / tiles:8.0f} cyc/tile  {100 * c_ / tot:5.1f}%")
print(f"MMA thread (wg 0): issue {v[16] / tiles:.0f} cyc/tile, wait {v[17] / tiles:.0f} cyc/tile")
