"""Dev check of pf_edgeconv_tc against pf_edgeconv (FFMA) on one block's real weights (run on the GPU box)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200 import ops
from puflow_b200.engine import PUFlowEngine
from tools import synth

Bp = int(sys.argv[1]) if len(sys.argv) > 1 else 8
sd = torch.load("tests/golden/puflow-x4-pu1k.pt", map_location="cpu")
eng = PUFlowEngine(sd, "cuda:0")
xyz = torch.from_numpy(synth.make_patches(Bp, 256, seed=1, shape="surface")).cuda()
idx = eng.knn(xyz, 16)
c = xyz.reshape(-1, 3)
for i in range(6):
    f = eng.w["feat"][i]
    from puflow_b200.packing import FEAT_CFG
    _, g, Co = FEAT_CFG[i]
    pq = ops.linear(c, f["wpq"], f["bpq"])
    ref = ops.edgeconv(pq, idx, f["wd"], f["wo"], Bp, 256, g, Co, 0.05)
    if "wtc" in f:
        got = ops.edgeconv_tc(pq, idx, f["wtc"], Bp, 256, g, Co, 0.05)
        torch.cuda.synchronize()
        err = (got - ref).abs().max().item()
        print(f"block {i}: max|tc-ffma| = {err:.3e}  max|ref| = {ref.abs().max().item():.3e}  nan={torch.isnan(got).any().item()}")
        if err > 1e-3 * ref.abs().max().item():
            bad = ((got - ref).abs() > 1e-3 * ref.abs().max()).nonzero()
            print("  first bad (row, ch):", bad[:8].tolist(), " n_bad", len(bad))
            print("  got", got[bad[0][0], :8].tolist()); print("  ref", ref[bad[0][0], :8].tolist())
    c = ref
if Bp >= 1024:
    f = eng.w["feat"][3]; pq = ops.linear(c, f["wpq"], f["bpq"])
    for name, fn in (("ffma", lambda: ops.edgeconv(pq, idx, f["wd"], f["wo"], Bp, 256, 32, 128, 0.05)),
                     ("tc", lambda: ops.edgeconv_tc(pq, idx, f["wtc"], Bp, 256, 32, 128, 0.05))):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): fn()
        e1.record(); torch.cuda.synchronize()
        print(f"{name}: {e0.elapsed_time(e1)/5:.3f} ms")
import ctypes
from puflow_b200 import _native
if hasattr(_native.lib, "pf_tc_timing"):
    buf = (ctypes.c_longlong * 64)()
    _native.lib.pf_tc_timing(buf)           # reset
    f = eng.w["feat"][3]; pq = ops.linear(c, f["wpq"], f["bpq"])
    ops.edgeconv_tc(pq, idx, f["wtc"], Bp, 256, 32, 128, 0.05)
    _native.lib.pf_tc_timing(buf)
    ntile = Bp * 32 / (148 * 2)
    names = ["tile top", "L0 epilogue", "L0 wait+bar", "Ll issue", "Ll readq/addp", "Ll mbar wait", "Ll tmem ld", "Ll epilogue",
             "Ll cpasync wait", "Ll fence+bar", "out issue", "out qprefetch", "out mbar wait", "out epilogue"]
    tot = sum(buf[:14])
    for k, nm in enumerate(names):
        print(f"  {nm:18s} {buf[k]/ntile:9.0f} cyc/tile  {100*buf[k]/tot:5.1f}%")
    print(f"  total {tot/ntile:.0f} cyc/tile")
