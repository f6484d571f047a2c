"""Dev tool: FFMA vs tensor-core flow kernels at config-2 size (timing + max difference)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200 import ops, packing

M = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
dev = torch.device("cuda:0")
sd = torch.load("tests/golden/puflow-x4-pu1k.pt", map_location="cpu")
w = packing.pack_state_dict(sd, dev)
g = torch.Generator(device=dev).manual_seed(0)
x = torch.randn(M, 3, device=dev, generator=g) * 0.5
raw = torch.randn(6, M, 192, device=dev, generator=g) * 0.5
zin = torch.randn(M * 4, 3, device=dev, generator=g) * 0.5


def timed(name, fn, n=5):
    out = fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        out = fn()
    e1.record(); torch.cuda.synchronize()
    print(f"{name:16s} {e0.elapsed_time(e1) / n:8.3f} ms", flush=True)
    return out


sb0 = timed("cond ffma", lambda: ops.flow_cond(raw, w["cond"]))
sb1 = timed("cond tc", lambda: ops.flow_cond(raw, w["cond_tc"]))
print("  sb maxdiff", (sb0 - sb1).abs().max().item(), "max", sb0.abs().max().item())
z0, s0 = timed("fwd ffma", lambda: ops.flow_fwd(x, raw, sb0, w["flow"]))
z1, s1 = timed("fwd tc", lambda: ops.flow_fwd(x, raw, sb0, w["flow_tc"]))
print("  z maxdiff", (z0 - z1).abs().max().item(), "max", z0.abs().max().item())
x0 = timed("inv ffma", lambda: ops.flow_inv(zin, raw, sb0, w["flow"], 4))
x1 = timed("inv tc", lambda: ops.flow_inv(zin, raw, sb0, w["flow_tc"], 4))
fin = torch.isfinite(x0).all(1)
print("  x maxdiff", (x0 - x1)[fin].abs().max().item(), "max", x0[fin].abs().max().item(), "finite frac", fin.float().mean().item())
