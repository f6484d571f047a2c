"""Dev tool: function evaluations per point of the inverse CNF blocks, and what lock-step lanes cost (warp / CTA maxima)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200 import ops
from puflow_b200.engine import CNFEngine
from tools import synth
sd = torch.load("tests/golden/puflow-x4-cnf-pu1k.pt", map_location="cpu")
E = CNFEngine(sd, "cuda:0")
xyz = torch.from_numpy(synth.make_patches(256, 256, seed=1, shape="surface")).cuda()
r = E.forward(xyz, 4, return_all=True)
ctxs = E.contexts(r["cs"])
y = r["zup"].reshape(-1, 3).contiguous()
tot = totw = totc = 0.0
for i in reversed(range(len(E.w["cnf"]))):
    b = E.w["cnf"][i]
    y, nfe = ops.cnf_flow(y, ctxs[i], b["blob"], b["T"], 0.0, 4, E.atol, E.rtol, return_nfe=True)
    n = nfe.float()
    w = n.view(-1, 32).max(1).values
    c = n.view(-1, 128).max(1).values
    q = n.view(-1, 4).max(1).values                      # the 4 points that share a conditioning row
    print(f"   mean of row-quad maxima {q.mean():.1f}; quad sums per CTA-sized chunk of 590 rows: max/mean "
          f"{(q[:q.numel() // 590 * 590].view(-1, 590).sum(1).max() / q[:q.numel() // 590 * 590].view(-1, 590).sum(1).mean()):.3f}")
    print(f"inverse block {i}: nfe mean {n.mean():.1f} max {n.max():.0f}; mean of warp maxima {w.mean():.1f}; of CTA maxima {c.mean():.1f}")
    tot += n.mean().item(); totw += w.mean().item(); totc += c.mean().item()
print(f"all blocks: lanes useful {tot / totw:.3f} of the warp time; warp-max / CTA-max {totw / totc:.3f}")

# Would ordering the rows by the forward pass's evaluation count (same block, same conditioning row) help the inverse?
print("--- rows ordered by the forward pass's nfe")
yf = xyz.reshape(-1, 3).contiguous()
fn = []
for i in range(len(E.w["cnf"])):
    b = E.w["cnf"][i]
    yf, nf = ops.cnf_flow(yf, ctxs[i], b["blob"], 0.0, b["T"], 1, E.atol, E.rtol, return_nfe=True)
    fn.append(nf)
y = r["zup"].reshape(-1, 3).contiguous()
tot = totw = totc = 0.0
for i in reversed(range(len(E.w["cnf"]))):
    b = E.w["cnf"][i]
    y, nfe = ops.cnf_flow(y, ctxs[i], b["blob"], b["T"], 0.0, 4, E.atol, E.rtol, return_nfe=True)
    n = nfe.float().view(-1, 4)
    order = torch.argsort(fn[i])
    ns = n[order].reshape(-1)
    w = ns.view(-1, 32).max(1).values
    c = ns.view(-1, 128).max(1).values
    cc = torch.corrcoef(torch.stack([fn[i].float(), n.max(1).values]))[0, 1]
    print(f"inverse block {i}: corr(forward nfe, quad max) {cc:.2f}; sorted: mean {ns.mean():.1f}, warp maxima {w.mean():.1f}, CTA maxima {c.mean():.1f}")
    tot += ns.mean().item(); totw += w.mean().item(); totc += c.mean().item()
print(f"sorted, all blocks: lanes useful {tot / totw:.3f} of the warp time; warp-max / CTA-max {totw / totc:.3f}")
