"""BASELINE.json configs[3] and configs[4] as a microbenchmark sweep (dev / profiling tool, not the bench line).

  config 4: kNN + EdgeConv sweep: k = 8 / 16 / 32, 256 - 4,096 pts per patch, batch 64 - 1,024
            (kNN: bytes-based GB/s AND pair-distances/s, the unit that binds; EdgeConv with block-3 weights, C = 128,
             growth 32: tensor-core kernel for k = 16, FP32 kernel otherwise)
  config 5: flow stack forward + inverse, FP32 kernels vs tcgen05 split-fp16 kernels, with the Chamfer distance between
            the two outputs as the validation figure

    python tools/sweep.py > gpurun_out/sweep.md
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from puflow_b200 import ops, packing
from puflow_b200.engine import PUFlowEngine
from tools import synth

dev = torch.device("cuda:0")
sd = torch.load("tests/golden/puflow-x4-pu1k.pt", map_location="cpu")
w = packing.pack_state_dict(sd, dev)
HBM = 6549.1


def timed(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


print("## config 4: kNN top-k sweep (B200, fp32 distances, bit-exact indices)\n")
print("| batch | pts/patch | k | ms | G pair-dist/s | algorithmic GB/s (12+4k B/query) | % of 6,549 GB/s |")
print("|---|---|---|---|---|---|---|")
for B, N in ((1024, 256), (256, 512), (256, 1024), (64, 2048), (64, 4096), (1024, 1024)):
    x = torch.from_numpy(synth.make_patches(min(B, 64), N, seed=1, shape="uniform")).to(dev).repeat(B // min(B, 64), 1, 1).contiguous()
    for K in (8, 16, 32):
        ms = timed(lambda: ops.knn_topk(x, x, K, return_dist=False))
        pairs = B * N * N / ms / 1e6
        gbs = B * N * (12 + 4 * K) / ms / 1e6
        print(f"| {B} | {N} | {K} | {ms:.3f} | {pairs:.0f} | {gbs:.1f} | {100 * gbs / HBM:.2f} |")

print("\n## config 4: EdgeConv sweep (block-3 weights: C_in = 128, growth 32, C_out = 128)\n")
print("| batch | pts/patch | k | kernel | ms | TFLOP/s (hidden-channel MACs x2) | algorithmic GB/s | % of 6,549 GB/s |")
print("|---|---|---|---|---|---|---|---|")
f = w["feat"][3]
for B, N in ((1024, 256), (256, 512), (256, 1024), (64, 2048), (64, 4096), (4096, 256)):
    x = torch.from_numpy(synth.make_patches(min(B, 64), N, seed=1, shape="uniform")).to(dev).repeat(B // min(B, 64), 1, 1).contiguous()
    pq = torch.randn(B * N, 512, device=dev)
    for K in (8, 16, 32):
        idx = ops.knn_topk(x, x, K, return_dist=False)[1]
        if K == 16:
            name, fn = "tcgen05", (lambda: ops.edgeconv_tc(pq, idx, f["wtc"], B, N, 32, 128, 0.05))
        else:
            name, fn = "fp32", (lambda: ops.edgeconv(pq, idx, f["wd"], f["wo"], B, N, 32, 128, 0.05))
        ms = timed(fn, 3)
        fl = 2.0 * (6 * 32 * 32 + 4 * 32 * 128) * B * N * K / ms / 1e9
        gbs = B * N * (2048 + 4 * K + 512) / ms / 1e6
        print(f"| {B} | {N} | {K} | {name} | {ms:.3f} | {fl:.1f} | {gbs:.1f} | {100 * gbs / HBM:.1f} |")

print("\n## config 5: flow stack forward + inverse, FP32 kernels vs tcgen05 split-fp16 kernels\n")
print("| input points | kernel set | cond ms | fwd ms | inv (x4) ms | fwd GB/s | inv GB/s | max abs diff of x vs FP32 | Chamfer(x_tc, x_fp32) |")
print("|---|---|---|---|---|---|---|---|---|")
eng = PUFlowEngine(sd, dev)
for Bp in (64, 1024, 4096):
    xyz = torch.from_numpy(synth.make_patches(min(Bp, 64), 256, seed=1, shape="surface")).to(dev).repeat(Bp // min(Bp, 64), 1, 1).contiguous()
    M = Bp * 256
    idx = eng.knn(xyz, 16)
    cs = eng.feat_extract(xyz, idx)
    raw = torch.empty((6, M, 192), device=dev)
    for i in range(6):
        ops.linear(cs[i].reshape(M, -1), w["wc"][i], None, 0, out=raw[i])
    x0 = xyz.reshape(M, 3)
    res = {}
    for name, cw, fw in (("fp32", w["cond"], w["flow"]), ("tcgen05", w["cond_tc"], w["flow_tc"])):
        tc_ms = timed(lambda: ops.flow_cond(raw, cw))
        sb = ops.flow_cond(raw, cw)
        tf = timed(lambda: ops.flow_fwd(x0, raw, sb, fw))
        z, _ = ops.flow_fwd(x0, raw, sb, fw)
        zup = z.repeat_interleave(4, 0) + 0.01 * torch.randn(M * 4, 3, device=dev, generator=torch.Generator(device=dev).manual_seed(0))
        ti = timed(lambda: ops.flow_inv(zup, raw, sb, fw, 4))
        res[name] = ops.flow_inv(zup, raw, sb, fw, 4)
        fb, ib = M * (24 + 4 * 6 * 72), M * (96 + 4 * 6 * 72)
        diff = cd = 0.0
        if name == "tcgen05":
            diff = (res["tcgen05"] - res["fp32"]).abs().max().item()
            nn_ = min(M * 4, 262144)
            a, b = res["tcgen05"][:nn_].view(-1, 4096, 3).contiguous(), res["fp32"][:nn_].view(-1, 4096, 3).contiguous()
            d1, d2, _, _ = ops.chamfer_nn(a, b)
            cd = (d1.mean() + d2.mean()).item()
        print(f"| {M} | {name} | {tc_ms:.3f} | {tf:.3f} | {ti:.3f} | {fb / tf / 1e6:.0f} | {ib / ti / 1e6:.0f} | {diff:.2e} | {cd:.2e} |")
