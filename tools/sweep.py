"""BASELINE.json configs[3] and configs[4] as a microbenchmark sweep (dev / profiling tool, not the bench line).

  config 4: kNN + EdgeConv sweep: k = 8 / 16 / 32, 256 - 4,096 pts per patch, batch 64 - 1,024
            (kNN: bytes-based GB/s AND pair-distances/s, the unit that binds; EdgeConv with block-3 weights, C = 128,
             growth 32: FP32 FFMA kernel, tcgen05 split-fp16 (fp32-class) and tcgen05 single-pass fp16 for every k)
  config 5: flow stack forward + inverse: FP32 FFMA kernels vs tcgen05 split-fp16 (fp32-class) vs tcgen05 single-pass fp16
            (TF32-class mantissa), with the max difference and the Chamfer distance to the FP32 output as validation

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
import json
try:
    HBM = float(json.load(open('MEASURED_PEAKS.json'))['hbm_gbs'])
except Exception:
    HBM = 6650.0


def timed(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


print("## config 4: kNN top-k sweep (B200, fp32 distances, bit-exact indices)\n")
print(f"| batch | pts/patch | k | ms | G pair-dist/s | algorithmic GB/s (12+4k B/query) | % of {HBM:.0f} GB/s |")
print("|---|---|---|---|---|---|---|")
for B, N in ((1024, 256), (256, 512), (256, 1024), (64, 2048), (64, 4096), (1024, 1024)):
    x = torch.from_numpy(synth.make_patches(min(B, 64), N, seed=1, shape="uniform")).to(dev).repeat(B // min(B, 64), 1, 1).contiguous()
    for K in (8, 16, 32):
        ms = timed(lambda: ops.knn_topk(x, x, K, return_dist=False))
        pairs = B * N * N / ms / 1e6
        gbs = B * N * (12 + 4 * K) / ms / 1e6
        print(f"| {B} | {N} | {K} | {ms:.3f} | {pairs:.0f} | {gbs:.1f} | {100 * gbs / HBM:.2f} |")

print("\n## config 4: EdgeConv sweep (block-3 weights: C_in = 128, growth 32, C_out = 128)\n")
print(f"| batch | pts/patch | k | kernel | ms | TFLOP/s (hidden-channel MACs x2) | algorithmic GB/s | % of {HBM:.0f} GB/s |")
print("|---|---|---|---|---|---|---|---|")
f = w["feat"][3]
for B, N in ((1024, 256), (256, 512), (256, 1024), (64, 2048), (64, 4096), (4096, 256)):
    x = torch.from_numpy(synth.make_patches(min(B, 64), N, seed=1, shape="uniform")).to(dev).repeat(B // min(B, 64), 1, 1).contiguous()
    pq = torch.randn(B * N, 512, device=dev)
    for K in (8, 16, 32):
        idx = ops.knn_topk(x, x, K, return_dist=False)[1]
        variants = [("tcgen05 split-fp16 (fp32-class)", lambda: ops.edgeconv_tc(pq, idx, f["wtc"], B, N, 32, 128, 0.05)),
                    ("tcgen05 single-pass fp16", lambda: ops.edgeconv_tc(pq, idx, f["wtc"], B, N, 32, 128, 0.05, precision=1))]
        if B * N <= 262144:
            variants.insert(0, ("fp32 FFMA", lambda: ops.edgeconv(pq, idx, f["wd"], f["wo"], B, N, 32, 128, 0.05)))
        for name, fn in variants:
            ms = timed(fn, 3)
            fl = 2.0 * (6 * 32 * 32 + 4 * 32 * 128) * B * N * K / ms / 1e9
            gbs = B * N * (2048 + 4 * K + 512) / ms / 1e6
            print(f"| {B} | {N} | {K} | {name} | {ms:.3f} | {fl:.1f} | {gbs:.1f} | {100 * gbs / HBM:.1f} |")

print("\n## config 5: flow stack forward + inverse: FP32 FFMA vs tcgen05 split-fp16 (fp32-class) vs tcgen05 single-pass fp16\n")
print("| input points | arithmetic | cond ms | fwd ms | inv (x4) ms | fwd GB/s | inv GB/s | max abs diff of x vs FP32 | Chamfer(x, x_fp32) |")
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
    del cs
    x0 = xyz.reshape(M, 3)
    res = {}
    for name, cw, fw, prec in (("fp32 FFMA", w["cond"], w["flow"], 0), ("tcgen05 split-fp16", w["cond_tc"], w["flow_tc"], 0),
                               ("tcgen05 single-pass fp16", w["cond_tc"], w["flow_tc"], 1)):
        tc_ms = timed(lambda: ops.flow_cond(raw, cw, precision=prec))
        sb = ops.flow_cond(raw, cw, precision=prec)
        tf = timed(lambda: ops.flow_fwd(x0, raw, sb, fw, precision=prec))
        z, _ = ops.flow_fwd(x0, raw, sb, fw, precision=prec)
        if name == "fp32 FFMA":
            zup = z.repeat_interleave(4, 0) + 0.01 * torch.randn(M * 4, 3, device=dev, generator=torch.Generator(device=dev).manual_seed(0))
        ti = timed(lambda: ops.flow_inv(zup, raw, sb, fw, 4, precision=prec))
        res[name] = ops.flow_inv(zup, raw, sb, fw, 4, precision=prec)
        fb, ib = M * (24 + 4 * 6 * 72), M * (96 + 4 * 6 * 72)
        diff = cd = 0.0
        if name != "fp32 FFMA":
            diff = (res[name] - res["fp32 FFMA"]).abs().max().item()
            nn_ = min(M * 4, 262144)
            a, b = res[name][:nn_].view(-1, 4096, 3).contiguous(), res["fp32 FFMA"][:nn_].view(-1, 4096, 3).contiguous()
            d1, d2, _, _ = ops.chamfer_nn(a, b)
            cd = (d1.mean() + d2.mean()).item()
        print(f"| {M} | {name} | {tc_ms:.3f} | {tf:.3f} | {ti:.3f} | {fb / tf / 1e6:.0f} | {ib / ti / 1e6:.0f} | {diff:.2e} | {cd:.2e} |")

print("\n## whole network, precision axis (pf_upsample_patches, 4,096 patches of 256 points)\n")
print("| arithmetic | ms | generated points/s | x normwise error vs reference fixture (pu1k, 3 patches) |")
print("|---|---|---|---|")
xyz = torch.from_numpy(synth.make_patches(64, 256, seed=1, shape="surface")).to(dev).repeat(64, 1, 1).contiguous()
gold = np.load("tests/golden/net_pu1k.npz")
for prec in ("fp32", "fp16"):
    e = PUFlowEngine(sd, dev, precision=prec)
    ms = timed(lambda: e.forward_c(xyz, 4), 3)
    x = e.forward(torch.from_numpy(gold["xyz"]).to(dev), 4)[0].cpu().numpy()
    err = (np.abs(x - gold["x"]).reshape(x.shape[0], -1).max(1) / np.abs(gold["x"]).reshape(x.shape[0], -1).max(1)).max()
    print(f"| {'tcgen05 split-fp16 (fp32-class)' if prec == 'fp32' else 'tcgen05 single-pass fp16'} | {ms:.2f} | {4096 * 1024 / ms * 1e3:.3e} | {err:.2e} |")
