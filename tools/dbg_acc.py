"""Accumulation error of the split-fp16 tensor-core GEMM against FFMA and float64."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from puflow_b200 import ops, packing
dev = "cuda:0"
torch.manual_seed(0)
for K in (32, 64, 128):
    for kind in ("normal", "positive"):
        M, N = 8192, 128
        a = torch.randn(M, K); w = torch.randn(K, N) * 0.1
        if kind == "positive":
            a, w = a.abs(), w.abs()
        ref = a.double() @ w.double()
        rms = ref.pow(2).mean().sqrt()
        o_f = ops.linear(a.to(dev), w.to(dev), None, 0).cpu().double()
        o_t = ops.linear_tc(a.to(dev), packing.pack_linear_tc(w).to(dev), N, None, 0).cpu().double()
        o_32 = (a.to(dev) @ w.to(dev)).cpu().double()
        for name, o in (("ffma", o_f), ("tc", o_t), ("torch", o_32)):
            e = o - ref
            print(f"K={K} {kind:8s} {name:5s} bias {e.mean() / rms:+.2e} rms {e.pow(2).mean().sqrt() / rms:.2e} max {e.abs().max() / rms:.2e}")
