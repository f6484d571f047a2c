"""GPU parity tests: every C-ABI kernel against the CPU oracle (oracle/) / the golden fixtures of the
unmodified reference (tests/golden), through puflow_b200.ops (ctypes -> C-ABI).

Tolerances (tests/parity.py): indices bit-exact; fp32 tensors normwise per batch item, 1e-4 on final
outputs and per-kernel outputs, 1e-3 on the ill-conditioned latent intermediates (z, zup)."""
import numpy as np
import pytest
import torch

from oracle import puflow_oracle as po
from oracle import synth
from tests import emulate, parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def ops():
    from puflow_b200 import ops as _ops
    return _ops


@pytest.fixture(scope="module")
def packed(state_dict):
    from puflow_b200 import packing
    return packing.pack_state_dict(state_dict, "cpu")


def _to(o, dev):
    if isinstance(o, torch.Tensor):
        return o.to(dev)
    if isinstance(o, dict):
        return {k: _to(v, dev) for k, v in o.items()}
    if isinstance(o, list):
        return [_to(v, dev) for v in o]
    return o


# ---- (1) kNN ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("B,N,K,shape", [(8, 256, 16, "surface"), (3, 256, 8, "uniform"), (2, 1024, 32, "uniform"),
                                         (2, 2500, 16, "uniform"), (3, 250, 16, "uniform"), (2, 37, 4, "uniform"),
                                         (1, 4096, 16, "uniform"), (5, 16, 16, "uniform")])
def test_knn_bit_exact(ops, dev, B, N, K, shape):
    if shape == "surface":
        p = synth.make_patches(B, N, seed=B * N + K, shape="surface")
    else:
        p = np.random.default_rng(B * N + K).uniform(-1, 1, (B, N, 3)).astype(np.float32)
    d_ref, i_ref = po.knn(p, p, K)
    t = torch.from_numpy(p).to(dev)
    d, i = ops.knn_topk(t, t, K)
    assert np.array_equal(i.cpu().numpy().astype(np.int64), i_ref)
    assert np.array_equal(d.cpu().numpy(), d_ref)


def test_knn_lattice_ties(ops, dev):
    p = np.stack([synth.make_cloud(343, s, "lattice") for s in (3, 4)])
    _, i_ref = po.knn(p, p, 16)
    t = torch.from_numpy(p).to(dev)
    _, i = ops.knn_topk(t, t, 16)
    assert np.array_equal(i.cpu().numpy().astype(np.int64), i_ref)


def test_knn_query_set_differs_from_points(ops, dev):
    rng = np.random.default_rng(0)
    q = rng.uniform(-1, 1, (4, 77, 3)).astype(np.float32)
    p = rng.uniform(-1, 1, (4, 1500, 3)).astype(np.float32)
    d_ref, i_ref = po.knn(q, p, 8)
    d, i = ops.knn_topk(torch.from_numpy(q).to(dev), torch.from_numpy(p).to(dev), 8)
    assert np.array_equal(i.cpu().numpy().astype(np.int64), i_ref) and np.array_equal(d.cpu().numpy(), d_ref)


def test_knn_golden_and_errors(ops, dev, golden):
    g = golden("net_pu1k.npz")
    t = torch.from_numpy(g["xyz"]).to(dev)
    d, i = ops.knn_topk(t, t, 16)
    assert np.array_equal(i.cpu().numpy(), g["knn_idx"]) and np.array_equal(d.cpu().numpy(), g["knn_d"])
    from puflow_b200._native import PuflowError
    with pytest.raises(PuflowError):
        ops.knn_topk(t[:, :8].contiguous(), t[:, :8].contiguous(), 16)          # K > np
    with pytest.raises(PuflowError):
        ops.knn_topk(t.cpu(), t.cpu(), 16)                                       # no CPU fallback
    e = torch.empty((0, 256, 3), device=dev)
    assert ops.knn_topk(e, e, 16)[1].shape == (0, 256, 16)                       # empty batch


# ---- (5) Chamfer -----------------------------------------------------------------------------------
def test_chamfer_known_answer(ops, dev, golden):
    """The reference's own unit test (metric/PyTorchCD/unit_test.py:14-35): MSE < 1e-8, indices identical."""
    g = golden("chamfer_unit.npz")
    d1, d2, i1, i2 = [t.cpu().numpy() for t in ops.chamfer_nn(torch.from_numpy(g["a"]).to(dev),
                                                              torch.from_numpy(g["b"]).to(dev))]
    assert np.mean((d1 - g["dist1"]) ** 2) + np.mean((d2 - g["dist2"]) ** 2) < 1e-8
    assert np.array_equal(i1, g["idx1"]) and np.array_equal(i2, g["idx2"])


@pytest.mark.parametrize("B,n,m", [(2, 8216, 2048), (3, 101, 57), (1, 5000, 20024), (4, 1, 3)])
def test_chamfer_bit_exact(ops, dev, B, n, m):
    rng = np.random.default_rng(n + m)
    a = rng.uniform(-1, 1, (B, n, 3)).astype(np.float32)
    b = rng.uniform(-1, 1, (B, m, 3)).astype(np.float32)
    if n > 100:
        b[:, : min(m, 50)] = a[:, : min(m, 50)]            # exact zeros + duplicate candidates (ties)
        b[:, 1] = b[:, 0]
    ref = po.chamfer_nn(a, b)
    out = ops.chamfer_nn(torch.from_numpy(a).to(dev), torch.from_numpy(b).to(dev))
    for o, r in zip(out, ref):
        assert np.array_equal(o.cpu().numpy(), r)


@pytest.mark.parametrize("case", ["surface", "plane", "line", "one_point", "lattice", "far_queries", "clusters", "nonfinite",
                                  "np256", "np1000", "np8192"])
def test_chamfer_grid_pruning_is_exact(ops, dev, case):
    """pf_chamfer_nn prunes with a grid from 256 candidates per cloud (csrc/nn.cu nn_grid_kernel); pf_knn_topk(K = 1) is the
    brute-force kernel.  Same distances and indices, bit for bit (chamfer3D.cu:36-69,126: strict <, the lowest index wins),
    on the geometries that stress the pruning bound: degenerate bounding boxes, exact ties, queries outside the box,
    clustered candidates (empty cells), non-finite coordinates; and against the C oracle on the smaller ones."""
    rng = np.random.default_rng(len(case))
    B, n, m = 3, 4096, 2048
    if case == "surface":
        b = synth.make_clouds(B, m, seed=3)
        a = np.repeat(b, 2, axis=1) + rng.normal(0, 0.01, (B, 2 * m, 3)).astype(np.float32)
    elif case == "plane":
        b = rng.uniform(-1, 1, (B, m, 3)).astype(np.float32); b[..., 2] = 0.25
        a = rng.uniform(-1, 1, (B, n, 3)).astype(np.float32)
    elif case == "line":
        b = np.zeros((B, m, 3), np.float32); b[..., 0] = rng.uniform(-1, 1, (B, m))
        a = rng.uniform(-1, 1, (B, n, 3)).astype(np.float32)
    elif case == "one_point":
        b = np.full((B, m, 3), 0.5, np.float32)
        a = rng.uniform(-1, 1, (B, n, 3)).astype(np.float32)
    elif case == "lattice":                                   # equal distances everywhere: the tie order must survive the pruning
        g = np.stack(np.meshgrid(*[np.arange(16)] * 3, indexing="ij"), -1).reshape(-1, 3).astype(np.float32)
        b = np.stack([g[rng.permutation(4096)][:m] for _ in range(B)])
        a = np.stack([g[rng.permutation(4096)] + np.float32(0.5) for _ in range(B)])
    elif case == "far_queries":
        b = rng.uniform(-1, 1, (B, m, 3)).astype(np.float32)
        a = (rng.uniform(-1, 1, (B, n, 3)) * np.array([1000.0, 3.0, 1.0])).astype(np.float32)
    elif case == "clusters":
        c = rng.uniform(-1, 1, (B, 8, 1, 3)); b = (c + rng.normal(0, 0.01, (B, 8, m // 8, 3))).reshape(B, m, 3).astype(np.float32)
        a = rng.uniform(-1.5, 1.5, (B, n, 3)).astype(np.float32)
    elif case == "nonfinite":
        b = rng.uniform(-1, 1, (B, m, 3)).astype(np.float32)
        b[0, 5, 1] = np.nan; b[1, 7, 0] = np.inf                # cloud 0: a NaN candidate is never chosen; cloud 1: unbounded box
        a = rng.uniform(-1, 1, (B, n, 3)).astype(np.float32); a[2, 3, 2] = np.nan
    else:
        m = int(case[2:])
        b = rng.uniform(-1, 1, (B, m, 3)).astype(np.float32)
        a = rng.uniform(-1, 1, (B, 1500, 3)).astype(np.float32)
        a[:, :40] = b[:, :40]                                   # exact zeros
    ta, tb = torch.from_numpy(a).to(dev), torch.from_numpy(b).to(dev)
    d1, _, i1, _ = ops.chamfer_nn(ta, tb, both=False)
    dk, ik = ops.knn_topk(ta, tb, 1)
    assert torch.equal(i1, ik[..., 0]) and torch.equal(d1.view(torch.int32), dk[..., 0].view(torch.int32))
    if case in ("surface", "lattice", "np256", "np1000"):
        ref = po.chamfer_nn(a, b)
        assert np.array_equal(d1.cpu().numpy(), ref[0]) and np.array_equal(i1.cpu().numpy(), ref[2])


# ---- linear ----------------------------------------------------------------------------------------
@pytest.mark.parametrize("M,K,N,act", [(512, 3, 128, 0), (300, 32, 256, 0), (1000, 128, 512, 0), (256, 128, 64, 1),
                                       (130, 64, 16, 1), (257, 160, 192, 2), (128, 3, 640, 0), (1, 16, 32, 0),
                                       (1001, 3, 128, 2), (77, 4, 32, 1), (300, 1, 8, 0), (65, 2, 1024, 0), (200000, 3, 128, 0)])
def test_linear(ops, dev, M, K, N, act):
    gen = torch.Generator().manual_seed(M + K + N)
    a = torch.randn(M, K, generator=gen)
    w = torch.randn(K, N, generator=gen) / K ** 0.5
    b = torch.randn(N, generator=gen)
    ref = a.double() @ w.double() + b.double()
    if act == 1:
        ref = torch.relu(ref)
    elif act == 2:
        ref = torch.nn.functional.leaky_relu(ref, 0.05)
    out = ops.linear(a.to(dev), w.to(dev), b.to(dev), act, 0.05).cpu()
    assert (out.double() - ref).abs().max() <= 2e-6 * max(1.0, ref.abs().max())
    # row-strided input view, no bias
    big = torch.randn(M, K + 4 * 3, generator=gen)
    out2 = ops.linear(big.to(dev)[:, 4:4 + K], w.to(dev)).cpu()
    ref2 = big[:, 4:4 + K].double() @ w.double()
    assert (out2.double() - ref2).abs().max() <= 2e-6 * max(1.0, ref2.abs().max())


@pytest.mark.parametrize("M,K,N,act", [(1000, 128, 512, 0), (300, 64, 192, 2), (129, 32, 16, 1), (128, 16, 32, 0),
                                       (5000, 128, 64, 1), (70000, 64, 128, 0)])
def test_linear_tensor_core(ops, dev, M, K, N, act):
    """pf_linear_tc (tcgen05, fp16 hi/lo split, 3 MMAs per product).  STATED TOLERANCE: 4e-6 * max|ref| against a
    float64 reference (the FFMA pf_linear is checked at 2e-6 above): fp32-class."""
    from puflow_b200.packing import pack_linear_tc
    gen = torch.Generator().manual_seed(M + K + N)
    a = torch.randn(M, K, generator=gen) * 3.0
    w = torch.randn(K, N, generator=gen) / K ** 0.5
    b = torch.randn(N, generator=gen)
    wtc = pack_linear_tc(w).to(dev)
    ref = a.double() @ w.double() + b.double()
    if act == 1:
        ref = torch.relu(ref)
    elif act == 2:
        ref = torch.nn.functional.leaky_relu(ref, 0.05)
    out = ops.linear_tc(a.to(dev), wtc, N, b.to(dev), act, 0.05).cpu()
    assert (out.double() - ref).abs().max() <= 4e-6 * max(1.0, ref.abs().max())
    # row-strided input view, no bias, output into a row-strided slice
    big = torch.randn(M, K + 8, generator=gen)
    dst = torch.zeros(M, N + 16, device=dev)
    ops.linear_tc(big.to(dev)[:, 4:4 + K], wtc, N, out=dst[:, 16:])
    ref2 = big[:, 4:4 + K].double() @ w.double()
    assert (dst[:, 16:].cpu().double() - ref2).abs().max() <= 4e-6 * max(1.0, ref2.abs().max())
    assert float(dst[:, :16].abs().max()) == 0.0


# ---- (2) EdgeConv ----------------------------------------------------------------------------------
@pytest.mark.parametrize("blk,Bp,N", [(0, 3, 256), (1, 2, 256), (3, 3, 256), (5, 1, 250), (2, 300, 64)])
def test_edgeconv_blocks_real_weights(ops, dev, packed, blk, Bp, N):
    from puflow_b200.packing import FEAT_CFG
    C, g, Co = FEAT_CFG[blk]
    gen = torch.Generator().manual_seed(blk * 1000 + N)
    x = torch.randn(Bp * N, C, generator=gen)
    p = synth.make_patches(Bp, N, seed=blk, shape="uniform")
    idx = torch.from_numpy(po.knn(p, p, 16)[1]).int()
    f = packed["feat"][blk]
    pq = emulate.linear(x, f["wpq"], f["bpq"])
    ref = emulate.edgeconv(pq, idx, f["wd"], f["wo"], N, g, Co)
    out = ops.edgeconv(pq.to(dev), idx.to(dev), f["wd"].to(dev), f["wo"].to(dev), Bp, N, g, Co).cpu()
    ok, err = parity.close_norm(out.view(Bp, N, Co).numpy(), ref.view(Bp, N, Co).numpy(), 1e-5)
    assert ok, err


@pytest.mark.parametrize("K", [8, 32])
def test_edgeconv_other_k(ops, dev, packed, K):
    """config 4 of BASELINE.json sweeps k = 8/16/32 with block-3 weights."""
    Bp, N, (C, g, Co) = 2, 512, (128, 32, 128)
    gen = torch.Generator().manual_seed(K)
    x = torch.randn(Bp * N, C, generator=gen)
    p = synth.make_patches(Bp, N, seed=K, shape="uniform")
    idx = torch.from_numpy(po.knn(p, p, K)[1]).int()
    f = packed["feat"][3]
    pq = emulate.linear(x, f["wpq"], f["bpq"])
    ref = emulate.edgeconv(pq, idx, f["wd"], f["wo"], N, g, Co)
    out = ops.edgeconv(pq.to(dev), idx.to(dev), f["wd"].to(dev), f["wo"].to(dev), Bp, N, g, Co).cpu()
    ok, err = parity.close_norm(out.view(Bp, N, Co).numpy(), ref.view(Bp, N, Co).numpy(), 1e-5)
    assert ok, err
    tc = ops.edgeconv_tc(pq.to(dev), idx.to(dev), f["wtc"].to(dev), Bp, N, g, Co).cpu()      # tcgen05 kernel, same K
    ok, err = parity.close_norm(tc.view(Bp, N, Co).numpy(), ref.view(Bp, N, Co).numpy(), 2e-5)
    assert ok, err


@pytest.mark.parametrize("K,blk,Bp,N", [(8, 3, 5, 250), (32, 3, 3, 130), (8, 0, 3, 256), (32, 1, 2, 256), (8, 1, 40, 64)])
def test_edgeconv_tensor_core_other_k(ops, dev, packed, K, blk, Bp, N):
    """pf_edgeconv_tc for K = 8 / 32 (tile = 16 / 4 points), ragged N, all three (G, C_out) shapes; 2e-5 normwise against
    the float64-accumulated emulation like the K = 16 test below."""
    from puflow_b200.packing import FEAT_CFG
    C, g, Co = FEAT_CFG[blk]
    gen = torch.Generator().manual_seed(K * 10 + blk)
    x = torch.randn(Bp * N, C, generator=gen)
    p = synth.make_patches(Bp, N, seed=K + blk, shape="uniform")
    idx = torch.from_numpy(po.knn(p, p, K)[1]).int()
    f = packed["feat"][blk]
    pq = emulate.linear(x, f["wpq"], f["bpq"])
    nref = min(Bp, 3)
    ref = emulate.edgeconv(pq[:nref * N], idx[:nref], f["wd"], f["wo"], N, g, Co)
    out = ops.edgeconv_tc(pq.to(dev), idx.to(dev), f["wtc"].to(dev), Bp, N, g, Co).cpu()
    ok, err = parity.close_norm(out[:nref * N].view(nref, N, Co).numpy(), ref.view(nref, N, Co).numpy(), 2e-5)
    assert ok, err
    ffma = ops.edgeconv(pq.to(dev), idx.to(dev), f["wd"].to(dev), f["wo"].to(dev), Bp, N, g, Co).cpu()
    ok, err = parity.close_norm(out.view(Bp, N, Co).numpy(), ffma.view(Bp, N, Co).numpy(), 2e-5)
    assert ok, err


@pytest.mark.parametrize("blk,Bp,N", [(2, 3, 256), (3, 2, 256), (5, 1, 250), (4, 37, 64), (3, 600, 256),
                                      (0, 3, 256), (1, 3, 256), (0, 1, 250), (1, 37, 64), (0, 700, 256), (1, 700, 256)])
def test_edgeconv_tensor_core(ops, dev, packed, blk, Bp, N):
    """pf_edgeconv_tc (tcgen05, fp16 hi/lo split operands, 3 MMAs per product, fp32 accumulate).
    STATED TOLERANCE: 2e-5 normwise per patch against the float64-accumulated emulation (the split drops the
    lo*lo term, 2^-22 relative per product; measured 2e-7..1e-6), i.e. the same class as the FP32 FFMA kernel,
    which is checked at 1e-5 above.  Also: bit-identical to itself on a sub-batch (tile scheduling independent)."""
    from puflow_b200.packing import FEAT_CFG
    C, g, Co = FEAT_CFG[blk]
    gen = torch.Generator().manual_seed(blk * 1000 + N)
    scale = [1.0, 1.0, 2.0, 5.0, 20.0, 90.0][blk]                       # input magnitudes seen with pu1k weights
    x = torch.randn(Bp * N, C, generator=gen) * scale
    p = synth.make_patches(Bp, N, seed=blk, shape="uniform")
    idx = torch.from_numpy(po.knn(p, p, 16)[1]).int()
    f = packed["feat"][blk]
    pq = emulate.linear(x, f["wpq"], f["bpq"])
    out = ops.edgeconv_tc(pq.to(dev), idx.to(dev), f["wtc"].to(dev), Bp, N, g, Co).cpu()
    nref = min(Bp, 3)
    ref = emulate.edgeconv(pq[:nref * N], idx[:nref], f["wd"], f["wo"], N, g, Co)
    ok, err = parity.close_norm(out[:nref * N].view(nref, N, Co).numpy(), ref.view(nref, N, Co).numpy(), 2e-5)
    assert ok, err
    ffma = ops.edgeconv(pq.to(dev), idx.to(dev), f["wd"].to(dev), f["wo"].to(dev), Bp, N, g, Co).cpu()
    ok, err = parity.close_norm(out.view(Bp, N, Co).numpy(), ffma.view(Bp, N, Co).numpy(), 2e-5)
    assert ok, err
    if Bp > 2:
        sub = ops.edgeconv_tc(pq[N:3 * N].contiguous().to(dev), idx[1:3].contiguous().to(dev), f["wtc"].to(dev), 2, N, g,
                              Co).cpu()
        assert torch.equal(sub, out[N:3 * N])


def test_edgeconv_tensor_core_rejects_other_shapes(ops, dev, packed):
    f = packed["feat"][3]
    pq = torch.zeros(256, 512, device=dev)
    idx4 = torch.zeros(1, 256, 4, dtype=torch.int32, device=dev)
    with pytest.raises(Exception):       # K outside {8, 16, 32}
        ops.edgeconv_tc(pq, idx4, f["wtc"].to(dev), 1, 256, 32, 128)
    idx16 = torch.zeros(1, 256, 16, dtype=torch.int32, device=dev)
    with pytest.raises(Exception):       # blob of another block shape
        ops.edgeconv_tc(pq, idx16, packed["feat"][0]["wtc"].to(dev), 1, 256, 32, 128)


@pytest.mark.parametrize("K", [32, 128])
def test_linear_tc_accumulation_accuracy(ops, dev, K):
    """The split-fp16 tensor-core product must be as accurate as the FP32 FFMA kernel: weights carry a 2^10 factor so no
    lo half is an fp16 subnormal, and the full-magnitude hi . hi pass is issued last because the accumulator adder
    truncates (the opposite order measured 3.5x the FFMA rms error and a -1e-6 bias on all-positive data)."""
    from puflow_b200 import packing
    g = torch.Generator().manual_seed(5)
    M, N = 8192, 128
    for positive in (False, True):
        a, w = torch.randn(M, K, generator=g), torch.randn(K, N, generator=g) * 0.03
        if positive:
            a, w = a.abs(), w.abs()
        ref = a.double() @ w.double()
        rms = ref.pow(2).mean().sqrt()
        e_f = (ops.linear(a.to(dev), w.to(dev), None, 0).cpu().double() - ref)
        e_t = (ops.linear_tc(a.to(dev), packing.pack_linear_tc(w).to(dev), N, None, 0).cpu().double() - ref)
        rf, rt = e_f.pow(2).mean().sqrt() / rms, e_t.pow(2).mean().sqrt() / rms
        assert rt <= (4.0 if positive else 1.5) * rf + 2e-8, (K, positive, float(rt), float(rf))
        assert e_t.abs().max() / rms <= 4e-6


def _split_ref(x):
    """[M, K] fp32 -> uint8 tiles in the pre-split operand layout (include/puflow_b200.h, PF_SPLIT_TILE_BYTES)."""
    M, K = x.shape
    T = (M + 127) // 128
    xp = torch.zeros(T * 128, K)
    xp[:M] = x.clamp(-65504.0, 65504.0)
    hi = xp.to(torch.float16)
    lo = (xp - hi.float()).to(torch.float16)

    def lay(h):     # [T*128, K] -> [T][K/8][128][8]
        return h.view(T, 128, K // 8, 8).permute(0, 2, 1, 3).contiguous().view(T, -1)
    return torch.cat([lay(hi), lay(lo)], 1).contiguous().view(torch.uint8).view(T, 512 * K)


@pytest.mark.parametrize("M,K,N,act", [(1000, 128, 512, 0), (300, 64, 512, 2), (129, 32, 256, 1), (128, 128, 16, 0),
                                       (40000, 128, 512, 0), (5000, 64, 48, 0)])
def test_linear_tensor_core_presplit(ops, dev, M, K, N, act):
    """pf_linear_tc_split (TMA-fed, input pre-split in the operand layout) is bit-identical to pf_linear_tc on the same
    values: same hi/lo operands, same MMA order."""
    from puflow_b200.packing import pack_linear_tc
    gen = torch.Generator().manual_seed(M + K + N)
    a = torch.randn(M, K, generator=gen) * 3
    W = torch.randn(K, N, generator=gen) / K ** 0.5
    b = torch.randn(N, generator=gen)
    wtc = pack_linear_tc(W).to(dev)
    ref = ops.linear_tc(a.to(dev), wtc, N, b.to(dev), act, 0.05)
    got = ops.linear_tc_split(_split_ref(a).to(dev), M, K, wtc, N, b.to(dev), act, 0.05)
    assert torch.equal(ref, got)
    got2 = ops.linear_tc_split(_split_ref(a).to(dev), M, K, wtc, N, None, 0)
    ref2 = ops.linear_tc(a.to(dev), wtc, N, None, 0)
    assert torch.equal(ref2, got2)


@pytest.mark.parametrize("blk,Bp,N", [(0, 3, 256), (1, 5, 256), (3, 2, 256), (4, 3, 250)])
def test_edgeconv_presplit_output(ops, dev, packed, blk, Bp, N):
    """pf_edgeconv_tc_split writes the same values twice: fp32 rows and fp16 hi/lo tiles in the operand layout."""
    from puflow_b200.packing import FEAT_CFG
    C, g, Co = FEAT_CFG[blk]
    gen = torch.Generator().manual_seed(blk)
    x = torch.randn(Bp * N, C, generator=gen)
    p = synth.make_patches(Bp, N, seed=blk, shape="uniform")
    idx = torch.from_numpy(po.knn(p, p, 16)[1]).int()
    f = packed["feat"][blk]
    pq = emulate.linear(x, f["wpq"], f["bpq"]).to(dev)
    split = ops.split_tiles(Bp * N, Co, dev)
    split.zero_()
    out = ops.edgeconv_tc(pq, idx.to(dev), f["wtc"].to(dev), Bp, N, g, Co, split=split)
    plain = ops.edgeconv_tc(pq, idx.to(dev), f["wtc"].to(dev), Bp, N, g, Co)
    assert torch.equal(out, plain)
    assert torch.equal(split.cpu(), _split_ref(out.cpu()))


# ---- (3) flow --------------------------------------------------------------------------------------
def _flow_inputs(packed, M, seed):
    gen = torch.Generator().manual_seed(seed)
    x = torch.randn(M, 3, generator=gen) * 0.5
    raw = torch.stack([emulate.linear(torch.randn(M, packed["wc"][i].shape[0], generator=gen), packed["wc"][i])
                       for i in range(6)])
    return x, raw


@pytest.mark.parametrize("M", [256, 1000, 77])
def test_flow_cond_fwd_inv(ops, dev, packed, M):
    x, raw = _flow_inputs(packed, M, M)
    sb_ref = emulate.flow_cond(raw, packed["cond"])
    sb = ops.flow_cond(raw.to(dev), packed["cond"].to(dev))
    assert (sb.cpu() - sb_ref).abs().max() <= 1e-5 * max(1.0, sb_ref.abs().max())
    z_ref, ss_ref = emulate.flow_fwd(x, raw, sb_ref, packed["flow"])
    z, ss = ops.flow_fwd(x.to(dev), raw.to(dev), sb, packed["flow"].to(dev))
    ok, err = parity.close_norm(z.cpu().numpy()[None], z_ref.numpy()[None], 1e-4)
    assert ok, err
    assert (ss.cpu() - ss_ref).abs().max() <= 1e-4 * max(1.0, ss_ref.abs().max())
    R = 4
    zin = torch.randn(M * R, 3, generator=torch.Generator().manual_seed(1)) * 0.5
    xi_ref = emulate.flow_inv(zin, raw, sb_ref, packed["flow"], R)
    xi = ops.flow_inv(zin.to(dev), raw.to(dev), sb, packed["flow"].to(dev), R)
    ok, err = parity.close_norm(xi.cpu().numpy()[None], xi_ref.numpy()[None], 1e-4)
    assert ok, err
    # round trip: inverse(forward(x)) == x  (R = 1)
    back = ops.flow_inv(z, raw.to(dev), sb, packed["flow"].to(dev), 1).cpu()
    assert (back - x).abs().max() < 1e-4      # random conditioning features: exp(s) is far larger than with real cs


def test_flow_blocks_golden(ops, dev, packed, golden):
    """Single FlowBlock forward / inverse against the unmodified reference (flows_pu1k.npz)."""
    fl = golden("flows_pu1k.npz")
    for bi in (0, 1):
        x = torch.from_numpy(fl[f"b{bi}_x"]).reshape(-1, 3)
        c = torch.from_numpy(fl[f"b{bi}_c"]).reshape(x.shape[0], -1)
        raw = ops.linear(c.to(dev), packed["wc"][bi].to(dev))[None].contiguous()
        cw, fw = packed["cond"][bi:bi + 1].to(dev), packed["flow"][bi:bi + 1].to(dev)
        sb = ops.flow_cond(raw, cw)
        y, _ = ops.flow_fwd(x.to(dev), raw, sb, fw)
        ok, err = parity.close_norm(y.cpu().view(2, 64, 3).numpy(), fl[f"b{bi}_block_fwd"], 1e-4)
        assert ok, (bi, err)
        xi = ops.flow_inv(x.to(dev), raw, sb, fw, 1)
        ok, err = parity.close_norm(xi.cpu().view(2, 64, 3).numpy(), fl[f"b{bi}_block_inv"], 1e-4)
        assert ok, (bi, err)


@pytest.mark.parametrize("M", [256, 1000, 77, 60000])
def test_flow_tensor_core(ops, dev, packed, M):
    """tcgen05 flow kernels (fp16 hi/lo split, fp32 accumulation).  Stated tolerance: the same 1e-4 normwise bound as
    the FP32 kernels against the float64-accumulated emulation (measured ~1e-6), 1e-5 on sb."""
    x, raw = _flow_inputs(packed, M, M)
    rawd, ctc, ftc = raw.to(dev), packed["cond_tc"].to(dev), packed["flow_tc"].to(dev)
    if M <= 1000:
        sb_ref = emulate.flow_cond(raw, packed["cond"])
        z_ref, ss_ref = emulate.flow_fwd(x, raw, sb_ref, packed["flow"])
    else:       # large case: the FP32 FFMA kernels (checked against the emulation above) are the reference
        sbf = ops.flow_cond(rawd, packed["cond"].to(dev))
        zf, ssf = ops.flow_fwd(x.to(dev), rawd, sbf, packed["flow"].to(dev))
        sb_ref, z_ref, ss_ref = sbf.cpu(), zf.cpu(), ssf.cpu()
    sb = ops.flow_cond(rawd, ctc)
    assert (sb.cpu() - sb_ref).abs().max() <= 1e-5 * max(1.0, sb_ref.abs().max())
    assert (sb.cpu()[..., 6:] == 0).all()
    z, ss = ops.flow_fwd(x.to(dev), rawd, sb, ftc)
    ok, err = parity.close_norm(z.cpu().numpy()[None], z_ref.numpy()[None], 1e-4)
    assert ok, err
    assert (ss.cpu() - ss_ref).abs().max() <= 1e-4 * max(1.0, ss_ref.abs().max())
    R = 4
    zin = torch.randn(M * R, 3, generator=torch.Generator().manual_seed(1)) * 0.5
    if M <= 1000:
        xi_ref = emulate.flow_inv(zin, raw, sb_ref, packed["flow"], R)
    else:
        xi_ref = ops.flow_inv(zin.to(dev), rawd, sb_ref.to(dev), packed["flow"].to(dev), R).cpu()
    xi = ops.flow_inv(zin.to(dev), rawd, sb, ftc, R)
    ok, err = parity.close_norm(xi.cpu().numpy()[None], xi_ref.numpy()[None], 1e-4)
    assert ok, err
    # round trip; random conditioning features give exp(s) up to e^40, so a few points in 60,000 amplify rounding
    back = (ops.flow_inv(z, rawd, sb, ftc, 1).cpu() - x).abs().flatten()
    assert back.quantile(0.99) < 1e-5 and back.quantile(0.999) < 3e-5 and back.max() < (1e-4 if M <= 1000 else 2e-3)


def test_flow_tensor_core_golden_and_subsets(ops, dev, packed, golden):
    """Single FlowBlock forward / inverse on the tensor-core kernels against the unmodified reference."""
    fl = golden("flows_pu1k.npz")
    for bi in (0, 1):
        x = torch.from_numpy(fl[f"b{bi}_x"]).reshape(-1, 3)
        c = torch.from_numpy(fl[f"b{bi}_c"]).reshape(x.shape[0], -1)
        raw = ops.linear(c.to(dev), packed["wc"][bi].to(dev))[None].contiguous()
        cw, fw = packed["cond_tc"][bi:bi + 1].to(dev), packed["flow_tc"][bi:bi + 1].to(dev)
        sb = ops.flow_cond(raw, cw)
        y, _ = ops.flow_fwd(x.to(dev), raw, sb, fw)
        ok, err = parity.close_norm(y.cpu().view(2, 64, 3).numpy(), fl[f"b{bi}_block_fwd"], 1e-4)
        assert ok, (bi, err)
        xi = ops.flow_inv(x.to(dev), raw, sb, fw, 1)
        ok, err = parity.close_norm(xi.cpu().view(2, 64, 3).numpy(), fl[f"b{bi}_block_inv"], 1e-4)
        assert ok, (bi, err)
    # empty batch
    e = ops.flow_inv(torch.empty(0, 3, device=dev), torch.empty(6, 0, 192, device=dev), torch.empty(6, 0, 8, device=dev),
                     packed["flow_tc"].to(dev), 4)
    assert e.shape == (0, 3)


@pytest.mark.parametrize("blk,M", [(0, 300), (1, 128), (2, 1000), (5, 20000)])
def test_cond_chain_tensor_core(ops, dev, packed, blk, M):
    """Fused FeatMergeUnit + conditioning first layers + coupling2 nets (pf_cond_chain_tc) against the un-fused,
    un-composed float64 chain.  Stated tolerance 1e-4 normwise on raw1, 1e-5 * max(1, |sb|) on sb (measured ~1e-6)."""
    from puflow_b200.packing import FEAT_CFG
    C = FEAT_CFG[blk][2]
    gen = torch.Generator().manual_seed(blk * 7 + 1)
    c = torch.randn(M, C, generator=gen).abs() * 0.7                       # EdgeConv outputs are max-pooled, mostly > 0
    m = packed["merge"][blk]
    h = emulate.linear(c.double(), m["w1"].double(), m["b1"].double(), "relu")
    cs = emulate.linear(h, m["w2"].double())
    raw = emulate.linear(cs, packed["wc"][blk].double()).float()
    sb_ref = emulate.flow_cond(raw[None], packed["cond"][blk:blk + 1])[0]
    raw1 = torch.empty(M, 64, device=dev)
    sb = torch.empty(M, 8, device=dev)
    ops.cond_chain(c.to(dev), packed["chain_tc"][blk].to(dev), raw1, sb)
    ok, err = parity.close_norm(raw1.cpu().numpy()[None], raw[:, :64].numpy()[None], 1e-4)
    assert ok, err
    assert (sb.cpu() - sb_ref).abs().max() <= 1e-5 * max(1.0, sb_ref.abs().max())
    assert (sb.cpu()[:, 6:] == 0).all()
    # the compact raw1 layout drives the flow kernels exactly like the 192-wide one
    x = torch.randn(M, 3, generator=gen) * 0.5
    raw_full = torch.zeros(1, M, 192, device=dev)
    raw_full[0, :, :64] = raw1
    ftc = packed["flow_tc"][blk:blk + 1].to(dev)
    za, _ = ops.flow_fwd(x.to(dev), raw_full, sb[None].contiguous(), ftc)
    zb, _ = ops.flow_fwd(x.to(dev), raw1[None].contiguous(), sb[None].contiguous(), ftc)
    assert torch.equal(za, zb)


# ---- (4) interpolation -----------------------------------------------------------------------------
@pytest.mark.parametrize("Bp,N,R", [(3, 256, 4), (2, 100, 2), (1, 256, 16)])
def test_interp(ops, dev, packed, Bp, N, R):
    p = synth.make_patches(Bp, N, seed=N + R, shape="surface")
    xyz = torch.from_numpy(p)
    z = torch.randn(Bp, N, 3, generator=torch.Generator().manual_seed(R)) * 0.6
    idx = torch.from_numpy(po.knn(p, p, 16)[1]).int()
    wi = packed["interp"]
    pq = emulate.linear(xyz.reshape(-1, 3), wi["wpq"], wi["bpq"])
    ref, _ = emulate.interp(xyz, z, idx, pq, wi["blob"], R)
    out = ops.interp(xyz.to(dev), z.to(dev), idx.to(dev), pq.to(dev), wi["blob"].to(dev), R).cpu()
    ok, err = parity.close_norm(out.numpy(), ref.numpy(), 1e-4)
    assert ok, err


@pytest.mark.parametrize("Bp,N,R", [(3, 256, 4), (2, 100, 2), (1, 256, 16), (40, 256, 4), (1, 64, 32)])
def test_interp_tensor_core(ops, dev, packed, Bp, N, R):
    """pf_interp_tc (tcgen05, fp16 hi/lo split, unfolded geo weights).  STATED TOLERANCE: 1e-4 normwise per patch
    against the float64-accumulated emulation of the FFMA formulation (same bound as pf_interp above)."""
    p = synth.make_patches(Bp, N, seed=N + R, shape="surface")
    xyz = torch.from_numpy(p)
    z = torch.randn(Bp, N, 3, generator=torch.Generator().manual_seed(R)) * 0.6
    idx = torch.from_numpy(po.knn(p, p, 16)[1]).int()
    wi = packed["interp"]
    nref = min(Bp, 3)
    pq = emulate.linear(xyz[:nref].reshape(-1, 3), wi["wpq"], wi["bpq"])
    ref, _ = emulate.interp(xyz[:nref], z[:nref], idx[:nref], pq, wi["blob"], R)
    out = ops.interp_tc(xyz.to(dev), z.to(dev), idx.to(dev), wi["tc"].to(dev), R).cpu()
    ok, err = parity.close_norm(out[:nref].numpy(), ref.numpy(), 1e-4)
    assert ok, err
    assert torch.isfinite(out).all()
    if Bp > 2:
        sub = ops.interp_tc(xyz[1:3].contiguous().to(dev), z[1:3].contiguous().to(dev), idx[1:3].contiguous().to(dev),
                            wi["tc"].to(dev), R).cpu()
        assert torch.equal(sub, out[1:3])


def test_interp_tensor_core_tile_pipeline_is_deterministic(ops, dev, packed):
    """The interpolation kernel overlaps a tile's tail with the next tile's head and schedules its side work by polling
    (csrc/interp_tc.cu): a CTA that runs MANY tiles must produce, bit for bit, what one-tile-per-CTA launches produce, run after
    run.  (A scheduling bug of an early version only showed in the second tile of a CTA, and only as run-to-run differences.)"""
    Bp, N, R = 96, 256, 4                                  # 1,536 tiles on <= 148 CTAs: 10-11 tiles per CTA
    p = synth.make_patches(Bp, N, seed=5, shape="surface")
    xyz = torch.from_numpy(p).to(dev)
    z = (torch.randn(Bp, N, 3, generator=torch.Generator().manual_seed(1)) * 0.6).to(dev)
    idx = torch.from_numpy(po.knn(p, p, 16)[1]).int().to(dev)
    wtc = packed["interp"]["tc"].to(dev)
    first = ops.interp_tc(xyz, z, idx, wtc, R)
    for _ in range(5):
        assert torch.equal(ops.interp_tc(xyz, z, idx, wtc, R), first)
    for b in (0, 1, 37, 95):                               # 16 tiles each: one tile per CTA, no cross-tile overlap
        one = ops.interp_tc(xyz[b:b + 1].contiguous(), z[b:b + 1].contiguous(), idx[b:b + 1].contiguous(), wtc, R)
        assert torch.equal(one[0], first[b])


# ---- whole path ------------------------------------------------------------------------------------
def test_network_stages_match_reference(dev, state_dict, golden):
    from puflow_b200.engine import PUFlowEngine
    g = golden("net_pu1k.npz")
    eng = PUFlowEngine(state_dict, dev)
    r = eng.forward(torch.from_numpy(g["xyz"]).to(dev), 4, return_all=True)
    assert np.array_equal(r["knn_idx"].cpu().numpy(), g["knn_idx"])
    for i in range(6):
        ok, err = parity.close_norm(r["feats"][i].cpu().numpy(), g[f"feat{i}"], 1e-4)
        assert ok, ("feat", i, err)
        ok, err = parity.close_norm(r["cs"][i].cpu().numpy(), g[f"cs{i}"], 1e-4)
        assert ok, ("cs", i, err)
    ok, err = parity.close_norm(r["z"].cpu().numpy(), g["z"], 1e-3)
    assert ok, ("z", err)
    fz = r["zup"].cpu().view(3, 256, 4, 3).permute(0, 1, 3, 2).numpy()
    ok, err = parity.close_norm(fz, g["fz"], 1e-3)
    assert ok, ("fz", err)
    ok, err = parity.close_norm(r["x"].cpu().numpy(), g["x"], 1e-4)
    assert ok, ("x", err)
    assert abs(float(r["logp_x"]) - float(g["logp_x"][0])) <= 1e-4 * abs(float(g["logp_x"][0]))
    for b in range(3):
        assert parity.chamfer_distance(r["x"][b].cpu().numpy(), g["x"][b]) <= 1e-5


@pytest.mark.parametrize("tag", ["pugan", "pugeo"])
def test_network_other_checkpoints(dev, golden, tag):
    """Same stage-by-stage comparison with the reference's pugan / pugeo weight sets (activations up to ~400 with pugeo:
    the fp16 hi/lo operands of the tensor-core kernels stay far from saturation, and the tolerances are unchanged)."""
    import os
    from puflow_b200.engine import PUFlowEngine
    sd = torch.load(os.path.join(os.path.dirname(__file__), "golden", f"puflow-x4-{tag}.pt"), map_location="cpu")
    g = golden(f"net_{tag}.npz")
    r = PUFlowEngine(sd, dev).forward(torch.from_numpy(g["xyz"]).to(dev), 4, return_all=True)
    assert np.array_equal(r["knn_idx"].cpu().numpy(), g["knn_idx"])
    for i in range(6):
        ok, err = parity.close_norm(r["feats"][i].cpu().numpy(), g[f"feat{i}"], 1e-4)
        assert ok, ("feat", i, err)
        ok, err = parity.close_norm(r["cs"][i].cpu().numpy(), g[f"cs{i}"], 1e-4)
        assert ok, ("cs", i, err)
    ok, err = parity.close_norm(r["z"].cpu().numpy(), g["z"], 1e-3)
    assert ok, ("z", err)
    ok, err = parity.close_norm(r["x"].cpu().numpy(), g["x"], 1e-4)
    assert ok, ("x", err)
    assert abs(float(r["logp_x"]) - float(g["logp_x"][0])) <= 1e-4 * abs(float(g["logp_x"][0]))
    for b in range(3):
        assert parity.chamfer_distance(r["x"][b].cpu().numpy(), g["x"][b]) <= 1e-5


def test_network_ffma_and_tensor_core_paths_agree(dev, state_dict, golden):
    """The FP32-FFMA kernels (PUFlowEngine(kernels="ffma")) and the default tensor-core path both meet the reference."""
    from puflow_b200.engine import PUFlowEngine
    g = golden("net_pu1k.npz")
    xyz = torch.from_numpy(g["xyz"]).to(dev)
    x_tc = PUFlowEngine(state_dict, dev).forward(xyz, 4)[0]
    eng = PUFlowEngine(state_dict, dev, kernels="ffma")
    assert not eng.use_tc
    x_ff = eng.forward(xyz, 4)[0]
    for x in (x_tc, x_ff):
        ok, err = parity.close_norm(x.cpu().numpy(), g["x"], 1e-4)
        assert ok, err
    ok, err = parity.close_norm(x_tc.cpu().numpy(), x_ff.cpu().numpy(), 1e-4)
    assert ok, err


def test_network_vs_oracle_one_cloud(dev, state_dict):
    """32 real kNN-256 patches of one synthetic 2,048-pt cloud (config 1 of BASELINE.json)."""
    from puflow_b200.engine import PUFlowEngine
    pc = torch.from_numpy(synth.make_cloud(2048, 2021, "torus"))[None]
    pcn, _, _ = po.normalize_pc(pc)
    seeds = torch.from_numpy(po.fps(pcn.numpy(), 32)).long()
    _, pidx = po.knn(pcn[0][seeds[0]][None].numpy(), pcn.numpy(), 256)
    patches, _, _ = po.normalize_pc(pcn[0][torch.from_numpy(pidx[0])])
    ref = po.PUFlowOracle(state_dict).forward(patches, 4, return_all=True)
    eng = PUFlowEngine(state_dict, dev)
    r = eng.forward(patches.to(dev), 4, return_all=True)
    assert np.array_equal(r["knn_idx"].cpu().numpy().astype(np.int64), ref["knn_idx"].numpy())
    ok, err = parity.close_norm(r["x"].cpu().numpy(), ref["x"].numpy(), 1e-4)
    assert ok, err
    assert abs(float(r["logp_x"]) - float(ref["logp_x"])) <= 1e-4 * abs(float(ref["logp_x"]))


def test_full_size_properties(dev, state_dict):
    """BASELINE.json config 2 size (4,096 patches): size-independent properties instead of the oracle."""
    from puflow_b200.engine import PUFlowEngine
    Bp = 4096
    p = synth.make_patches(Bp, 256, seed=2021, shape="surface")
    xyz = torch.from_numpy(p).to(dev)
    eng = PUFlowEngine(state_dict, dev)
    r = eng.forward(xyz, 4, return_all=True)
    idx = r["knn_idx"]
    assert (idx[:, :, 0] == torch.arange(256, device=dev)[None]).all()           # self is the nearest
    assert int(idx.min()) >= 0 and int(idx.max()) < 256
    assert torch.isfinite(r["x"]).all() and r["x"].shape == (Bp, 1024, 3)
    # the batch is processed patch by patch: a slice of the batch gives bit-identical results
    sub = eng.forward(xyz[1000:1003].contiguous(), 4)[0]
    assert torch.equal(sub, r["x"][1000:1003])
    # oracle on 2 of the 4,096 patches
    ref = po.PUFlowOracle(state_dict).forward(torch.from_numpy(p[[7, 4000]]), 4)[0]
    ok, err = parity.close_norm(r["x"][[7, 4000]].cpu().numpy(), ref.numpy(), 1e-4)
    assert ok, err
    # flow round trip at full size: g(f(x)) with R = 1 returns the input
    cond = eng.conditioning(r["cs"])
    back = eng.g(r["z"], cond, 1)
    assert (back - xyz).abs().max() < 5e-5


# ---- caller side -----------------------------------------------------------------------------------
@pytest.mark.parametrize("B,n,npoint", [(3, 2048, 32), (2, 5000, 78), (2, 40960, 700), (1, 12345, 300), (4, 100, 100),
                                        (40, 2048, 32), (64, 700, 50), (40, 3000, 40), (130, 33, 33)])   # B >= 38: one CTA per cloud
def test_fps_bit_exact(ops, dev, B, n, npoint):
    rng = np.random.default_rng(n)
    p = rng.uniform(-1, 1, (B, n, 3)).astype(np.float32)
    ref = po.fps(p, npoint)
    out = ops.fps(torch.from_numpy(p).to(dev), npoint).cpu().numpy()
    assert np.array_equal(out, ref)


def test_fps_points_at_the_centroid_follow_fps_py(ops, dev):
    """Semantics pin: this FPS follows the reference's OWN modules/utils/fps.py:40-73, under which points at /
    near the origin are ordinary candidates.  Upstream pointnet2_ops.furthest_point_sample (un-vendored; what patch.py:102,156
    imports) additionally never selects candidates with |p|^2 <= 1e-3 -- a documented deviation (DESIGN.md section 2;
    tools/fps_near_origin.py: 0 such points in the surface clouds of the bench, 1 of 2,048 in a volumetric cloud, none
    selected).  Here: a centred cloud with a cluster of 40 points inside radius 0.03 of the centroid; every mode returns the
    oracle's indices, which DO include near-origin points."""
    rng = np.random.default_rng(3)
    shell = rng.standard_normal((4056, 3))
    shell /= np.linalg.norm(shell, axis=1, keepdims=True)
    core = rng.uniform(-0.015, 0.015, (40, 3))
    p = np.concatenate([core, shell], 0)[rng.permutation(4096)].astype(np.float32)[None]
    ref = po.fps(p, 600)
    assert ((p[0, ref[0]] ** 2).sum(-1) <= 1e-3).sum() >= 1
    x = torch.from_numpy(p).to(dev)
    for mode in ("brute", "bucket", "cluster"):
        assert np.array_equal(ops.fps(x, 600, mode=mode).cpu().numpy(), ref)


def test_fps_pointnet2_near_origin_rule_is_available(ops, dev):
    """The opt-in: ops.fps(skip_near_origin=1e-3) (pf_fps_park_near_origin + the same kernels) selects what an FPS that skips
    candidates with |p|^2 <= 1e-3 selects -- upstream pointnet2_ops' rule as restated in oracle.fps_pointnet2 (unpinned: the
    package is not in the reference tree).  Same centred cloud as above: no near-origin point is ever chosen, every mode
    agrees, and the default (fps.py) result differs.  modules.utils.patch.FPS_SKIP_NEAR_ORIGIN routes PatchHelper through it."""
    rng = np.random.default_rng(3)
    shell = rng.standard_normal((4056, 3))
    shell /= np.linalg.norm(shell, axis=1, keepdims=True)
    core = rng.uniform(-0.015, 0.015, (40, 3))
    p = np.concatenate([core, shell], 0)[rng.permutation(4096)].astype(np.float32)[None]
    p = np.concatenate([p, p[:, ::-1]], 0).copy()                      # a second cloud: another start point
    ref = po.fps_pointnet2(p, 600)
    x = torch.from_numpy(p).to(dev)
    for mode in ("brute", "bucket", "cluster", "queue"):
        got = ops.fps(x, 600, mode=mode, skip_near_origin=1e-3).cpu().numpy()
        assert np.array_equal(got, ref)
        assert (((p[np.arange(2)[:, None], got] ** 2).sum(-1) <= 1e-3)[:, 1:]).sum() == 0
    assert not np.array_equal(ops.fps(x, 600).cpu().numpy(), ref)
    from modules.utils import patch
    assert patch.FPS_SKIP_NEAR_ORIGIN == 0.0                          # the default stays the reference's own fps.py
    patch.FPS_SKIP_NEAR_ORIGIN = 1e-3
    try:
        assert np.array_equal(patch.furthest_point_sample(x, 600).cpu().numpy(), ref)
    finally:
        patch.FPS_SKIP_NEAR_ORIGIN = 0.0


def test_fps_lattice_ties(ops, dev):
    p = np.stack([synth.make_cloud(1000, s, "lattice") for s in (1, 2)])
    assert np.array_equal(ops.fps(torch.from_numpy(p).to(dev), 200).cpu().numpy(), po.fps(p, 200))


@pytest.mark.parametrize("B,n,npoint,kind", [(3, 2048, 32, "uniform"), (2, 5000, 600, "surface"), (2, 40960, 1500, "surface"),
                                             (1, 12345, 700, "uniform"), (4, 100, 100, "uniform"), (2, 1000, 400, "lattice"),
                                             (1, 99840, 400, "surface"), (2, 3000, 3000, "dup"),
                                             (1, 41000, 300, "uniform"), (1, 40929, 300, "uniform")])   # either side of the
                                             # register-resident / shared-memory metadata switch (64 buckets per warp)
def test_fps_bucketed_bit_exact(ops, dev, B, n, npoint, kind):
    """The spatially pruned FPS (pf_fps_bucketed) returns exactly the indices of the brute-force definition
    (oracle C restatement of modules/utils/fps.py:40-73), including exact ties (lattice, duplicated points) and
    npoint == n, and agrees with the brute-force CUDA kernel."""
    rng = np.random.default_rng(n + npoint)
    if kind == "uniform":
        p = rng.uniform(-1, 1, (B, n, 3)).astype(np.float32)
    elif kind == "surface":
        p = np.stack([synth.make_cloud(n, 11 + b, synth.SHAPES[b % 3]) for b in range(B)])
    elif kind == "lattice":
        p = np.stack([synth.make_cloud(n, 5 + b, "lattice") for b in range(B)])
    else:   # every point appears three times
        base = rng.uniform(-1, 1, (B, n // 3, 3)).astype(np.float32)
        p = np.concatenate([base, base, base], 1)[:, rng.permutation(n // 3 * 3)]
        n = p.shape[1]
        npoint = min(npoint, n)
    ref = po.fps(p, npoint)
    x = torch.from_numpy(np.ascontiguousarray(p)).to(dev)
    out = ops.fps(x, npoint, mode="bucket").cpu().numpy()
    assert np.array_equal(out, ref)
    if n <= 102400:
        assert np.array_equal(ops.fps(x, npoint, mode="brute").cpu().numpy(), ref)
    assert np.array_equal(ops.fps(x, npoint, mode="cluster").cpu().numpy(), ref)      # buckets spread over a CTA cluster
    assert np.array_equal(ops.fps(x, npoint, mode="queue").cpu().numpy(), ref)        # threshold queue, one warp runs the chain


@pytest.mark.parametrize("B,n,npoint,kind", [(1, 99840, 2500, "surface"), (8, 99840, 300, "surface"), (1, 40960, 8216, "surface"),
                                             (3, 20000, 900, "uniform"), (1, 131000, 300, "uniform"), (2, 9000, 9000, "dup"),
                                             (2, 8000, 500, "lattice"), (1, 30000, 200, "blob"), (1, 6000, 100, "same")])
def test_fps_clustered_bit_exact(ops, dev, B, n, npoint, kind):
    """pf_fps_clustered (one thread-block cluster per cloud: DSMEM candidate exchange, remote mbarrier arrives) against the
    brute-force definition: the config-3 merge size, 8 clouds at once, a full 8,216-round run, ties (lattice, triplicated
    points), and degenerate clouds whose cells overflow a CTA's capacity ("blob": 90 % of the points in one grid cell,
    "same": all points identical) -- these take the in-stream fallback to the single-CTA kernel."""
    rng = np.random.default_rng(n + npoint)
    if kind == "uniform":
        p = rng.uniform(-1, 1, (B, n, 3)).astype(np.float32)
    elif kind == "surface":
        p = np.stack([synth.make_cloud(n, 21 + b, synth.SHAPES[b % 3]) for b in range(B)])
    elif kind == "lattice":
        p = np.stack([synth.make_cloud(n, 5 + b, "lattice") for b in range(B)])
    elif kind == "dup":
        base = rng.uniform(-1, 1, (B, n // 3, 3)).astype(np.float32)
        p = np.concatenate([base, base, base], 1)[:, rng.permutation(n // 3 * 3)]
        npoint = min(npoint, p.shape[1])
    elif kind == "blob":
        p = rng.uniform(-1, 1, (B, n, 3)).astype(np.float32)
        p[:, : n * 9 // 10] = p[:, : n * 9 // 10] * 1e-3 + 0.5
    else:
        p = np.full((B, n, 3), 0.25, np.float32)
    ref = po.fps(p, npoint)
    x = torch.from_numpy(np.ascontiguousarray(p)).to(dev)
    assert np.array_equal(ops.fps(x, npoint, mode="cluster").cpu().numpy(), ref)
    assert np.array_equal(ops.fps(x, npoint, mode="queue").cpu().numpy(), ref)


def test_fps_queue_tie_search(ops, dev):
    """pf_fps_queued when more points tie with the maximum than the queue holds (400 exact copies of each of six far
    points): the block-wide tie search draws the lowest original index, like the brute-force definition."""
    rng = np.random.default_rng(7)
    B, n, npoint = 2, 6000, 3000
    p = rng.uniform(-1, 1, (B, n, 3)).astype(np.float32)
    for j in range(6):
        p[:, 400 * j:400 * (j + 1)] = rng.uniform(2, 3, (B, 1, 3)).astype(np.float32) * (1 if j % 2 else -1)
    p = np.ascontiguousarray(p[:, rng.permutation(n)])
    ref = po.fps(p, npoint)
    assert np.array_equal(ops.fps(torch.from_numpy(p).to(dev), npoint, mode="queue").cpu().numpy(), ref)


@pytest.mark.parametrize("kind,B,n,npoint", [("surface", 1, 99840, 20024), ("surface", 3, 70001, 1500), ("lattice", 2, 80000, 600),
                                             ("ties", 2, 66000, 3000), ("dup", 1, 90000, 2500), ("same", 2, 66000, 50)])
def test_fps_cluster_queue_bit_exact(ops, dev, kind, B, n, npoint):
    """The threshold queue with a cloud's buckets spread over an 8-CTA cluster (fps_cqueue_kernel: what pf_fps_queued runs for few
    clouds of >= 65,536 candidates) against the brute-force definition: config 3's full merge (20,024 of 99,840), a ragged size,
    lattice ties, more points tying with the maximum than the queue holds (the cluster-wide tie search), triplicated points, all
    points identical."""
    rng = np.random.default_rng(n + npoint)
    if kind == "surface":
        p = np.stack([synth.make_cloud(n, 41 + b, synth.SHAPES[b % 3]) for b in range(B)])
    elif kind == "lattice":
        p = rng.integers(0, 40, (B, n, 3)).astype(np.float32)
    elif kind == "ties":
        p = rng.uniform(-1, 1, (B, n, 3)).astype(np.float32)
        for j in range(6):
            p[:, 400 * j:400 * (j + 1)] = rng.uniform(2, 3, (B, 1, 3)).astype(np.float32) * (1 if j % 2 else -1)
        p = np.ascontiguousarray(p[:, rng.permutation(n)])
    elif kind == "dup":
        base = rng.uniform(-1, 1, (B, n // 3, 3)).astype(np.float32)
        p = np.concatenate([base, base, base], 1)[:, rng.permutation(n // 3 * 3)]
    else:
        p = np.full((B, n, 3), 0.25, np.float32)
    ref = po.fps(p, npoint)
    out = ops.fps(torch.from_numpy(np.ascontiguousarray(p)).to(dev), npoint, mode="queue").cpu().numpy()
    assert np.array_equal(out, ref)


@pytest.mark.parametrize("B,nq,np_,k", [(2, 32, 2048, 256), (1, 78, 5000, 256), (2, 5, 300, 256), (1, 3, 1000, 1000)])
def test_knn_patch_bit_exact(ops, dev, B, nq, np_, k):
    rng = np.random.default_rng(np_)
    p = rng.uniform(-1, 1, (B, np_, 3)).astype(np.float32)
    q = p[:, :nq].copy()
    ref = po.knn(q, p, k)[1]
    out = ops.knn_patch(torch.from_numpy(q).to(dev), torch.from_numpy(p).to(dev), k).cpu().numpy()
    assert np.array_equal(out.astype(np.int64), ref)


@pytest.mark.parametrize("kind", ["lattice", "duplicates", "identical"])
def test_knn_patch_ties(ops, dev, kind):
    """The select-then-sort path keeps the (distance, index) order under ties at the k-th distance; more than 1,024 keys at or
    below it (every point identical) take the full-sort fallback."""
    rng = np.random.default_rng(17)
    B, n, nq, k = 2, 2048, 5, 256
    if kind == "lattice":
        p = rng.integers(0, 8, (B, n, 3)).astype(np.float32)
    elif kind == "duplicates":
        p = np.repeat(rng.uniform(-1, 1, (B, n // 4, 3)), 4, axis=1).astype(np.float32)
    else:
        p = np.full((B, n, 3), 0.25, np.float32)
    q = p[:, :nq].copy()
    ref = po.knn(q, p, k)[1]
    out = ops.knn_patch(torch.from_numpy(q).to(dev), torch.from_numpy(p).to(dev), k).cpu().numpy()
    assert np.array_equal(out.astype(np.int64), ref)
