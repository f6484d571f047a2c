"""Reference harness — TEST INFRASTRUCTURE, runs ONLY in the build container.

Imports the *unmodified* reference modules from /root/reference (read-only) in THIS interpreter and
runs them on CPU, to (a) generate the golden fixtures committed under tests/golden/ and (b) time the
reference's CPU path for orientation.  Nothing under puflow_b200/ or modules/ imports this file, and
it never runs on the GPU box (/root/reference does not exist there).

Why a separate interpreter: the drop-in API means this repo also has a top-level `modules` package;
the reference uses absolute imports (`from modules.flows.coupling import ...`,
/root/reference/modules/discrete/interpflow.py:12-17).  This script therefore must be started with
/root/repo NOT on sys.path (run it by file path: `python oracle/ref/run_reference.py golden`).

The reference's heavy primitives live in un-vendored third-party CUDA packages (pytorch3d, knn_cuda,
pointnet2_ops; /root/reference/docker/Dockerfile:47-50).  Stand-ins with the semantics SURVEY.md
§8(c) fixes are injected into sys.modules before the reference is imported:

  pytorch3d.ops.knn_points   squared L2 by direct differences ((dx*dx+dy*dy)+dz*dz) in fp32, K
                             smallest by lexicographic (distance, index), sorted ascending, int64
  pytorch3d.ops.knn_gather   x[b, idx]
  knn_cuda.KNN               sqrt distances ascending, ties -> lower index, [B,k,M] layout
  pointnet2_ops.pointnet2_utils.furthest_point_sample
                             = the reference's own modules/utils/fps.py:40-73 (starts at index 0)
  pointnet2_ops.pointnet2_utils.gather_operation   torch.gather on [B,C,N]
"""
import argparse
import importlib.util
import os
import sys
import time
import types

import numpy as np
import torch

REF = os.environ.get("PUFLOW_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))


def _load_synth():
    spec = importlib.util.spec_from_file_location("pf_synth", os.path.join(REPO, "tools", "synth.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


# ------------------------------------------------------------------------------------------------
# stand-ins for the un-vendored third-party packages
def _sqdist(p1, p2):
    d = p1[:, :, None, :] - p2[:, None, :, :]
    return (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]


def knn_points(p1, p2, K, return_nn=False, return_sorted=True, **kw):
    d = _sqdist(p1, p2)
    dist, idx = torch.sort(d, dim=-1, stable=True)
    return dist[..., :K].contiguous(), idx[..., :K].contiguous(), None


def knn_gather(x, idx):
    B = x.shape[0]
    return x[torch.arange(B)[:, None, None], idx]


class KNN:
    def __init__(self, k, transpose_mode=False):
        assert not transpose_mode
        self.k = k

    def __call__(self, ref, query):
        d = _sqdist(query.transpose(1, 2), ref.transpose(1, 2))        # [B, M, N]
        dist, idx = torch.sort(d, dim=-1, stable=True)
        return dist[..., :self.k].sqrt().transpose(1, 2).contiguous(), idx[..., :self.k].transpose(1, 2).contiguous()


def gather_operation(feat, idx):
    return torch.gather(feat, 2, idx.long()[:, None, :].expand(-1, feat.shape[1], -1))


def install_shims():
    sys.path.insert(0, REF)
    sys.path.insert(1, os.path.join(REF, "metric", "PyTorchCD"))
    assert not any(os.path.abspath(p) == REPO for p in sys.path if p), "repo root must not be on sys.path"
    p3d, ops = types.ModuleType("pytorch3d"), types.ModuleType("pytorch3d.ops")
    ops.knn_points, ops.knn_gather = knn_points, knn_gather
    ops.sample_farthest_points = lambda *a, **k: (_ for _ in ()).throw(NotImplementedError())
    p3d.ops = ops
    sys.modules["pytorch3d"], sys.modules["pytorch3d.ops"] = p3d, ops
    kc = types.ModuleType("knn_cuda")
    kc.KNN = KNN
    sys.modules["knn_cuda"] = kc
    from modules.utils.fps import farthest_point_sampling
    pn, pu = types.ModuleType("pointnet2_ops"), types.ModuleType("pointnet2_ops.pointnet2_utils")
    pu.furthest_point_sample = lambda xyz, n: farthest_point_sampling(xyz, n).int()
    pu.gather_operation = gather_operation
    pn.pointnet2_utils = pu
    sys.modules["pointnet2_ops"], sys.modules["pointnet2_ops.pointnet2_utils"] = pn, pu
    # GaussianDistribution hard-codes device='cuda:0' (modules/utils/probs.py:51,56)
    import modules.utils.probs as probs
    orig = probs.GaussianDistribution.__init__

    def patched(self, pc_channel, mu, vars, temperature=1.0, device="cpu"):
        orig(self, pc_channel, mu, vars, temperature, device="cpu")
    probs.GaussianDistribution.__init__ = patched


def load_network(name="puflow-x4-pu1k.pt"):
    from modules.discrete.interpflow import PointInterpFlow
    net = PointInterpFlow(3)
    sd = torch.load(os.path.join(REF, "pretrain", name), map_location="cpu")
    missing = net.load_state_dict(sd)
    assert not missing.missing_keys and not missing.unexpected_keys
    net.set_to_initialized_state()
    net.eval()
    return net, sd


def surface_patches(synth, n_cloud=2048, seed=2021, shape="torus"):
    """Real kNN-256 patches of a synthetic cloud, through the reference's own PatchHelper."""
    from modules.utils.patch import PatchHelper
    pc = torch.from_numpy(synth.make_cloud(n_cloud, seed, shape))[None]
    ph = PatchHelper(256, patch_expand_ratio=4)
    pcn, _, _ = PatchHelper.normalize_pc(pc)
    patches = ph.extract_knn_patch(pcn, ph.knn, 256, 4)          # [1, n_patch, 256, 3]
    patches = patches.reshape(-1, 256, 3)
    patches, _, _ = PatchHelper.normalize_pc(patches)
    return patches.contiguous()


# ------------------------------------------------------------------------------------------------
@torch.no_grad()
def cmd_golden_extra(args):
    """The other two discrete checkpoints of the reference (pretrain/puflow-x4-pugan.pt, -pugeo.pt): network
    stages on 3 real surface patches each, so that parity (and the fp16 range assumptions of the tensor-core kernels)
    is pinned for every released weight set, not only pu1k."""
    synth = _load_synth()
    out = args.out
    for tag in ("pugan", "pugeo"):
        torch.manual_seed(2021)
        net, sd = load_network(f"puflow-x4-{tag}.pt")
        torch.save(sd, os.path.join(out, f"puflow-x4-{tag}.pt"))
        xs = [surface_patches(synth, 2048, 3021 + i, s)[7 * i + 2] for i, s in enumerate(synth.SHAPES)]
        xyz = torch.stack(xs, 0).contiguous()
        g = {"xyz": xyz}
        d16, idx16, _ = knn_points(xyz, xyz, K=16)
        g["knn_idx"] = idx16.int()
        c = xyz
        for i in range(net.num_blocks):
            c = net.feat_convs[i](c, knn_idx=idx16)
            g[f"feat{i}"] = c
            g[f"cs{i}"] = net.merge_convs[i](c)
        cs = [g[f"cs{i}"] for i in range(6)]
        g["z"], g["log_det_J"] = net.f(xyz, cs)
        g["fz"] = net.interp(g["z"], xyz, 4)
        x, logp = net(xyz, 4)
        g["x"], g["logp_x"] = x, logp.reshape(1)
        np.savez_compressed(os.path.join(out, f"net_{tag}.npz"), **{k: v.numpy() for k, v in g.items()})
        print(tag, "max |feat5|", float(g["feat5"].abs().max()), "max |x|", float(x.abs().max()), flush=True)


@torch.no_grad()
def cmd_golden(args):
    synth = _load_synth()
    out = args.out
    os.makedirs(out, exist_ok=True)
    torch.manual_seed(2021)
    net, sd = load_network()

    # (1) the checkpoint, re-serialised (data asset; 408 keys) so GPU-box tests use real weights
    torch.save(sd, os.path.join(out, "puflow-x4-pu1k.pt"))

    # (2) network stages on 3 real surface patches (torus / sphere / cube)
    xs = [surface_patches(synth, 2048, 2021 + i, s)[5 * i + 1] for i, s in enumerate(synth.SHAPES)]
    xyz = torch.stack(xs, 0).contiguous()
    g = {"xyz": xyz}
    d16, idx16, _ = knn_points(xyz, xyz, K=16)
    g["knn_idx"], g["knn_d"] = idx16.int(), d16
    c = xyz
    for i in range(net.num_blocks):
        c = net.feat_convs[i](c, knn_idx=idx16)
        g[f"feat{i}"] = c
        g[f"cs{i}"] = net.merge_convs[i](c)
    cs = [g[f"cs{i}"] for i in range(6)]
    z, logdet = net.f(xyz, cs)
    g["z"], g["log_det_J"] = z, logdet
    _, logp = net.log_prob(xyz, cs)
    g["logp_x"] = logp.reshape(1)
    ctx, idx8 = net.interp.knn_context(xyz)
    g["idx8"] = idx8.view(3, 256, 8).int()
    w = net.interp.weight_unit(ctx)
    g["interp_w"] = torch.softmax(w[:, :, :4], dim=-1)            # [B,N,4,8]
    fz = net.interp(z, xyz, 4)
    g["fz"] = fz
    g["x"] = net.g(fz, cs, 4)
    x2, logp2 = net(xyz, 4)
    assert torch.equal(x2, g["x"]) and torch.equal(logp2.reshape(1), g["logp_x"])
    np.savez_compressed(os.path.join(out, "net_pu1k.npz"), **{k: v.numpy() for k, v in g.items()})

    # (3) flow layers one by one (block 0 = even split, block 1 = odd split)
    fl = {}
    gen = torch.Generator().manual_seed(7)
    x = torch.randn(2, 64, 3, generator=gen) * 0.5
    for bi in (0, 1):
        fb = net.flow_blocks[bi]
        cdim = fb.coupling2.bias_net.layers[0].weight.shape[1]
        cc = torch.randn(2, 64, cdim, generator=gen)
        fl[f"b{bi}_x"], fl[f"b{bi}_c"] = x, cc
        for nm, mod in (("actnorm", fb.actnorm), ("permutate1", fb.permutate1), ("coupling1", fb.coupling1),
                        ("permutate2", fb.permutate2), ("coupling2", fb.coupling2)):
            y, ld = mod(x, cc)
            xi, ldi = mod.inverse(x, cc)
            fl[f"b{bi}_{nm}_fwd"], fl[f"b{bi}_{nm}_inv"] = y, xi
            if ld is not None:
                fl[f"b{bi}_{nm}_fwd_ld"] = torch.as_tensor(ld).reshape(-1)
                fl[f"b{bi}_{nm}_inv_ld"] = torch.as_tensor(ldi).reshape(-1)
        y, ld = fb(x, cc)
        fl[f"b{bi}_block_fwd"], fl[f"b{bi}_block_fwd_ld"] = y, ld.reshape(-1)
        fl[f"b{bi}_block_inv"] = fb.inverse(x, cc)
    np.savez_compressed(os.path.join(out, "flows_pu1k.npz"), **{k: v.numpy() for k, v in fl.items()})

    # (4) Chamfer: the reference's own known-answer test shape (metric/PyTorchCD/unit_test.py:14-35)
    import chamfer_python
    gen = torch.Generator().manual_seed(11)
    a, b = torch.rand(4, 100, 3, generator=gen), torch.rand(4, 200, 3, generator=gen)
    d1, d2, i1, i2 = chamfer_python.distChamfer(a, b)
    np.savez_compressed(os.path.join(out, "chamfer_unit.npz"), a=a.numpy(), b=b.numpy(), dist1=d1.numpy(),
                        dist2=d2.numpy(), idx1=i1.numpy(), idx2=i2.numpy())

    # (5) end to end (PatchHelper.upsample + remove_outliers, upsample.py:46-54) on small clouds
    from modules.utils.patch import PatchHelper
    import modules.utils.patch as patch_mod

    def remove_outliers(sr, lr, n_out):            # patch.py:198-214 with the fp64 torch Chamfer
        d1, _, _, _ = chamfer_python.distChamfer(sr, lr)
        order = torch.argsort(d1, dim=-1, descending=True)
        keep = torch.ones(sr.shape[:2], dtype=torch.bool)
        keep[torch.arange(sr.shape[0])[:, None], order[:, :n_out]] = False
        return sr[keep].view(sr.shape[0], -1, 3), d1

    for n_cloud, shape, seed in ((512, "torus", 2021), (2048, "sphere", 2021)):
        pc = torch.from_numpy(synth.make_cloud(n_cloud, seed, shape))[None].contiguous()
        ph = PatchHelper(256, patch_expand_ratio=4)
        e = {"pc": pc}
        # intermediates, re-deriving what PatchHelper.upsample does (patch.py:35-79)
        pcn, cen, rad = PatchHelper.normalize_pc(pc)
        e["pc_norm"], e["centroid"], e["radius"] = pcn, cen, rad
        n_patch = int(n_cloud / 256 * 4)
        seeds = patch_mod.furthest_point_sample(pcn, n_patch)
        e["seed_idx"] = seeds
        _, idxp = ph.knn(pcn.transpose(1, 2).contiguous(), patch_mod.gather_operation(pcn.transpose(1, 2).contiguous(), seeds))
        e["patch_idx"] = idxp.transpose(1, 2).int()               # [1, n_patch, 256]
        t0 = time.time()
        pred = ph.upsample(net, pc, npoint=n_cloud * 4 + 24, upratio=4, jitter=False)
        e["pred_raw"] = pred                                      # [1, 4N+24, 3]
        final, d1 = remove_outliers(pred, pc, 24)
        e["pred"], e["dist1"] = final, d1
        print(f"e2e {n_cloud}: {time.time() - t0:.1f}s", final.shape)
        np.savez_compressed(os.path.join(out, f"e2e_{n_cloud}.npz"), **{k: v.numpy() for k, v in e.items()})
    print("golden fixtures written to", out)


@torch.no_grad()
def cmd_time(args):
    synth = _load_synth()
    torch.set_num_threads(os.cpu_count())
    net, _ = load_network()
    x = surface_patches(synth, 2048, 2021, "sphere")
    print("patches", tuple(x.shape), "threads", torch.get_num_threads())
    net(x, 4)
    ts = []
    for _ in range(args.reps):
        t0 = time.time()
        y, _ = net(x, 4)
        ts.append(time.time() - t0)
    t = float(np.median(ts))
    print(f"reference PointInterpFlow.forward {x.shape[0]} patches: {t:.3f}s -> {y.shape[0] * y.shape[1] / t:.0f} new pts/s")


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    sub = ap.add_subparsers(dest="cmd", required=True)
    g = sub.add_parser("golden")
    g.add_argument("--out", default=os.path.join(REPO, "tests", "golden"))
    ge = sub.add_parser("golden-extra")
    ge.add_argument("--out", default=os.path.join(REPO, "tests", "golden"))
    t = sub.add_parser("time")
    t.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    install_shims()
    {"golden": cmd_golden, "golden-extra": cmd_golden_extra, "time": cmd_time}[a.cmd](a)
