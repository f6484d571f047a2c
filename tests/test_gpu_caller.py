"""GPU tests of the caller-side kernels (csrc/caller.cu: SURVEY.md section 8 rows a18, a19, f3) and of the one-call cloud
entry point pf_upsample_clouds, against the torch expressions of the reference's PatchHelper
(/root/reference/modules/utils/patch.py:82-93,142-158,168-178,198-214) evaluated on the same inputs."""
import os
import threading

import numpy as np
import pytest
import torch

from tests import parity
from tools import synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
CKPT = os.path.join(os.path.dirname(__file__), "golden", "puflow-x4-pu1k.pt")


@pytest.fixture(scope="module")
def ops():
    from puflow_b200 import ops as o
    return o


@pytest.fixture(scope="module")
def net():
    from modules.discrete.interpflow import PointInterpFlow
    n = PointInterpFlow(3)
    n.load_state_dict(torch.load(CKPT, map_location="cpu"))
    n.set_to_initialized_state()
    return n.to(DEV).eval()


def _normalize_ref(pc):            # patch.py:168-178, float64 so that the fp32 kernel is within rounding of it
    pc = pc.double()
    cen = pc.mean(1, keepdim=True)
    q = pc - cen
    rad = (q ** 2).sum(-1, keepdim=True).sqrt().max(1, keepdim=True)[0]
    return q / rad, cen, rad


@pytest.mark.parametrize("B,n", [(3, 2048), (5, 256), (2, 5000), (1, 7)])
def test_normalize_pc(ops, B, n):
    pc = torch.from_numpy(np.stack([synth.make_cloud(n, 11 + b, "torus") * (1 + b) + b for b in range(B)], 0)).float()
    out, cen, rad = ops.normalize_pc(pc.to(DEV))
    ro, rc, rr = _normalize_ref(pc)
    assert cen.shape == (B, 1, 3) and rad.shape == (B, 1, 1)
    assert (cen.cpu().double() - rc).abs().max() <= 1e-6 * max(1.0, float(rc.abs().max()))
    assert ((rad.cpu().double() - rr).abs() <= 4e-7 * rr).all()
    assert (out.cpu().double() - ro).abs().max() <= 1e-6
    assert float(out.norm(dim=-1).max(1)[0].sub(1).abs().max()) <= 2e-7          # the farthest point lands on the unit sphere


def test_gather_normalize_and_points(ops):
    B, n, items, k = 3, 1000, 7, 256
    g = torch.Generator().manual_seed(3)
    pc = torch.randn(B, n, 3, generator=g)
    idx = torch.randint(0, n, (B, items, k), generator=g, dtype=torch.int32)
    out, cen, rad = ops.gather_normalize(pc.to(DEV), idx.to(DEV))
    patches = pc[torch.arange(B).view(-1, 1), idx.long().view(B, -1)].view(B * items, k, 3)
    same, c2, r2 = ops.normalize_pc(patches.to(DEV))                       # gather + normalise == normalise of the gathered
    assert torch.equal(out, same) and torch.equal(cen, c2) and torch.equal(rad, r2)
    gp = ops.gather_points(pc.to(DEV), idx.view(B, -1).to(DEV))
    assert torch.equal(gp.cpu(), patches.view(B, items * k, 3))
    scale, shift = torch.rand(B, 1, 1) + 0.5, torch.randn(B, 1, 3)
    gs = ops.gather_points(pc.to(DEV), idx.view(B, -1).to(DEV), scale.to(DEV), shift.to(DEV))
    assert torch.equal(gs.cpu(), patches.view(B, items * k, 3) * scale + shift)   # separately rounded * then +, like torch


def test_denorm_concat(ops):
    g = torch.Generator().manual_seed(5)
    x, p = torch.randn(6, 1024, 3, generator=g), torch.randn(6, 256, 3, generator=g)
    cen, rad = torch.randn(6, 1, 3, generator=g), torch.rand(6, 1, 1, generator=g) + 0.1
    out = ops.denorm_concat(x.to(DEV), p.to(DEV), cen.to(DEV), rad.to(DEV))
    assert torch.equal(out.cpu(), torch.cat([x, p], 1) * rad + cen)          # patch.py:90-91


def _remove_outliers_ref(sr, d, k):        # patch.py:205-214 with a stable argsort
    order = torch.argsort(d, dim=-1, descending=True, stable=True)
    keep = torch.ones_like(d, dtype=torch.bool)
    keep[torch.arange(d.shape[0]).view(-1, 1), order[:, :k]] = False
    return torch.stack([sr[b][keep[b]] for b in range(d.shape[0])], 0)


@pytest.mark.parametrize("B,n,k,kind", [(3, 8216, 24, "random"), (2, 20024, 24, "random"), (4, 1000, 24, "ties"), (2, 300, 299, "random"),
                                        (2, 5000, 1, "ties"), (1, 70000, 24, "ties"), (2, 2048, 24, "zeros")])
def test_remove_outliers_kernel(ops, B, n, k, kind):
    g = torch.Generator().manual_seed(n + k)
    sr = torch.randn(B, n, 3, generator=g)
    if kind == "random":
        d = torch.rand(B, n, generator=g) ** 4
    elif kind == "ties":                       # few distinct values: the cut falls inside a run of equal distances
        d = torch.randint(0, 6, (B, n), generator=g).float() * 0.125
    else:
        d = torch.zeros(B, n)
    out = ops.remove_outliers(sr.to(DEV), d.to(DEV), k)
    assert torch.equal(out.cpu(), _remove_outliers_ref(sr, d, k))


def test_one_call_clouds_matches_the_staged_path(net):
    """pf_upsample_clouds (one C call, also replayed from a CUDA graph) == PatchHelper's stage-by-stage path (the same
    kernels launched one by one from Python), bit for bit."""
    from modules.utils.patch import PatchHelper
    from puflow_b200 import ops
    B, n = 3, 1024
    pc = torch.from_numpy(synth.make_clouds(B, n, seed=77)).to(DEV)
    ph = PatchHelper(256, patch_expand_ratio=4)
    eng = net.engine(pc.device)
    n_patch = int(n / 256 * 4)
    one = eng.upsample_clouds(pc, 4, n * 4 + 24, 0, 256, n_patch)
    prev = ops.set_launch_log(ops.LaunchLog(events=True))          # forces PatchHelper.upsample onto the staged path
    try:
        staged = ph.upsample(net, pc, npoint=n * 4 + 24, upratio=4, jitter=False)
    finally:
        ops.set_launch_log(prev)
    assert torch.equal(one, staged)
    assert torch.equal(ph.upsample(net, pc, npoint=n * 4 + 24, upratio=4, jitter=False), one)      # default = the one call
    full = eng.upsample_clouds(pc, 4, n * 4 + 24, 24, 256, n_patch)
    assert torch.equal(full, PatchHelper.remove_outliers(one, pc, 24))
    graphed = eng.upsample_clouds(pc, 4, n * 4 + 24, 24, 256, n_patch, graph=True)
    again = eng.upsample_clouds(pc.flip(0).contiguous(), 4, n * 4 + 24, 24, 256, n_patch, graph=True)   # replay, new data
    assert torch.equal(graphed, full) and torch.equal(again, full.flip(0))


def test_upsample_array_host_buffers(net):
    from modules.discrete.upsample import upsample_array
    from modules.utils.patch import PatchHelper
    pc = synth.make_clouds(2, 512, seed=5)
    ph = PatchHelper(256, patch_expand_ratio=4)
    a = upsample_array(net, ph, pc, 4, 24, None, DEV)
    b = upsample_array(net, ph, pc, 4, 24, None, DEV, graph=False)
    out = torch.empty((2, 2048, 3)).pin_memory()
    upsample_array(net, ph, torch.from_numpy(pc).pin_memory(), 4, 24, None, DEV, out=out)
    assert a.shape == (2, 2048, 3) and np.array_equal(a, b) and np.array_equal(a, out.numpy())


def test_upsample_arrays_pipelined_loop(net):
    """upsample_arrays (copies of neighbouring batches on copy streams under the kernels) yields, in order, exactly what
    one synchronous upsample_array call per batch returns: different data per batch, two cloud sizes (two CUDA graphs),
    pinned result buffers used round-robin, numpy results without them, an empty sequence."""
    from modules.discrete.upsample import upsample_array, upsample_arrays
    from modules.utils.patch import PatchHelper
    ph = PatchHelper(256, patch_expand_ratio=4)
    batches = [synth.make_clouds(2, 512, seed=s) for s in (1, 2, 3, 4, 5)] + [synth.make_clouds(2, 1024, seed=6)]
    want = [upsample_array(net, ph, b, 4, 24, None, DEV) for b in batches]
    got = [r.copy() for r in upsample_arrays(net, ph, batches, 4, 24, None, DEV)]
    assert len(got) == len(want) and all(np.array_equal(g, w) for g, w in zip(got, want))
    outs = [torch.empty((2, 2048, 3)).pin_memory() for _ in range(2)]
    pinned = (torch.from_numpy(b).pin_memory() for b in batches[:5])
    n = 0
    for i, r in enumerate(upsample_arrays(net, ph, pinned, 4, 24, None, DEV, outs=outs)):
        assert r is outs[i % 2] and np.array_equal(r.numpy(), want[i])       # valid until two batches later
        n += 1
    assert n == 5
    assert list(upsample_arrays(net, ph, [], 4, 24, None, DEV)) == []
    with pytest.raises(ValueError):
        list(upsample_arrays(net, ph, batches[:1], 4, 24, None, DEV, outs=outs[:1]))


def test_two_host_threads_two_streams(net):
    """One host thread per stream through pf_upsample_patches (profiling off): no shared scratch, no shared state
    (SURVEY.md section 8b 'kernels re-entrant'); both results equal the single-threaded one."""
    eng = net.engine(torch.device(DEV))
    xs = [torch.from_numpy(synth.make_patches(40, 256, seed=s, shape="surface")).to(DEV) for s in (1, 2)]
    want = [eng.forward_c(x, 4)[0].clone() for x in xs]
    torch.cuda.synchronize()
    got, errs = [None, None], []

    def work(i):
        try:
            st = torch.cuda.Stream(DEV)
            with torch.cuda.stream(st):
                for _ in range(5):
                    y = eng.forward_c(xs[i], 4)[0]
                st.synchronize()
            got[i] = y
        except Exception as e:      # noqa: BLE001
            errs.append(e)

    th = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    assert torch.equal(got[0], want[0]) and torch.equal(got[1], want[1])
