"""Pins the CPU oracle (oracle/) against the golden fixtures produced by the unmodified reference
(oracle/ref/run_reference.py, run in the build container).  CPU only."""
import numpy as np
import pytest
import torch

from oracle import puflow_oracle as po
from oracle import synth
from tests import parity


def _close(a, b, rtol=1e-4):
    """Normwise per batch item: max|a-b| <= rtol * max|b| (see tests/parity.py for why)."""
    return parity.close_norm(a, b, rtol)


def test_knn_matches_reference_standin(golden):
    g = golden("net_pu1k.npz")
    d, idx = po.knn(g["xyz"], g["xyz"], 16)
    assert np.array_equal(idx, g["knn_idx"].astype(np.int64))
    assert np.array_equal(d, g["knn_d"])           # bit-exact distances
    _, idx8 = po.knn(g["xyz"], g["xyz"], 8)
    assert np.array_equal(idx8, g["idx8"].astype(np.int64))
    assert np.array_equal(idx8, idx[:, :, :8])     # K=8 list == first 8 of the sorted K=16 list


def test_knn_lattice_ties():
    p = synth.make_cloud(343, 3, "lattice")[None]
    d, idx = po.knn(p, p, 16)
    dd = ((p[0][:, None, :] - p[0][None, :, :]) ** 2).sum(-1)
    order = np.lexsort((np.arange(343)[None, :].repeat(343, 0), dd), axis=-1)[:, :16]
    assert np.array_equal(idx[0], order)


def test_network_stages_match_reference(golden, state_dict):
    g = golden("net_pu1k.npz")
    o = po.PUFlowOracle(state_dict)
    r = o.forward(torch.from_numpy(g["xyz"]), 4, return_all=True)
    assert np.array_equal(r["knn_idx"].numpy(), g["knn_idx"])
    for i in range(6):
        ok, err = _close(r["feats"][i].numpy(), g[f"feat{i}"])
        assert ok, (i, err)
        ok, err = _close(r["cs"][i].numpy(), g[f"cs{i}"])
        assert ok, (i, err)
    for k in ("z", "fz", "x"):
        ok, err = _close(r[k].numpy(), g[k])
        assert ok, (k, err)
    assert abs(float(r["logp_x"]) - float(g["logp_x"][0])) <= 1e-4 * abs(float(g["logp_x"][0]))


@pytest.mark.parametrize("tag", ["pugan", "pugeo"])
def test_network_stages_match_reference_other_checkpoints(golden, tag):
    """The reference's other two discrete weight sets (pretrain/puflow-x4-pugan.pt, -pugeo.pt; fixtures by
    oracle/ref/run_reference.py golden-extra)."""
    import os
    sd = torch.load(os.path.join(os.path.dirname(__file__), "golden", f"puflow-x4-{tag}.pt"), map_location="cpu")
    g = golden(f"net_{tag}.npz")
    r = po.PUFlowOracle(sd).forward(torch.from_numpy(g["xyz"]), 4, return_all=True)
    assert np.array_equal(r["knn_idx"].numpy(), g["knn_idx"])
    for i in range(6):
        ok, err = _close(r["feats"][i].numpy(), g[f"feat{i}"])
        assert ok, (i, err)
        ok, err = _close(r["cs"][i].numpy(), g[f"cs{i}"])
        assert ok, (i, err)
    for k in ("z", "x"):
        ok, err = _close(r[k].numpy(), g[k])
        assert ok, (k, err)


def test_fp64_oracle_is_consistent(golden, state_dict):
    g = golden("net_pu1k.npz")
    o = po.PUFlowOracle(state_dict, dtype=torch.float64)
    x, _ = o.forward(torch.from_numpy(g["xyz"]), 4)
    ok, err = _close(x.numpy(), g["x"])
    assert ok, err


def test_flow_blocks_match_reference(golden, state_dict):
    fl = golden("flows_pu1k.npz")
    o = po.PUFlowOracle(state_dict)
    for bi in (0, 1):
        x, c = torch.from_numpy(fl[f"b{bi}_x"]), torch.from_numpy(fl[f"b{bi}_c"])
        y, ld = o.flow_fwd(bi, x, c)
        assert _close(y.numpy(), fl[f"b{bi}_block_fwd"])[0]
        assert _close(ld.numpy(), fl[f"b{bi}_block_fwd_ld"])[0]
        assert _close(o.flow_inv(bi, x, c).numpy(), fl[f"b{bi}_block_inv"])[0]
        # round trip
        assert _close(o.flow_inv(bi, y, c).numpy(), x.numpy(), rtol=1e-5)[0]


def test_chamfer_known_answer(golden):
    """The reference's own unit test (metric/PyTorchCD/unit_test.py:14-35): MSE < 1e-8, idx equal."""
    g = golden("chamfer_unit.npz")
    d1, d2, i1, i2 = po.chamfer_nn(g["a"], g["b"])
    assert np.mean((d1 - g["dist1"]) ** 2) + np.mean((d2 - g["dist2"]) ** 2) < 1e-8
    assert np.array_equal(i1, g["idx1"]) and np.array_equal(i2, g["idx2"])


def test_fps_and_patch_extraction_match_reference(golden):
    for name in ("e2e_512.npz", "e2e_2048.npz"):
        e = golden(name)
        n_patch = e["seed_idx"].shape[1]
        assert np.array_equal(po.fps(e["pc_norm"], n_patch), e["seed_idx"])
        seeds = e["pc_norm"][0][e["seed_idx"][0]][None]
        _, pidx = po.knn(seeds, e["pc_norm"], 256)
        assert np.array_equal(pidx, e["patch_idx"].astype(np.int64))


def test_fps_pointnet2_restatement_reduces_to_fps_py():
    """oracle.fps_pointnet2 (upstream pointnet2_ops' candidate rule, unpinned) with the rule switched off is the pinned
    fps.py restatement (nn_oracle.c); with it on, no candidate within the threshold is selected after the start point and
    clouds without such points are unchanged."""
    rng = np.random.default_rng(11)
    p = rng.uniform(-1, 1, (2, 700, 3)).astype(np.float32)
    assert np.array_equal(po.fps_pointnet2(p, 96, thresh2=-1.0), po.fps(p, 96))
    p[:, 5:25] *= np.float32(0.01)                                   # twenty points within 0.0174 of the origin
    sel = po.fps_pointnet2(p, 96)
    assert ((p[np.arange(2)[:, None], sel[:, 1:]] ** 2).sum(-1) > 1e-3).all()
    shell = p / np.linalg.norm(p, axis=-1, keepdims=True)
    assert np.array_equal(po.fps_pointnet2(shell, 96), po.fps(shell, 96))


def test_end_to_end_small_cloud(golden, state_dict):
    """FPS over the merged candidates is chaotic under 1e-7 perturbations (patches overlap, so the
    candidate set holds near-duplicate points): rows need not line up, the point SET must.
    north_star: Chamfer distance to the reference output <= 1e-5."""
    e = golden("e2e_512.npz")
    o = po.PUFlowOracle(state_dict)
    pc = torch.from_numpy(e["pc"])
    pred = po.upsample_cloud(o, pc, 512 * 4 + 24)
    assert pred.shape == e["pred_raw"].shape
    assert parity.chamfer_distance(pred[0].numpy(), e["pred_raw"][0]) <= 1e-5
    out = po.remove_outliers(pred, pc, 24)
    assert out.shape == e["pred"].shape
    assert parity.chamfer_distance(out[0].numpy(), e["pred"][0]) <= 1e-5
    # given the reference's own candidates, outlier removal is exact
    out2 = po.remove_outliers(torch.from_numpy(e["pred_raw"]), pc, 24)
    assert np.array_equal(out2.numpy(), e["pred"])
