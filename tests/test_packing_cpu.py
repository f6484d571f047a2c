"""CPU: the packed / factored arithmetic the kernels implement (tests/emulate.py on
puflow_b200.packing output) reproduces the unmodified reference's stage outputs (tests/golden)."""
import numpy as np
import torch

from puflow_b200 import packing
from tests import emulate, parity


def test_blob_sizes(state_dict):
    w = packing.pack_state_dict(state_dict, "cpu")
    from puflow_b200 import _native as N
    assert w["flow"].shape == (6, N.PF_FLOW_BLOB) and w["cond"].shape == (6, N.PF_COND_BLOB)
    assert w["interp"]["blob"].numel() == N.PF_INTERP_BLOB and w["interp"]["wpq"].shape == (3, 640)
    for i, (C, g, Co) in enumerate(packing.FEAT_CFG):
        f = w["feat"][i]
        assert f["wpq"].shape == (C, 2 * (4 * g + Co)) and f["wd"].numel() == 6 * g * g and f["wo"].shape == (4 * g, Co)
        assert w["wc"][i].shape[1] == 192
    assert all(t.dtype == torch.float32 for t in (w["flow"], w["cond"], w["interp"]["blob"]))


def test_factored_path_matches_reference(golden, state_dict):
    g = golden("net_pu1k.npz")
    w = packing.pack_state_dict(state_dict, "cpu")
    xyz = torch.from_numpy(g["xyz"])
    r = emulate.forward(w, xyz, torch.from_numpy(g["knn_idx"]), 4)
    for i in range(6):
        ok, err = parity.close_norm(r["feats"][i].numpy(), g[f"feat{i}"], 1e-4)
        assert ok, ("feat", i, err)
        ok, err = parity.close_norm(r["cs"][i].numpy(), g[f"cs{i}"], 1e-4)
        assert ok, ("cs", i, err)
    ok, err = parity.close_norm(r["z"].numpy(), g["z"], 1e-3)
    assert ok, ("z", err)
    ok, err = parity.close_norm(r["logdet"].numpy(), g["log_det_J"], 1e-4)
    assert ok, ("logdet", err)
    assert abs(float(r["logp_x"]) - float(g["logp_x"][0])) <= 1e-4 * abs(float(g["logp_x"][0]))
    # interpolation weights: golden is [B,N,R,k]; ours [M,k,R]
    ours = r["interp_w"].reshape(3, 256, 8, 4).permute(0, 1, 3, 2).numpy()
    assert np.abs(ours - g["interp_w"]).max() < 2e-5
    fz = r["zup"].reshape(3, 256, 4, 3).permute(0, 1, 3, 2).numpy()          # -> [B,N,3,R]
    ok, err = parity.close_norm(fz, g["fz"], 1e-3)
    assert ok, ("fz", err)
    ok, err = parity.close_norm(r["x"].numpy(), g["x"], 1e-4)
    assert ok, ("x", err)


def test_flow_blocks_match_reference_layers(golden, state_dict):
    """Block 0 (even split) and block 1 (odd split) against the reference's FlowBlock.forward/inverse on
    random x / c (flows_pu1k.npz), through the packed blobs."""
    fl = golden("flows_pu1k.npz")
    w = packing.pack_state_dict(state_dict, "cpu")
    for bi in (0, 1):
        x = torch.from_numpy(fl[f"b{bi}_x"]).reshape(-1, 3)
        c = torch.from_numpy(fl[f"b{bi}_c"]).reshape(x.shape[0], -1)
        raw = emulate.linear(c, w["wc"][bi])[None]
        sb = emulate.flow_cond(raw, w["cond"][bi:bi + 1])
        y, ssum = emulate.flow_fwd(x, raw, sb, w["flow"][bi:bi + 1])
        ok, err = parity.close_norm(y.reshape(2, 64, 3).numpy(), fl[f"b{bi}_block_fwd"], 1e-4)
        assert ok, (bi, "fwd", err)
        xi = emulate.flow_inv(x, raw, sb, w["flow"][bi:bi + 1], 1)
        ok, err = parity.close_norm(xi.reshape(2, 64, 3).numpy(), fl[f"b{bi}_block_inv"], 1e-4)
        assert ok, (bi, "inv", err)
        back = emulate.flow_inv(y, raw, sb, w["flow"][bi:bi + 1], 1)
        assert (back - x).abs().max() < 1e-5
