"""CPU: the packed / factored arithmetic the kernels implement (tests/emulate.py on
puflow_b200.packing output) reproduces the unmodified reference's stage outputs (tests/golden)."""
import numpy as np
import pytest
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


# ---- tensor-core blobs: decode the operand layout on the CPU and compare with the float64 source ---------------
def _umma_decode(buf, rows, Kt):
    """inverse of packing.umma_kmajor_fp16's layout: uint8 -> [rows, Kt] float32 (one part)."""
    h = buf.view(torch.float16).view(rows // 8, Kt // 8, 8, 8).permute(0, 2, 1, 3).reshape(rows, Kt)
    return h.float() / packing.TC_WSCALE


def test_tensor_core_blobs_decode_to_the_weights(state_dict):
    w = packing.pack_state_dict(state_dict, "cpu")
    from puflow_b200 import _native as N
    # flow: W2 hi + lo == W2 to fp16-pair precision; fp32 tail is the head of the fp32 blob
    for i in range(6):
        b = w["flow_tc"][i]
        assert b.numel() == N.PF_FLOW_TC_BLOCK
        W2 = w["flow"][i][384:4480].view(64, 64).t()
        dec = _umma_decode(b[:8192], 64, 64) + _umma_decode(b[8192:16384], 64, 64)
        assert (dec - W2).abs().max() <= 2.0 ** -21 * W2.abs().max()
        # the 2^10 factor keeps the lo halves out of the fp16 subnormals: 22 good bits for every weight above 2^-12
        big = W2.abs() >= 2.0 ** -12
        assert ((dec - W2).abs()[big] <= 2.0 ** -21 * W2.abs()[big]).all()
        assert torch.equal(b[16384:].view(torch.float32), w["flow"][i][:384])
        c = w["cond_tc"][i]
        assert c.numel() == N.PF_COND_TC_BLOCK
        assert torch.equal(c[32768:].view(torch.float32), w["cond"][i][8192:])
    # chain: composed layer W2c = (conv2^T . wc)^T
    for i, (C, g, Co) in enumerate(packing.FEAT_CFG):
        b = w["chain_tc"][i]
        assert b.numel() == N.chain_tc_bytes(Co)
        H1 = Co // 2
        W1 = state_dict[f"merge_convs.{i}.conv1.weight"].double()
        dec = _umma_decode(b[:Co * Co], H1, Co) + _umma_decode(b[Co * Co:2 * Co * Co], H1, Co)
        assert (dec.double() - W1).abs().max() <= 2.0 ** -21 * W1.abs().max()
        W2c = (w["wc"][i].double().t() @ state_dict[f"merge_convs.{i}.conv2.weight"].double())
        o = 2 * Co * Co
        dec = _umma_decode(b[o:o + 192 * Co], 192, H1) + _umma_decode(b[o + 192 * Co:o + 384 * Co], 192, H1)
        assert (dec.double() - W2c).abs().max() <= 1e-6 * W2c.abs().max()
    # EdgeConv: K slices padded to 16 for g = 8, Wo rows repeated for C_out < 128
    for i, (C, g, Co) in enumerate(packing.FEAT_CFG):
        f = w["feat"][i]
        gp = max(g, 16)
        assert f["wtc"].numel() == packing.edgeconv_tc_blob_bytes(g)
        Wl, woff = {}, 0
        for l in range(1, 4):
            Wl[l] = f["wd"][woff:woff + g * l * g].view(g * l, g).t()   # [g out, g*l in]
            woff += g * l * g
        off = 0
        for m in range(3):                                             # K-slice m against layers m+1..3, rows stacked
            rows = gp * (3 - m)
            part = 2 * rows * gp
            dec = _umma_decode(f["wtc"][off:off + part], rows, gp) + _umma_decode(f["wtc"][off + part:off + 2 * part], rows, gp)
            for n, l in enumerate(range(m + 1, 4)):
                blk = dec[n * gp:(n + 1) * gp]
                assert (blk[:g, :g] - Wl[l][:, m * g:(m + 1) * g]).abs().max() <= 2.0 ** -21 * Wl[l].abs().max()
                assert (blk[:g, g:] == 0).all() and (blk[g:] == 0).all()
            off += 2 * part
        KT = 4 * gp
        hi = f["wtc"][off:off + 128 * KT * 2].view(torch.float16).view(128, KT).float() / packing.TC_WSCALE
        lo = f["wtc"][off + 128 * KT * 2:].view(torch.float16).view(128, KT).float() / packing.TC_WSCALE
        wo = f["wo"].t()                                               # [C_out, 4g]
        for r in (0, Co - 1, Co % 128, 127):
            for j in range(4):
                assert ((hi + lo)[r, j * gp:j * gp + g] - wo[r % Co, j * g:(j + 1) * g]).abs().max() <= 2.0 ** -21 * wo.abs().max()


def test_model_blob_layout(state_dict):
    """The single blob of pf_upsample_patches: the layout is enumerated by the C library (pf_model_entry), entries are
    256-byte aligned, in order, non-overlapping, and each holds exactly the bytes of the piece the per-kernel path uses."""
    from puflow_b200 import _native
    w = packing.pack_state_dict(state_dict, "cpu")
    blob = packing.pack_model_blob(w)
    ents = _native.model_entries()
    assert len(ents) == _native.PF_MODEL_ENTRIES and blob.numel() == _native.lib.pf_model_bytes()
    end = 0
    for name, off, nbytes in ents:
        assert off % 256 == 0 and off >= end and nbytes > 0, (name, off, nbytes)
        end = off + nbytes
    assert end <= blob.numel() < end + 256
    by = {n: (o, b) for n, o, b in ents}
    o, b = by["feat3.wtc"]
    assert torch.equal(blob[o:o + b], w["feat"][3]["wtc"])
    o, b = by["feat0.wpq"]
    assert torch.equal(blob[o:o + b].view(torch.float32).view(3, 128), w["feat"][0]["wpq"])
    o, b = by["chain5"]
    assert torch.equal(blob[o:o + b], w["chain_tc"][5])
    o, b = by["flow_tc"]
    assert torch.equal(blob[o:o + b].view(6, -1), w["flow_tc"])
    o, b = by["logdet_const"]
    assert blob[o:o + b].view(torch.float32).item() == pytest.approx(float(w["logdet_const"]), rel=1e-7)
    assert "feat0.wpq_tc" not in by and "feat1.wpq_tc" in by
    # workspace size: monotone, -1 for nonsense
    f = _native.lib.pf_upsample_workspace_bytes
    assert f(32, 256, 4) > 0 and f(64, 256, 4) > f(32, 256, 4) and f(-1, 256, 4) == -1
