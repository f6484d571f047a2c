"""The reduced-precision tensor-core variant (PF_PREC_FP16: single-pass fp16 operands, fp32 accumulation -- the mantissa
width of TF32) against the reference fixtures, BASELINE.json config 5 / north_star "any TF32/BF16 tensor-core variant must
have its tolerance stated".

STATED TOLERANCE of precision="fp16" (measured on B200 against the reference fixtures, tools/dev_precision.py:
x 3.5e-3 / 6.1e-2 / 1.3e-2 and z 0.16 / 0.21 / 0.11 normwise for the pu1k / pugan / pugeo checkpoints, where the default
split path reads 1.4e-5 / 6.0e-5 / 2.0e-5 and 2e-4 / 3e-4 / 1e-4 -- the flow is ill-conditioned, 11-bit operands cost two
to three orders of magnitude):
  * network output x: 1e-2 (pu1k), 3e-2 (pugeo), 1e-1 (pugan) normwise per patch against the unmodified reference
    (fp32-split path: 1e-4 for all three);
  * latent z: 3e-1 normwise (fp32-split path: 1e-3);
  * end to end (pu1k, one 2,048-pt cloud): Chamfer distance to the reference output <= 1e-4 -- it does NOT meet the
    1e-5 bar north_star sets for the product path (the default path does), which is why it is an option and not the
    default;
  * kNN indices are computed in exact fp32 arithmetic in every mode: still bit-exact.
The default (fp32-split) path is untouched by the switch: its results with precision="fp32" passed explicitly are
bit-identical to those without the argument."""
import os

import numpy as np
import pytest
import torch

from tests import parity
from tools import synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GOLD = os.path.join(os.path.dirname(__file__), "golden")
CD_FP16 = 1e-4          # stated end-to-end tolerance of the single-pass variant (see the module docstring)


def _net(precision):
    from modules.discrete.interpflow import PointInterpFlow
    n = PointInterpFlow(3)
    n.load_state_dict(torch.load(os.path.join(GOLD, "puflow-x4-pu1k.pt"), map_location="cpu"))
    n.set_to_initialized_state()
    n = n.to(DEV).eval()
    n.precision = precision
    return n


@pytest.mark.parametrize("tag", ["pu1k", "pugan", "pugeo"])
def test_single_pass_network_within_stated_tolerance(tag):
    from puflow_b200.engine import PUFlowEngine
    e = np.load(os.path.join(GOLD, f"net_{tag}.npz"))
    sd = torch.load(os.path.join(GOLD, f"puflow-x4-{tag}.pt"), map_location="cpu")
    xyz = torch.from_numpy(e["xyz"]).to(DEV)
    lo = PUFlowEngine(sd, DEV, precision="fp16").forward(xyz, 4, return_all=True)
    hi = PUFlowEngine(sd, DEV, precision="fp32").forward(xyz, 4, return_all=True)
    assert np.array_equal(lo["knn_idx"].cpu().numpy(), e["knn_idx"])
    ok, err = parity.close_norm(lo["x"].cpu().numpy(), e["x"], {"pu1k": 1e-2, "pugeo": 3e-2, "pugan": 1e-1}[tag])
    assert ok, err
    ok, err = parity.close_norm(lo["z"].cpu().numpy(), e["z"], 3e-1)
    assert ok, err
    ok_hi, _ = parity.close_norm(hi["x"].cpu().numpy(), e["x"], 1e-4)
    assert ok_hi
    # the variant really is a different arithmetic (not a silent alias of the split path)
    assert not torch.equal(lo["x"], hi["x"])
    # and the one-call C path takes the same switch
    xc, _ = PUFlowEngine(sd, DEV, precision="fp16").forward_c(xyz, 4)
    assert torch.equal(xc, lo["x"])


def test_single_pass_end_to_end_chamfer():
    from modules.utils.patch import PatchHelper
    e = np.load(os.path.join(GOLD, "e2e_2048.npz"))
    pc = torch.from_numpy(e["pc"]).to(DEV)
    net = _net("fp16")
    ph = PatchHelper(256, patch_expand_ratio=4)
    pred = ph.upsample(net, pc, npoint=2048 * 4 + 24, upratio=4, jitter=False)
    out = PatchHelper.remove_outliers(pred, pc, 24)
    cd = parity.chamfer_distance(out[0].cpu().numpy(), e["pred"][0])
    assert cd <= CD_FP16, cd
    net.precision = "fp32"                       # attribute switch re-packs the engine; default path again
    pred32 = ph.upsample(net, pc, npoint=2048 * 4 + 24, upratio=4, jitter=False)
    assert not torch.equal(pred32, pred)
    assert parity.chamfer_distance(pred32[0].cpu().numpy(), e["pred_raw"][0]) <= 1e-5


def test_precision_argument_is_checked_and_default_is_bit_identical():
    from puflow_b200 import ops
    from puflow_b200.engine import PUFlowEngine
    sd = torch.load(os.path.join(GOLD, "puflow-x4-pu1k.pt"), map_location="cpu")
    with pytest.raises(ValueError):
        PUFlowEngine(sd, DEV, precision="bf8")
    eng = PUFlowEngine(sd, DEV)
    xyz = torch.from_numpy(synth.make_patches(5, 256, seed=3, shape="surface")).to(DEV)
    a, _ = ops.upsample_patches(eng.model_blob(), xyz, 4)
    b, _ = ops.upsample_patches(eng.model_blob(), xyz, 4, precision=0)
    assert torch.equal(a, b)
    with pytest.raises(Exception):
        ops.upsample_patches(eng.model_blob(), xyz, 4, precision=7)
