"""pf_fps_queued_static: the allocation-free variant of pf_fps_queued (scratch in a static __device__ array of the
library; no counterpart in the reference).  Called through the C-ABI and
compared with the existing entry point, bit for bit (index work), on two calls in a row that share the static scratch.
Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest
import torch

from tools import synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _call_static(lib, x, npoint):
    B, n, _ = x.shape
    idx = torch.full((B, npoint), -1, dtype=torch.int32, device=x.device)      # inputs / outputs stay the caller's
    with torch.cuda.device(x.device):
        rc = lib.pf_fps_queued_static(x.data_ptr(), B, n, npoint, idx.data_ptr(), torch.cuda.current_stream().cuda_stream)
    return rc, idx


def test_fps_queued_static_matches_pf_fps_queued_twice():
    from puflow_b200 import ops
    from puflow_b200._native import lib, PF_FPS_STATIC_MAX_POINTS
    # call 1: three 40,960-candidate clouds (the merge FPS of config 2); call 2: smaller, ragged clouds over the stale scratch
    p1 = np.stack([synth.make_cloud(40960, 41 + b, synth.SHAPES[b % 3]) for b in range(3)])
    p2 = np.random.default_rng(3).uniform(-1, 1, (5, 7001, 3)).astype(np.float32)
    for p, npoint in ((p1, 2500), (p2, 700)):
        x = torch.from_numpy(np.ascontiguousarray(p)).to(DEV)
        ref = ops.fps(x, npoint, mode="queue")                  # the existing entry point, caller-owned scratch
        rc, out = _call_static(lib, x, npoint)
        torch.cuda.synchronize()
        assert rc == 0, lib.pf_last_error().decode()
        assert torch.equal(out, ref)
    # beyond the static array: a clear error, nothing launched
    n_big = PF_FPS_STATIC_MAX_POINTS // 2 + 32
    xb = torch.zeros((2, n_big, 3), device=DEV)
    rc, _ = _call_static(lib, xb, 4)
    assert rc != 0 and "PF_FPS_STATIC_MAX_POINTS" in lib.pf_last_error().decode()
