"""Reference harness for the CNF variant's dynamics -- TEST INFRASTRUCTURE, runs ONLY in the build container.

Loads the *unmodified* /root/reference/modules/continuous/{diffeq_layers,odefunc}.py (they import only torch), builds the
reference's ODEnet exactly as modules/continuous/interpflow.py:27 does (hidden (64, 64), input (3,), context_dim = cond
channels of the block, layer_type 'concatsquash', nonlinearity 'tanh'), loads flow_blocks.{0,1,2}.cnf.odefunc.diffeq.* of
pretrain/puflow-x4-cnf-pu1k.pt, evaluates it on seeded inputs and writes tests/golden/cnf_dynamics.npz.  The integrator
(torchdiffeq) is not in the reference tree and not installed: only the dynamics can be pinned (oracle/cnf_oracle.py).
Run by file path, with /root/repo NOT on sys.path:  python oracle/ref/run_reference_cnf.py
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("PUFLOW_REFERENCE", "/root/reference")
REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def load_pkg():
    pkg = types.ModuleType("refcont")
    pkg.__path__ = [os.path.join(REF, "modules", "continuous")]
    sys.modules["refcont"] = pkg
    for name in ("diffeq_layers", "odefunc"):
        spec = importlib.util.spec_from_file_location(f"refcont.{name}", os.path.join(REF, "modules", "continuous", name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[f"refcont.{name}"] = mod
        spec.loader.exec_module(mod)
    return sys.modules["refcont.odefunc"]


def main():
    odefunc = load_pkg()
    sd = torch.load(os.path.join(REF, "pretrain", "puflow-x4-cnf-pu1k.pt"), map_location="cpu")
    out = {}
    g = torch.Generator().manual_seed(20260)
    for blk, cdim in ((0, 32), (1, 64), (2, 128)):
        net = odefunc.ODEnet((64, 64), input_shape=(3,), context_dim=cdim, layer_type="concatsquash", nonlinearity="tanh")
        p = f"flow_blocks.{blk}.cnf.odefunc.diffeq."
        net.load_state_dict({k[len(p):]: v for k, v in sd.items() if k.startswith(p)})
        net.eval()
        M = 96
        y = torch.randn(1, M, 3, generator=g)
        c = torch.randn(1, M, cdim, generator=g) * 0.5
        t = torch.rand(1, M, 1, generator=g) * 0.35
        with torch.no_grad():
            dy = net(torch.cat([t, c], dim=-1), y)          # ODEfunc.forward's conditional branch, odefunc.py:136-140
        out[f"y{blk}"], out[f"c{blk}"], out[f"t{blk}"], out[f"dy{blk}"] = (a[0].numpy() for a in (y, c, t, dy))
    dst = os.path.join(REPO, "tests", "golden", "cnf_dynamics.npz")
    np.savez_compressed(dst, **out)
    print("wrote", dst, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
