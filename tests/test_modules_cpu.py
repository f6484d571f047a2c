"""CPU: host-side logic of the drop-in modules (no kernel launches): state_dict layout, API surface,
loud failure without CUDA, and that the C-ABI library exports every symbol include/puflow_b200.h declares."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "puflow_b200.h")).read()
    declared = set(re.findall(r"PF_API\s+(?:const\s+char\s*\*|int|int64_t|long\s+long)\s+(pf_\w+)\s*\(", hdr))
    assert len(declared) >= 12, declared
    lib = ctypes.CDLL(os.path.join(ROOT, "puflow_b200", "_lib", "libpuflow_b200.so"))
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    from puflow_b200 import _native
    assert set(_native.SIGNATURES) | {"pf_last_error", "pf_fps_work_floats", "pf_fps_clustered_work_floats", "pf_model_bytes", "pf_upsample_clouds_workspace_bytes",
                                      "pf_upsample_workspace_bytes"} == declared
    assert lib.pf_abi_version() == _native.PF_ABI_VERSION


def test_header_constants_match_binding():
    hdr = open(os.path.join(ROOT, "include", "puflow_b200.h")).read()
    from puflow_b200 import _native
    for name in ("PF_COND_BLOB", "PF_FLOW_BLOB", "PF_INTERP_BLOB", "PF_ABI_VERSION", "PF_COND_TC_BLOCK",
                 "PF_FLOW_TC_BLOCK", "PF_EDGECONV_TC_BLOB", "PF_INTERP_TC_BLOB", "PF_MODEL_ENTRIES", "PF_CNF_CTX", "PF_CNF_BLOB"):
        assert int(re.search(rf"#define {name} (\d+)", hdr).group(1)) == getattr(_native, name)


def test_state_dict_layout_matches_checkpoint(state_dict):
    from modules.discrete.interpflow import PointInterpFlow
    net = PointInterpFlow(3)
    own = net.state_dict()
    assert len(own) == 408 and list(own.keys()) == list(state_dict.keys())
    for k in own:
        assert own[k].shape == state_dict[k].shape and own[k].dtype == state_dict[k].dtype, k
    res = net.load_state_dict(state_dict)
    assert not res.missing_keys and not res.unexpected_keys
    net.set_to_initialized_state()
    assert all(fb.actnorm.is_inited for fb in net.flow_blocks)
    assert net.num_blocks == 6 and net.num_neighbors == 16 and net.interp.k == 8


def test_no_cpu_fallback(state_dict):
    from modules.discrete.interpflow import PointInterpFlow
    from modules.flows.normalize import ActNorm
    from modules.utils.patch import PatchHelper
    from puflow_b200 import ops
    from puflow_b200._native import PuflowError
    net = PointInterpFlow(3)
    net.load_state_dict(state_dict)
    net.set_to_initialized_state()
    x = torch.rand(2, 256, 3)
    with pytest.raises(RuntimeError):
        net(x)
    with pytest.raises(RuntimeError):
        ActNorm(3, dim=2)(x)
    with pytest.raises(PuflowError):
        ops.chamfer_nn(x, x)
    with pytest.raises(PuflowError):
        PatchHelper.remove_outliers(x, x, 4)


def test_flow_layer_api_surface():
    from modules.flows.coupling import AffineCouplingLayer, AffineInjectorLayer, AffineSpatialCouplingLayer
    from modules.flows.normalize import ActNorm
    from modules.flows.permutate import InvertibleConv1x1_1D, Permutation
    from modules.discrete.interpflow import FlowBlock, LinearA1D
    a = ActNorm(3, dim=2)
    assert a.logs.shape == (1, 1, 3) and a.bias.shape == (1, 1, 3) and a.eps == 1e-6 and not a.is_inited
    p = Permutation('reverse', 3, dim=2)
    assert p.permutater.direct_idx.tolist() == [2, 1, 0] and p.permutater.inverse_idx.tolist() == [2, 1, 0]
    assert Permutation('inv1x1', 3, dim=2).permutater.W.shape == (3, 3)
    assert isinstance(Permutation('inv1x1', 3, dim=2).permutater, InvertibleConv1x1_1D)
    fb = FlowBlock(idim=3, hdim=64, cdim=32, is_even=True)
    keys = set(fb.state_dict().keys())
    assert "coupling1.bias_net.layers.0.weight" in keys and "coupling2.scale_net.layers.4.bias" in keys
    assert fb.coupling1.bias_net.layers[0].weight.shape == (64, 33)
    assert isinstance(fb.coupling1, AffineSpatialCouplingLayer) and isinstance(fb.coupling2, AffineInjectorLayer)
    assert issubclass(AffineSpatialCouplingLayer, AffineCouplingLayer)
    assert LinearA1D(2, 64, 1, 128).layers[4].weight.abs().sum() == 0          # zero-initialised last layer


def test_upsample_entry_point_signature():
    import inspect
    from modules.discrete import upsample
    params = list(inspect.signature(upsample.upsampling).parameters)
    assert params[:8] == ["data_paths", "target_path", "checkpoint_path", "up_ratio", "num_outlier", "num_patch",
                          "num_upsampling", "seed"]


def _build_c_host(tmp_path):
    import subprocess
    exe = str(tmp_path / "upsample_patches")
    libdir = os.path.join(ROOT, "puflow_b200", "_lib")
    cmd = ["gcc", "-O2", "-I" + os.path.join(ROOT, "include"), "-I/usr/local/cuda/include",
           os.path.join(ROOT, "examples", "upsample_patches.c"), "-o", exe, "-L" + libdir, "-lpuflow_b200",
           "-L/usr/local/cuda/lib64", "-lcudart", "-Wl,-rpath," + libdir]
    subprocess.run(cmd, check=True, capture_output=True)
    return exe


def test_header_is_plain_c_and_the_c_host_links(tmp_path):
    """include/puflow_b200.h compiles as C and examples/upsample_patches.c (a host with no Python) links against the
    library; without arguments it prints its usage and exits 1 (no GPU needed)."""
    import subprocess
    exe = _build_c_host(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 1 and "usage" in r.stderr
