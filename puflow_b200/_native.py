"""ctypes binding of the C-ABI in include/puflow_b200.h.

The shared library is built in-tree (puflow_b200/_lib/libpuflow_b200.so) by
`make -C puflow_b200/csrc` / `__graft_entry__.build()`.  Importing this module without the library
raises ImportError -- there is deliberately no fallback implementation.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libpuflow_b200.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build the CUDA extension first (python -c 'import __graft_entry__ as g; g.build()' "
        f"or make -C {os.path.join(_HERE, 'csrc')}). puflow_b200 has no CPU / PyTorch fallback."
    )

lib = ctypes.CDLL(LIB_PATH)

_vp, _i, _i64, _f = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float

# name -> argtypes; every function returns int status except pf_last_error
SIGNATURES = {
    "pf_abi_version": [],
    "pf_knn_topk": [_vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp],
    "pf_chamfer_nn": [_vp, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp],
    "pf_linear": [_vp, _i64, _vp, _vp, _vp, _i64, _i64, _i, _i, _i, _f, _vp],
    "pf_linear_tc": [_vp, _i64, _vp, _vp, _vp, _i64, _i64, _i, _i, _i, _f, _i, _vp],
    "pf_edgeconv": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _vp],
    "pf_edgeconv_tc": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _i, _vp],
    "pf_edgeconv_tc_split": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _i, _vp],
    "pf_linear_tc_split": [_vp, _vp, _vp, _vp, _i64, _i64, _i, _i, _i, _f, _i, _vp],
    "pf_flow_cond": [_vp, _vp, _vp, _i64, _i, _vp],
    "pf_flow_fwd": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i, _vp],
    "pf_flow_inv": [_vp, _vp, _vp, _vp, _vp, _i64, _i, _i, _vp],
    "pf_flow_cond_tc": [_vp, _vp, _vp, _i64, _i, _i, _vp],
    "pf_flow_fwd_tc": [_vp, _vp, _i, _vp, _vp, _vp, _vp, _i64, _i, _i, _vp],
    "pf_flow_inv_tc": [_vp, _vp, _i, _vp, _vp, _vp, _i64, _i, _i, _i, _vp],
    "pf_cond_chain_tc": [_vp, _i, _vp, _vp, _vp, _i64, _i, _vp],
    "pf_interp": [_vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _i, _i, _vp],
    "pf_interp_tc": [_vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _vp],
    "pf_cnf_flow": [_vp, _vp, _i, _vp, _vp, _vp, _i, _i, _f, _f, _f, _f, _vp],
    "pf_fps": [_vp, _i, _i, _i, _vp, _vp, _vp],
    "pf_fps_bucketed": [_vp, _i, _i, _i, _vp, _vp, _vp],
    "pf_fps_clustered": [_vp, _i, _i, _i, _vp, _vp, _vp],
    "pf_fps_queued": [_vp, _i, _i, _i, _vp, _vp, _vp],
    "pf_fps_queued_max_points": [],
    "pf_fps_queued_static": [_vp, _i, _i, _i, _vp, _vp],
    "pf_knn_patch": [_vp, _vp, _i, _i, _i, _i, _vp, _vp],
    "pf_xyz_count": [ctypes.c_char_p, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int)],
    "pf_xyz_read": [ctypes.c_char_p, _vp, _i64, _i],
    "pf_xyz_write": [ctypes.c_char_p, _vp, _i64, _i],
    "pf_model_entry": [_i, ctypes.c_char_p, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int64)],
    "pf_upsample_patches": [_vp, _vp, _i, _i, _i, _vp, _i64, _vp, _vp, _i, _vp],
    "pf_normalize_pc": [_vp, _i, _i, _vp, _vp, _vp, _vp],
    "pf_gather_normalize": [_vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp],
    "pf_gather_points": [_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp],
    "pf_denorm_concat": [_vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp],
    "pf_remove_outliers": [_vp, _vp, _i, _i, _i, _vp, _vp],
    "pf_fps_park_near_origin": [_vp, _i, _i, _f, _vp, _vp],
    "pf_upsample_clouds": [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _i64, _vp, _i, _vp],
    "pf_profile_enable": [_i],
    "pf_profile_intervals": [],
    "pf_profile_interval": [_i, ctypes.c_char_p, ctypes.POINTER(ctypes.c_float)],
    "pf_profile_clear": [],
}
for _name, _args in SIGNATURES.items():
    _fn = getattr(lib, _name)
    _fn.argtypes = _args
    _fn.restype = ctypes.c_int
lib.pf_fps_work_floats.argtypes = [_i, _i]
lib.pf_fps_work_floats.restype = ctypes.c_longlong
lib.pf_fps_clustered_work_floats.argtypes = [_i, _i]
lib.pf_fps_clustered_work_floats.restype = ctypes.c_longlong
lib.pf_model_bytes.argtypes = []
lib.pf_model_bytes.restype = ctypes.c_int64
lib.pf_upsample_workspace_bytes.argtypes = [_i, _i, _i]
lib.pf_upsample_workspace_bytes.restype = ctypes.c_int64
lib.pf_upsample_clouds_workspace_bytes.argtypes = [_i, _i, _i, _i, _i, _i]
lib.pf_upsample_clouds_workspace_bytes.restype = ctypes.c_int64
lib.pf_last_error.argtypes = []
lib.pf_last_error.restype = ctypes.c_char_p

PF_ABI_VERSION = 10
PF_MODEL_ENTRIES = 27
PF_FPS_STATIC_MAX_POINTS = 262144
PF_COND_BLOB = 8840
PF_FLOW_BLOB = 4480
PF_COND_TC_BLOCK = 35360
PF_FLOW_TC_BLOCK = 17920


def chain_tc_bytes(C: int) -> int:
    return 2 * C * C + 2 * 192 * C + 4 * 8192 + (64 + 648) * 4

PF_INTERP_BLOB = 46304
PF_EDGECONV_TC_BLOB = 90112
PF_INTERP_TC_BLOB = 205440
PF_CNF_CTX = 264
PF_CNF_BLOB = 19520
ACT_NONE, ACT_RELU, ACT_LRELU = 0, 1, 2
PREC_FP32_SPLIT, PREC_FP16 = 0, 1          # PF_PREC_*: precision argument of the tensor-core entry points
PRECISIONS = {"fp32": PREC_FP32_SPLIT, "fp32_split": PREC_FP32_SPLIT, "fp16": PREC_FP16, 0: 0, 1: 1}

if lib.pf_abi_version() != PF_ABI_VERSION:
    raise ImportError(f"{LIB_PATH}: ABI version {lib.pf_abi_version()} != {PF_ABI_VERSION}; rebuild")


def model_entries():
    """[(name, offset, bytes)] of the packed-model blob, in blob order (the layout lives in csrc/model.cu)."""
    out = []
    for i in range(PF_MODEL_ENTRIES):
        name = ctypes.create_string_buffer(32)
        off, nbytes = ctypes.c_int64(), ctypes.c_int64()
        if lib.pf_model_entry(i, name, ctypes.byref(off), ctypes.byref(nbytes)) != 0:
            raise ImportError(f"pf_model_entry({i}) failed: {lib.pf_last_error().decode()}")
        out.append((name.value.decode(), off.value, nbytes.value))
    return out


class PuflowError(RuntimeError):
    pass


def check(rc: int, what: str):
    if rc != 0:
        raise PuflowError(f"{what} failed (status {rc}): {lib.pf_last_error().decode(errors='replace')}")
