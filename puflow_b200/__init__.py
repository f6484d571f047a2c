"""puflow_b200 -- B200-native (sm_100a) kernels for PU-Flow's discrete x4 upsampling hot path.

Layout:
  csrc/        hand-written CUDA kernels + the C-ABI (include/puflow_b200.h)
  _native.py   ctypes binding of libpuflow_b200.so (fails loudly when the library is missing)
  ops.py       torch.Tensor-checked wrappers around the C-ABI entry points
  packing.py   one-time weight packing from the reference's 408-key checkpoint layout
  engine.py    the hot path (PointInterpFlow.forward) as a sequence of kernel launches

There is no CPU fallback: every op raises if the tensors are not on a CUDA device or the native
library cannot be loaded.
"""
from . import _native  # noqa: F401  (raises ImportError if the shared library is absent)

__all__ = ["ops", "packing", "engine"]
