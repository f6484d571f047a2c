"""Runs ONE modules/flows layer on the fused flow kernel (puflow_b200/csrc/flow.cu): the layer's
parameters are placed in an otherwise-identity flow-step blob (puflow_b200.packing.identity_flow_blob),
so ActNorm, the 1x1 conv, the channel permutations and both coupling layers all execute the same
hand-written sm_100a code as the fused PointInterpFlow path.  Tensors must be [B, N, 3] fp32 on CUDA."""
import torch

from puflow_b200 import ops, packing


def check_input(x: torch.Tensor, who: str):
    if x.dim() != 3 or x.shape[-1] != 3:
        raise NotImplementedError(f"{who}: the B200 kernels implement the channel-last [B, N, 3] point layout "
                                  f"used by PointInterpFlow (got {tuple(x.shape)})")
    if not x.is_cuda:
        raise RuntimeError(f"{who}: tensor is on {x.device}; puflow_b200 has no CPU fallback")
    return x.contiguous().float()


class BlobCache:
    """Re-packs a layer's blob only when one of its parameters changed (or moved device).
    The key is (storage pointer, version counter) per tensor: in-place writes through `p.data` (or `p.data.copy_`) do NOT
    bump the version counter, so code that edits weights that way (EMA swaps, manual re-initialisation) must call
    invalidate() -- or write with `with torch.no_grad(): p.copy_(...)`, which does bump it."""

    def __init__(self):
        self.key, self.value = None, None

    def get(self, tensors, device, build, extra=()):
        key = (str(device),) + tuple(extra) + tuple((t.data_ptr(), t._version) for t in tensors)
        if key != self.key:
            self.key, self.value = key, build()
        return self.value

    def invalidate(self):
        self.key, self.value = None, None


def to_dev(t, device):
    return t.to(torch.float32).contiguous().to(device)


def run(x: torch.Tensor, flow_blob: torch.Tensor, inverse: bool, raw=None, sb=None):
    """x[B,N,3] through one flow step described by flow_blob[PF_FLOW_BLOB].  Returns (y[B,N,3], ssum[B*N] or None)."""
    B, N, _ = x.shape
    M = B * N
    if raw is None:
        raw = torch.zeros((1, M, 192), dtype=torch.float32, device=x.device)
    if sb is None:
        sb = torch.zeros((1, M, 8), dtype=torch.float32, device=x.device)
    fw = flow_blob.view(1, -1)
    if inverse:
        return ops.flow_inv(x.reshape(M, 3), raw, sb, fw, 1).view(B, N, 3), None
    y, ssum = ops.flow_fwd(x.reshape(M, 3), raw, sb, fw)
    return y.view(B, N, 3), ssum
