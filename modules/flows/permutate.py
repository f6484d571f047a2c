"""Permutation layers -- same classes / buffers / parameters as the reference's
modules/flows/permutate.py:15-126.  A channel permutation of 3 channels is a 3x3 0/1 matrix, so both the
shuffles and the invertible 1x1 conv run through the matrix slot of the fused sm_100a flow kernel."""
import numpy as np
import torch
import torch.nn as nn
from torch import Tensor

from puflow_b200 import packing
from modules.flows import _fused


# -----------------------------------------------------------------------------------------
class Permutation(nn.Module):

    def __init__(self, permutation: str, n_channel: int, dim: int):
        super(Permutation, self).__init__()
        assert permutation in ['reverse', 'random', 'inv1x1']
        assert dim in [-1, 1, 2, 3]
        if permutation == 'inv1x1':
            self.permutater = InvertibleConv1x1_1D(n_channel, dim)
        else:
            if dim == -1: self.permutater = _ShufflePermutationXD(permutation, n_channel)
            if dim ==  1: self.permutater = _ShufflePermutation1D(permutation, n_channel)
            if dim ==  2: self.permutater = _ShufflePermutation2D(permutation, n_channel)
            if dim ==  3: self.permutater = _ShufflePermutation3D(permutation, n_channel)

    def forward(self, x: Tensor, _c: Tensor):
        return self.permutater(x)

    def inverse(self, z: Tensor, _c: Tensor):
        return self.permutater.inverse(z)


# -----------------------------------------------------------------------------------------
class _ShufflePermutation(nn.Module):
    _channel_dim = None     # which tensor dim the subclass permutes

    def __init__(self, permutation: str, n_channel: int):
        super(_ShufflePermutation, self).__init__()
        if permutation == 'reverse':
            direct_idx = np.arange(n_channel - 1, -1, -1).astype(np.int64)
            inverse_idx = _ShufflePermutation.get_reverse(direct_idx, n_channel)
        if permutation == 'random':
            direct_idx = np.arange(n_channel - 1, -1, -1).astype(np.int64)
            np.random.shuffle(direct_idx)
            inverse_idx = _ShufflePermutation.get_reverse(direct_idx, n_channel)
        self.register_buffer('direct_idx', torch.from_numpy(direct_idx))
        self.register_buffer('inverse_idx', torch.from_numpy(inverse_idx))
        self._cache = _fused.BlobCache()

    @staticmethod
    def get_reverse(idx, n_channel: int):
        indices_inverse = np.zeros((n_channel,), dtype=np.int64)
        for i in range(n_channel):
            indices_inverse[idx[i]] = i
        return indices_inverse

    def _blob(self, device):
        def build():
            P = torch.zeros(3, 3, dtype=torch.float64)
            P[torch.arange(3), self.direct_idx.cpu()] = 1.0        # (P x)_i = x[direct_idx[i]]
            flow = packing.identity_flow_blob()
            packing.set_matrix(flow, P)
            return _fused.to_dev(flow, device)
        return self._cache.get((self.direct_idx,), device, build)

    def _run(self, x: Tensor, inverse: bool):
        if self._channel_dim not in (2, -1) or self.direct_idx.numel() != 3 or x.dim() != 3:
            raise NotImplementedError("channel shuffles on B200 kernels: channel-last [B, N, 3] only")
        x = _fused.check_input(x, type(self).__name__)
        return _fused.run(x, self._blob(x.device), inverse)[0], None

    def forward(self, x: Tensor):
        return self._run(x, False)

    def inverse(self, z: Tensor):
        return self._run(z, True)


class _ShufflePermutation1D(_ShufflePermutation):
    _channel_dim = 1

class _ShufflePermutation2D(_ShufflePermutation):
    _channel_dim = 2

class _ShufflePermutation3D(_ShufflePermutation):
    _channel_dim = 3

class _ShufflePermutationXD(_ShufflePermutation):
    _channel_dim = -1


# -----------------------------------------------------------------------------------------
class InvertibleConv1x1_1D(nn.Module):
    """z = x W^T (einsum 'ij,bnj->bni', reference permutate.py:112,118); the inverse matrix is computed
    once per parameter version in float64 instead of torch.inverse per call (:123)."""

    def __init__(self, channel: int, dim: int):
        super(InvertibleConv1x1_1D, self).__init__()
        w_init = np.random.randn(channel, channel)
        w_init = np.linalg.qr(w_init)[0].astype(np.float32)
        self.W = nn.Parameter(torch.from_numpy(w_init), requires_grad=True)
        if dim == 1:
            self.equation = 'ij,bjn->bin'
            self.Ndim = 2
        elif dim == 2 or dim == -1:
            self.equation = 'ij,bnj->bni'
            self.Ndim = 1
        else:
            raise NotImplementedError(f"Unsupport dim {dim} for InvertibleConv1x1 Layer.")
        self._cache = _fused.BlobCache()

    def _blob(self, device):
        def build():
            flow = packing.identity_flow_blob()
            packing.set_matrix(flow, self.W)
            return _fused.to_dev(flow, device)
        return self._cache.get((self.W,), device, build)

    def _check(self, x):
        if self.Ndim != 1 or self.W.shape != (3, 3):
            raise NotImplementedError("InvertibleConv1x1_1D on B200 kernels: channel-last with 3 channels only")
        return _fused.check_input(x, "InvertibleConv1x1_1D")

    def forward(self, x: Tensor):
        x = self._check(x)
        z, _ = _fused.run(x, self._blob(x.device), inverse=False)
        logdet = torch.slogdet(self.W)[1] * x.shape[self.Ndim]
        return z, logdet

    def inverse(self, z: Tensor):
        z = self._check(z)
        x, _ = _fused.run(z, self._blob(z.device), inverse=True)
        logdet = -torch.slogdet(self.W)[1] * x.shape[self.Ndim]
        return x, logdet
# -----------------------------------------------------------------------------------------
