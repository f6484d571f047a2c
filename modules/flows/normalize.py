"""ActNorm -- same class / constructor / parameters / return values as the reference's
modules/flows/normalize.py:9-54, computing on the fused sm_100a flow kernel."""
import torch
import torch.nn as nn
from torch import Tensor

from puflow_b200 import packing
from modules.flows import _fused


# -----------------------------------------------------------------------------------------
class ActNorm(nn.Module):
    """z = x * exp(logs) + bias (reference normalize.py:30-37); inverse (z - bias) * exp(-logs) (:39-43)."""

    def __init__(self, channel: int, dim=1):
        super(ActNorm, self).__init__()
        assert dim in [-1, 1, 2]
        self.dim = 2 if dim == -1 else dim
        if self.dim == 1:
            self.logs = nn.Parameter(torch.zeros((1, channel, 1)))
            self.bias = nn.Parameter(torch.zeros((1, channel, 1)))
            self.Ndim = 2
        if self.dim == 2:
            self.logs = nn.Parameter(torch.zeros((1, 1, channel)))
            self.bias = nn.Parameter(torch.zeros((1, 1, channel)))
            self.Ndim = 1
        self.eps = 1e-6
        self.is_inited = False
        self._cache = _fused.BlobCache()

    def _blob(self, device):
        def build():
            flow = packing.identity_flow_blob()
            packing.set_actnorm(flow, self.logs, self.bias)
            return _fused.to_dev(flow, device)
        return self._cache.get((self.logs, self.bias), device, build)

    def _check(self, x):
        if self.dim != 2 or self.logs.numel() != 3:
            raise NotImplementedError("ActNorm on B200 kernels: channel-last (dim=2) with 3 channels only")
        return _fused.check_input(x, "ActNorm")

    def forward(self, x: Tensor, _: Tensor = None):
        x = self._check(x)
        if not self.is_inited:
            self.__initialize(x)
        z, _ss = _fused.run(x, self._blob(x.device), inverse=False)
        logdet = torch.sum(self.logs) * x.shape[self.Ndim]
        return z, logdet

    def inverse(self, z: Tensor, _: Tensor = None):
        z = self._check(z)
        x, _ss = _fused.run(z, self._blob(z.device), inverse=True)
        logdet = -torch.sum(self.logs) * x.shape[self.Ndim]
        return x, logdet

    def __initialize(self, x: Tensor):
        # data-dependent initialisation (reference normalize.py:45-54); one-time parameter set-up
        with torch.no_grad():
            dims = [0, 1, 2]
            dims.remove(self.dim)
            bias = -torch.mean(x.detach(), dim=dims, keepdim=True)
            logs = -torch.log(torch.std(x.detach(), dim=dims, keepdim=True) + self.eps)
            self.bias.copy_(bias)               # under no_grad: bumps the version counter the blob cache keys on
            self.logs.copy_(logs)               # (a write through `.data` would leave a stale packed blob behind)
            self._cache.invalidate()
            self.is_inited = True
# -----------------------------------------------------------------------------------------
