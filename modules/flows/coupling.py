"""Coupling layers -- same classes / constructor signatures / child-module names (bias_net, scale_net)
as the reference's modules/flows/coupling.py:21-151, for the configurations PU-Flow's discrete model
instantiates (interpflow.py:55-63): the additive *spatial* coupling on 3 channels and the affine
*injector*, both with LinearA1D(… -> 64 -> 64 -> …) transform nets.  They run on the fused sm_100a flow
kernel; other configurations of the generic classes (chunk split, 'affineEx', the dead SplineCouplingLayer,
coupling.py:158-254) are outside the hot path and raise NotImplementedError."""
from typing import Dict

import torch
import torch.nn as nn
from torch import Tensor

from puflow_b200 import ops, packing
from modules.flows import _fused


# -----------------------------------------------------------------------------------------
class IdentityLayer(nn.Module):
    def forward(self, x: Tensor, **kwargs):
        return x


def _a1d_tensors(net):
    """(W0, W2, b2, W4, b4) of a LinearA1D-shaped transform net (interpflow.py:22-43)."""
    layers = getattr(net, "layers", None)
    if layers is None or len(layers) != 5 or layers[0].weight.shape[0] != 64 or layers[0].bias is not None:
        raise NotImplementedError("B200 flow kernels support LinearA1D transform nets with hidden width 64")
    return (layers[0].weight, layers[2].weight, layers[2].bias, layers[4].weight, layers[4].bias)


# -----------------------------------------------------------------------------------------
class AffineCouplingLayer(nn.Module):

    def __init__(self, coupling: str, transform_net: nn.Module, params: Dict, split_dim=1, clamp=None):
        super(AffineCouplingLayer, self).__init__()
        assert coupling in ['additive', 'affine', 'affineEx']
        assert split_dim in [-1, 1, 2, 3]
        self.coupling = coupling
        self.dim = split_dim
        self.bias_net = transform_net(**params)
        if self.coupling == 'affine':
            self.scale_net = transform_net(**params)
        elif self.coupling == 'affineEx':
            params_t = params
            params_t['in_channel'] = params['out_channel']
            params_t['out_channel'] = params['in_channel']
            self.g1 = transform_net(**params_t)
            self.g2 = transform_net(**params)
            self.g3 = transform_net(**params)
        if clamp is not None:
            raise NotImplementedError("clamp is None on the PU-Flow path (interpflow.py:57,62)")
        self.clamping = clamp or IdentityLayer()
        self._cache = _fused.BlobCache()

    def forward(self, x: Tensor, c: Tensor = None, **kwargs):
        raise NotImplementedError("generic chunk-split AffineCouplingLayer is not on PU-Flow's 3-channel path "
                                  "(interpflow.py:55 takes the idim == 3 branch)")

    def inverse(self, z: Tensor, c: Tensor = None, **kwargs):
        raise NotImplementedError("generic chunk-split AffineCouplingLayer is not on PU-Flow's 3-channel path")

    def channel_split(self, x: Tensor):
        return torch.chunk(x, 2, dim=self.dim)

    def channel_cat(self, h1: Tensor, h2: Tensor):
        return torch.cat([h1, h2], dim=self.dim)


# -----------------------------------------------------------------------------------------
class AffineSpatialCouplingLayer(AffineCouplingLayer):
    """h2 -/+= bias_net(cat[h1, c]) with split [1,2] (even) / [2,1] (odd) -- reference coupling.py:108-118,55-58,82-85."""

    def __init__(self, coupling, transform_net, params, is_even, split_dim, clamp=None):
        super().__init__(coupling, transform_net, params, split_dim=split_dim, clamp=clamp)
        self.is_even = is_even

    def channel_split(self, x: Tensor):
        if self.is_even:
            return torch.split(x, [1, 2], dim=self.dim)
        else:
            return torch.split(x, [2, 1], dim=self.dim)

    def _packed(self, device):
        t = _a1d_tensors(self.bias_net)

        def build():
            flow = packing.identity_flow_blob()
            w1c = packing.set_coupling1(flow, *t, 1 if self.is_even else 2)
            wc = torch.zeros(w1c.shape[0], 192, dtype=torch.float64)
            wc[:, :64] = w1c
            return _fused.to_dev(flow, device), _fused.to_dev(wc, device)
        return self._cache.get(t, device, build)

    def _run(self, x: Tensor, c: Tensor, inverse: bool):
        if self.coupling != 'additive' or self.dim not in (2, -1):
            raise NotImplementedError("AffineSpatialCouplingLayer on B200 kernels: 'additive', channel-last")
        x = _fused.check_input(x, "AffineSpatialCouplingLayer")
        flow, wc = self._packed(x.device)
        M = x.shape[0] * x.shape[1]
        raw = ops.linear(c.reshape(M, -1).contiguous().float(), wc).view(1, M, 192)
        y, _ = _fused.run(x, flow, inverse, raw=raw)
        return y, None

    def forward(self, x: Tensor, c: Tensor = None, **kwargs):
        return self._run(x, c, False)

    def inverse(self, z: Tensor, c: Tensor = None, **kwargs):
        return self._run(z, c, True)


# -----------------------------------------------------------------------------------------
class AffineInjectorLayer(AffineCouplingLayer):
    """x = (x - bias_net(c)) * exp(-scale_net(c)); inverse z * exp(scale) + bias -- reference coupling.py:127-151."""

    def __init__(self, coupling, transform_net, params, split_dim, clamp=None):
        super().__init__(coupling, transform_net, params, split_dim=split_dim, clamp=clamp)

    def _packed(self, device):
        ts, tb = _a1d_tensors(self.scale_net), _a1d_tensors(self.bias_net)

        def build():
            cond = packing.zero_cond_blob()
            ws0, wb0 = packing.set_injector(cond, ts, tb)
            wc = torch.zeros(ws0.shape[0], 192, dtype=torch.float64)
            wc[:, 64:128], wc[:, 128:192] = ws0, wb0
            return (_fused.to_dev(packing.identity_flow_blob(), device), _fused.to_dev(cond, device),
                    _fused.to_dev(wc, device))
        return self._cache.get(ts + tb, device, build)

    def _run(self, x: Tensor, c: Tensor, inverse: bool):
        if self.coupling != 'affine' or self.dim not in (2, -1):
            raise NotImplementedError("AffineInjectorLayer on B200 kernels: 'affine', channel-last")
        x = _fused.check_input(x, "AffineInjectorLayer")
        flow, cond, wc = self._packed(x.device)
        B, N, _ = x.shape
        raw = ops.linear(c.reshape(B * N, -1).contiguous().float(), wc).view(1, B * N, 192)
        sb = ops.flow_cond(raw, cond.view(1, -1))
        y, ssum = _fused.run(x, flow, inverse, raw=raw, sb=sb)
        s_sum = sb[0, :, 0:3].reshape(B, -1).sum(1)
        return y, (s_sum if inverse else -s_sum)

    def forward(self, x: Tensor, c: Tensor, **kwargs):
        return self._run(x, c, False)

    def inverse(self, z: Tensor, c: Tensor, **kwargs):
        return self._run(z, c, True)
# -----------------------------------------------------------------------------------------
