"""Drop-in for /root/reference/modules/continuous/interpflow.py (the CNF variant of PU-Flow), inference only.

Same class names, constructor arguments, parameter / buffer names (the 390 keys of pretrain/puflow-x4-cnf-pu1k.pt load with
strict=True) and call signatures as the reference; the compute runs on the B200 kernels: the conditioner and the interpolation
module are the discrete model's (reference continuous/interpflow.py:15 imports them from the discrete file), the flow blocks are
integrated by pf_cnf_flow (puflow_b200/csrc/cnf.cu).  What is different, and stated in DESIGN.md:
  * log p is not computed: forward() returns (x, None).  The reference estimates the divergence with a random Hutchinson probe
    through autograd (modules/continuous/odefunc.py:8-32); sample(), the inference entry point, discards it;
  * the adaptive Dormand-Prince solver controls every point's step on its own and ends exactly at T (the reference calls
    torchdiffeq.odeint, a dependency that is not in its tree: batch-wide error norm, interpolation at T);
  * no training: parameters are read, never differentiated.
"""
from typing import List, Tuple

import torch
import torch.nn as nn
from torch import Tensor

from modules.discrete.interpflow import FeatMergeUnit, FeatureExtractUnit, InterpolationModule, _idx32, check_input
from modules.flows._fused import BlobCache
from modules.utils.probs import GaussianDistribution
from puflow_b200.engine import CNFEngine


class ConcatSquashLinear(nn.Module):
    """diffeq_layers.py:72-86: layer(x) * sigmoid(hyper_gate([t, c])) + hyper_bias([t, c]).  Parameter holder."""

    def __init__(self, dim_in: int, dim_out: int, dim_c: int):
        super().__init__()
        self._layer = nn.Linear(dim_in, dim_out)
        self._hyper_bias = nn.Linear(1 + dim_c, dim_out, bias=False)
        self._hyper_gate = nn.Linear(1 + dim_c, dim_out)


class ODEnet(nn.Module):
    """odefunc.py:66-111 with layer_type='concatsquash', nonlinearity='tanh' (what FlowBlock builds)."""

    def __init__(self, hidden_dims: Tuple[int, ...], input_shape: Tuple[int, ...], context_dim: int, layer_type: str = "concatsquash",
                 nonlinearity: str = "tanh"):
        super().__init__()
        if layer_type != "concatsquash" or nonlinearity != "tanh" or tuple(hidden_dims) != (64, 64) or tuple(input_shape) != (3,):
            raise NotImplementedError("pf_cnf_flow implements the ODEnet the checkpoint holds: concatsquash, tanh, hidden (64, 64), 3 channels")
        dims = (input_shape[0],) + tuple(hidden_dims) + (input_shape[0],)
        self.layers = nn.ModuleList(ConcatSquashLinear(dims[k], dims[k + 1], context_dim) for k in range(3))


class ODEfunc(nn.Module):
    """odefunc.py:114-148.  Parameter / buffer holder (`_num_evals` is part of the checkpoint)."""

    def __init__(self, diffeq: ODEnet):
        super().__init__()
        self.diffeq = diffeq
        self.register_buffer("_num_evals", torch.tensor(0.))


class CNF(nn.Module):
    """cnf.py:33-50: T = sqrt_end_time ** 2 (train_T=True), solver dopri5, atol = rtol = 1e-5."""

    def __init__(self, odefunc: ODEfunc, conditional: bool = True, T: float = 1.0, train_T: bool = False, solver: str = "dopri5",
                 atol: float = 1e-5, rtol: float = 1e-5):
        super().__init__()
        if solver != "dopri5" or not conditional:
            raise NotImplementedError("pf_cnf_flow implements the conditional dopri5 CNF")
        self.train_T, self.T = train_T, T
        if train_T:
            self.register_parameter("sqrt_end_time", nn.Parameter(torch.sqrt(torch.tensor(T))))
        self.odefunc = odefunc
        self.atol, self.rtol = atol, rtol


class FlowBlock(nn.Module):
    """continuous/interpflow.py:19-50 (batch_norm=False as PointInterpFlow builds it: the two norm layers are identities)."""

    def __init__(self, idim, cdim, hdims: Tuple[int, ...] = (512, 512, 512), batch_norm=True, layer_type="concatsquash", nonlinearity="tanh"):
        super().__init__()
        if batch_norm:
            raise NotImplementedError("MovingBatchNorm1d is not on the path the checkpoint uses (batch_norm=False)")
        diffeq = ODEnet(hdims, input_shape=(idim,), context_dim=cdim, layer_type=layer_type, nonlinearity=nonlinearity)
        self.cnf = CNF(ODEfunc(diffeq), True, T=0.5, train_T=True, solver="dopri5", atol=1e-5, rtol=1e-5)


class PointInterpFlow(nn.Module):
    """continuous/interpflow.py:53-139."""

    def __init__(self, pc_channel: int):
        super().__init__()
        if pc_channel != 3:
            raise NotImplementedError("the B200 kernels implement the 3-channel (xyz) model the checkpoints hold")
        self.num_blocks = 6
        self.num_neighbors = 16
        self.dist = GaussianDistribution(pc_channel, mu=0.0, vars=1.0, temperature=1.0)
        self.interp = InterpolationModule(pc_channel=3, k=self.num_neighbors)
        feat_channels = [pc_channel, 32, 64] + [128] * (self.num_blocks - 2)
        growth_widths = [8, 16] + [32] * (self.num_blocks - 2)
        cond_channels = [32, 64] + [128] * (self.num_blocks - 2)
        self.feat_convs = nn.ModuleList(
            FeatureExtractUnit(feat_channels[i], feat_channels[i + 1], self.num_neighbors, growth_widths[i], is_dynamic=False)
            for i in range(self.num_blocks))
        self.merge_convs = nn.ModuleList(FeatMergeUnit(feat_channels[i + 1], cond_channels[i]) for i in range(self.num_blocks))
        self.flow_blocks = nn.ModuleList(
            FlowBlock(idim=pc_channel, cdim=cond_channels[i], hdims=(64, 64), batch_norm=False) for i in range(self.num_blocks))
        self._engine_cache = BlobCache()
        self.precision = "fp32"          # tensor-core contractions of the conditioner / interpolation module (see the discrete model)

    def invalidate_packed(self):
        self._engine_cache.invalidate()

    def engine(self, device=None) -> CNFEngine:
        tensors = list(self.parameters()) + list(self.buffers())
        device = device or tensors[0].device
        if torch.device(device).type != "cuda":
            raise RuntimeError(f"PointInterpFlow parameters are on {device}: move the model to a CUDA device "
                               f"(puflow_b200 has no CPU fallback)")
        c = self.flow_blocks[0].cnf
        return self._engine_cache.get(
            tensors, device,
            lambda: CNFEngine({k: v.detach() for k, v in self.state_dict().items()}, device, precision=self.precision,
                              atol=c.atol, rtol=c.rtol),
            extra=(self.precision, c.atol, c.rtol))

    def feat_extract(self, xyz: Tensor, knn_idx: Tensor):
        xyz = check_input(xyz, "PointInterpFlow.feat_extract")
        return self.engine(xyz.device).feat_extract(xyz, _idx32(knn_idx))

    def f(self, xyz: Tensor, cs: List[Tensor]):
        """-> (z, log_det_J = None): the divergence integral is not computed (module docstring)."""
        xyz = check_input(xyz, "PointInterpFlow.f")
        eng = self.engine(xyz.device)
        return eng.f(xyz, eng.contexts(cs)), None

    def g(self, z: Tensor, cs: List[Tensor], upratio: int):
        """z[B,N,3,R] -> [B, N*R, 3] (row n*R + r)."""
        eng = self.engine(z.device)
        zup = torch.flatten(z.transpose(2, 3), 1, 2).contiguous().float()
        return eng.g(zup, eng.contexts(cs), upratio)

    def set_to_initialized_state(self):
        pass

    def forward(self, xyz: Tensor, upratio=4):
        xyz = check_input(xyz, "PointInterpFlow")
        return self.engine(xyz.device).forward(xyz, upratio)

    def log_prob(self, xyz: Tensor, cs: List[Tensor]):
        raise NotImplementedError("log p of the continuous variant (Hutchinson divergence estimate, odefunc.py:8-32) is not built")

    def sample(self, sparse: Tensor, upratio=4):
        dense, _ = self(sparse, upratio)
        return dense
