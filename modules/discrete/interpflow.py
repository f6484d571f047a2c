"""PointInterpFlow and its sub-modules -- same class names, constructor signatures, attribute names and
state_dict keys (408 for PointInterpFlow(3)) as the reference's modules/discrete/interpflow.py:22-350, so
`load_state_dict(torch.load('pretrain/puflow-x4-*.pt'))` works unchanged.  The modules only HOLD the
parameters; every forward runs hand-written sm_100a kernels through puflow_b200 (C-ABI), never ATen
convolutions:

  PointInterpFlow.forward / sample      -> puflow_b200.engine.PUFlowEngine.forward (whole fused path)
  FeatureExtractUnit.forward            -> pf_linear (P|Q projections) + pf_edgeconv
  FeatMergeUnit.forward, LinearA1D      -> pf_linear
  FlowBlock.forward / inverse           -> pf_linear + pf_flow_cond + pf_flow_fwd / pf_flow_inv
  InterpolationModule.forward           -> pf_knn_topk + pf_linear + pf_interp

Inputs must be CUDA tensors; there is no CPU fallback.  On the hot path DistanceEncoder, KnnContextEncoder and
WeightEstimationUnit are evaluated fused inside pf_interp_tc; called on their own (as the reference's API allows:
net.interp.knn_context(xyz), net.interp.weight_unit(ctx)) they run layer by layer on pf_linear.
"""
from typing import List

import torch
import torch.nn as nn
import torch.nn.functional as F  # noqa: F401  (kept for API familiarity)
from torch import Tensor

from puflow_b200 import ops, packing
from puflow_b200._native import ACT_LRELU, ACT_NONE, ACT_RELU
from puflow_b200.engine import K_INTERP, PUFlowEngine

from modules.utils.probs import GaussianDistribution
from modules.flows.coupling import AffineCouplingLayer, AffineInjectorLayer, AffineSpatialCouplingLayer
from modules.flows.normalize import ActNorm
from modules.flows.permutate import Permutation
from modules.flows._fused import BlobCache, check_input, to_dev


def knn_points(p1: Tensor, p2: Tensor, K: int, return_nn: bool = False, return_sorted: bool = True, **kw):
    """pytorch3d.ops.knn_points-compatible wrapper over pf_knn_topk: (dists[B,P1,K], idx int64[B,P1,K], None)."""
    d, i = ops.knn_topk(p1.contiguous().float(), p2.contiguous().float(), K)
    return d, i.long(), None


def _conv1x1(rows: Tensor, conv: nn.Conv2d, bn: nn.BatchNorm2d = None, act: int = ACT_NONE, slope: float = 0.0):
    """One 1x1 Conv2d (+ eval-mode BatchNorm folded in, + activation) on per-edge rows[M, C_in] -> [M, C_out] through
    pf_linear (the FP32 FFMA kernel): the stand-alone forms of the interpolation sub-modules, whose arithmetic the hot
    path evaluates fused inside pf_interp_tc.  C_out must be a multiple of 8."""
    W = conv.weight.detach().double().flatten(1)                       # [C_out, C_in]
    b = conv.bias.detach().double() if conv.bias is not None else torch.zeros(W.shape[0], dtype=torch.float64, device=W.device)
    if bn is not None:
        g = bn.weight.detach().double() / torch.sqrt(bn.running_var.detach().double() + bn.eps)
        W, b = W * g[:, None], (b - bn.running_mean.detach().double()) * g + bn.bias.detach().double()
    return ops.linear(rows, W.t().contiguous().float(), b.float().contiguous(), act, slope)


def _rows(f: Tensor):
    """[B, C, N, K] -> ([B*N*K, C] contiguous fp32, (B, N, K))."""
    B, C, N, K = f.shape
    return f.permute(0, 2, 3, 1).reshape(B * N * K, C).contiguous().float(), (B, N, K)


def _unrows(r: Tensor, shp):
    B, N, K = shp
    return r.view(B, N, K, -1).permute(0, 3, 1, 2)                      # [B, C, N, K]


def _idx32(idx: Tensor):
    return idx if idx.dtype == torch.int32 and idx.is_contiguous() else idx.to(torch.int32).contiguous()


def _cuda_state(module: nn.Module, device):
    return {k: v.detach() for k, v in module.state_dict().items()}, device


# -----------------------------------------------------------------------------------------
class LinearA1D(nn.Module):
    """Linear(no bias) -> LeakyReLU -> Linear -> LeakyReLU -> Linear (zero-init) -- reference interpflow.py:22-43."""

    def __init__(self, dim_in: int, dim_h: int, dim_out: int, dim_c=None):
        super(LinearA1D, self).__init__()
        linear_zero = nn.Linear(dim_h, dim_out, bias=True)
        linear_zero.weight.data.zero_()
        linear_zero.bias.data.zero_()
        in_channel = dim_in if dim_c is None else dim_in + dim_c
        self.layers = nn.Sequential(
            nn.Linear(in_channel, dim_h, bias=False),
            nn.LeakyReLU(inplace=True),
            nn.Linear(dim_h, dim_h, bias=True),
            nn.LeakyReLU(inplace=True),
            linear_zero)
        self._cache = BlobCache()

    def forward(self, h: Tensor, c: Tensor = None):
        if c is not None:
            h = torch.cat([h, c], dim=-1)
        if not h.is_cuda:
            raise RuntimeError("LinearA1D: puflow_b200 has no CPU fallback")
        L = self.layers
        dev = h.device

        def build():
            w4 = torch.zeros(L[4].weight.shape[1], 4 * ((L[4].weight.shape[0] + 3) // 4), dtype=torch.float64)
            w4[:, :L[4].weight.shape[0]] = L[4].weight.detach().double().t().cpu()
            b4 = torch.zeros(w4.shape[1], dtype=torch.float64)
            b4[:L[4].bias.shape[0]] = L[4].bias.detach().double().cpu()
            return (to_dev(L[0].weight.detach().t(), dev), to_dev(L[2].weight.detach().t(), dev),
                    to_dev(L[2].bias.detach(), dev), to_dev(w4, dev), to_dev(b4, dev))
        w0, w2, b2, w4, b4 = self._cache.get((L[0].weight, L[2].weight, L[2].bias, L[4].weight, L[4].bias), dev, build)
        lead = h.shape[:-1]
        a = h.reshape(-1, h.shape[-1]).contiguous().float()
        a = ops.linear(a, w0, None, ACT_LRELU, 0.01)
        a = ops.linear(a, w2, b2, ACT_LRELU, 0.01)
        a = ops.linear(a, w4, b4, ACT_NONE)
        return a[:, :L[4].weight.shape[0]].reshape(*lead, -1)


# -----------------------------------------------------------------------------------------
class FlowBlock(nn.Module):
    """actnorm -> inv1x1 -> additive spatial coupling -> reverse -> affine injector (reference interpflow.py:46-82),
    one fused kernel per direction."""

    def __init__(self, idim, hdim, cdim, is_even):
        super(FlowBlock, self).__init__()
        self.actnorm = ActNorm(idim, dim=2)
        self.permutate1 = Permutation('inv1x1', idim, dim=2)
        self.permutate2 = Permutation('reverse', idim, dim=2)
        if idim == 3:
            tdim = 1 if is_even else 2
            self.coupling1 = AffineSpatialCouplingLayer('additive', LinearA1D, split_dim=2, is_even=is_even, clamp=None,
                params={'dim_in': tdim, 'dim_h': hdim, 'dim_out': idim - tdim, 'dim_c': cdim})
        else:
            self.coupling1 = AffineCouplingLayer('additive', LinearA1D, split_dim=2, clamp=None,
                params={'dim_in': idim // 2, 'dim_h': hdim, 'dim_out': idim - idim // 2, 'dim_c': cdim})
        self.coupling2 = AffineInjectorLayer('affine', LinearA1D, split_dim=2, clamp=None,
            params={'dim_in': cdim, 'dim_h': hdim, 'dim_out': idim, 'dim_c': None})
        self._is_even = is_even
        self._cache = BlobCache()

    def _packed(self, device):
        params = [p for p in self.parameters()]

        def build():
            if not self.actnorm.is_inited:
                raise RuntimeError("ActNorm is not initialised: load a checkpoint and call set_to_initialized_state()")
            sd = {k: v.detach() for k, v in self.state_dict().items()}
            fb = packing.pack_flow_block(sd, "", tdim=1 if self._is_even else 2)
            return dict(wc=to_dev(fb["wc"], device), flow=to_dev(fb["flow"], device).view(1, -1),
                        cond=to_dev(fb["cond"], device).view(1, -1), logdet_const=fb["logdet_const"])
        return self._cache.get(params, device, build)

    def _cond(self, w, c: Tensor, M: int):
        raw = ops.linear(c.reshape(M, -1).contiguous().float(), w["wc"]).view(1, M, 192)
        return raw, ops.flow_cond(raw, w["cond"])

    def forward(self, x: Tensor, c: Tensor = None):
        x = check_input(x, "FlowBlock")
        B, N, _ = x.shape
        w = self._packed(x.device)
        raw, sb = self._cond(w, c, B * N)
        z, ssum = ops.flow_fwd(x.reshape(B * N, 3), raw, sb, w["flow"])
        return z.view(B, N, 3), w["logdet_const"] * N - ssum.view(B, N).sum(1)

    def inverse(self, z: Tensor, c: Tensor = None):
        z = check_input(z, "FlowBlock")
        B, N, _ = z.shape
        w = self._packed(z.device)
        raw, sb = self._cond(w, c, B * N)
        return ops.flow_inv(z.reshape(B * N, 3), raw, sb, w["flow"], 1).view(B, N, 3)


# -----------------------------------------------------------------------------------------
class DistanceEncoder(nn.Module):
    """Parameter container (reference interpflow.py:85-120); evaluated inside pf_interp."""

    def __init__(self, dim_in: int, dim_out: int, k: int):
        super(DistanceEncoder, self).__init__()
        self.k = k
        self.mlp = nn.Sequential(
            nn.Conv2d(dim_in * 3 + 1, 64, [1, 1]),
            nn.BatchNorm2d(64),
            nn.LeakyReLU(inplace=True),
            nn.Conv2d(64, 64, [1, 1]),
            nn.BatchNorm2d(64),
            nn.LeakyReLU(inplace=True),
            nn.Conv2d(64, dim_out, [1, 1]))

    def distance_vec(self, xyz: Tensor):
        """reference interpflow.py:101-115: kNN (pf_knn_topk) + the 10-channel edge feature [x_i, x_j, x_i - x_j, |.|]."""
        xyz = check_input(xyz, "DistanceEncoder")
        B, N, C = xyz.shape
        idx = ops.knn_topk(xyz, xyz, self.k, return_dist=False)[1]                       # [B, N, k] int32, sorted
        neighbors = ops.gather_points(xyz, idx.view(B, N * self.k).contiguous()).view(B, N, self.k, C)
        pt_raw = xyz.unsqueeze(2).expand_as(neighbors)
        vec = pt_raw - neighbors
        dist = torch.sqrt(torch.sum(vec ** 2, dim=-1, keepdim=True))
        f = torch.cat([pt_raw, neighbors, vec, dist], dim=-1).permute(0, 3, 1, 2)        # [B, 10, N, k]
        return f, idx.long().view(B, -1)

    def forward(self, xyz: Tensor):
        """Stand-alone form (reference interpflow.py:117-120) -> (f[B, C_out, N, k], idx_knn[B, N*k]); three pf_linear
        launches on the per-edge rows.  The hot path evaluates the same layers fused inside pf_interp_tc."""
        f_dist, idx_knn = self.distance_vec(xyz)
        r, shp = _rows(f_dist)
        r = _conv1x1(r, self.mlp[0], self.mlp[1], ACT_LRELU, 0.01)
        r = _conv1x1(r, self.mlp[3], self.mlp[4], ACT_LRELU, 0.01)
        r = _conv1x1(r, self.mlp[6])
        return _unrows(r, shp), idx_knn


# -----------------------------------------------------------------------------------------
class KnnContextEncoder(nn.Module):
    """Parameter container (reference interpflow.py:123-134); evaluated inside pf_interp."""

    def __init__(self, pc_channel, k):
        super(KnnContextEncoder, self).__init__()
        self.distance_encoder = DistanceEncoder(dim_in=pc_channel, dim_out=128, k=k)
        self.feat_conv = FeatureExtractUnit(pc_channel, 128, k=k, growth_width=16, is_dynamic=False)

    def forward(self, xyz: Tensor):
        """Stand-alone form (reference interpflow.py:130-134) -> (context[B, 256, N, k], idx_knn[B, N*k])."""
        B, N, _ = xyz.shape
        dist, idx_knn = self.distance_encoder(xyz)
        feat = self.feat_conv(xyz, idx_knn.view(B, N, -1), is_pooling=False)
        return torch.cat([dist, feat], dim=1), idx_knn


# -----------------------------------------------------------------------------------------
class WeightEstimationUnit(nn.Module):
    """Parameter container (reference interpflow.py:137-159); evaluated inside pf_interp."""

    def __init__(self, feat_dim: int):
        super(WeightEstimationUnit, self).__init__()
        self.r_max = 32
        self.mlp = nn.Sequential(
            nn.Conv2d(feat_dim, 128, kernel_size=[1, 1]),
            nn.BatchNorm2d(128),
            nn.LeakyReLU(inplace=True),
            nn.Conv2d(128, 64, kernel_size=[1, 1]),
            nn.BatchNorm2d(64),
            nn.LeakyReLU(inplace=True),
            nn.Conv2d(64, self.r_max, kernel_size=[1, 1]))

    def forward(self, context: Tensor):
        """Stand-alone form (reference interpflow.py:153-159): context[B, C, N, K] -> [B, N, R_max, K]."""
        if not context.is_cuda:
            raise RuntimeError("WeightEstimationUnit: puflow_b200 has no CPU fallback")
        r, shp = _rows(context)
        r = _conv1x1(r, self.mlp[0], self.mlp[1], ACT_LRELU, 0.01)
        r = _conv1x1(r, self.mlp[3], self.mlp[4], ACT_LRELU, 0.01)
        r = _conv1x1(r, self.mlp[6])
        return _unrows(r, shp).permute(0, 2, 1, 3)


# -----------------------------------------------------------------------------------------
class InterpolationModule(nn.Module):
    """Latent kNN-weighted interpolation (reference interpflow.py:162-186): [B,N,3] -> [B,N,3,R]."""

    def __init__(self, pc_channel, k: int):
        super(InterpolationModule, self).__init__()
        k = 8                                   # the reference ignores its argument (interpflow.py:167)
        self.k = k
        self.knn_context = KnnContextEncoder(pc_channel, k)
        self.weight_unit = WeightEstimationUnit(feat_dim=256)
        self._cache = BlobCache()

    def _packed(self, device):
        tensors = list(self.parameters()) + list(self.buffers())

        def build():
            sd = {"interp." + k: v.detach() for k, v in self.state_dict().items()}
            return {k: to_dev(v, device) for k, v in packing.pack_interp(sd).items()}
        return self._cache.get(tensors, device, build)

    def forward(self, z: Tensor, xyz: Tensor, upratio: int, knn_idx: Tensor = None):
        xyz, z = check_input(xyz, "InterpolationModule"), check_input(z, "InterpolationModule")
        B, N, _ = xyz.shape
        w = self._packed(xyz.device)
        idx = _idx32(knn_idx) if knn_idx is not None else ops.knn_topk(xyz, xyz, self.k, return_dist=False)[1]
        pq = ops.linear(xyz.reshape(B * N, 3), w["wpq"], w["bpq"])
        out = ops.interp(xyz, z, idx, pq, w["blob"], upratio)            # [B, N*R, 3], row n*R + r
        return out.view(B, N, upratio, 3).transpose(2, 3)                # [B, N, 3, R] like the reference


# -----------------------------------------------------------------------------------------
class FeatureExtractUnit(nn.Module):
    """EdgeConv dense block (reference interpflow.py:190-248): gather + 4 x (conv1x1 + BN + LeakyReLU(0.05) + concat)
    + conv_out + max over K, as one pf_linear + one pf_edgeconv launch."""

    def __init__(self, idim: int, odim: int, k: int, growth_width: int, is_dynamic: bool):
        super(FeatureExtractUnit, self).__init__()
        assert (odim % growth_width == 0)
        self.k = k
        self.is_dynamic = is_dynamic
        self.num_conv = (odim // growth_width)
        self._cfg = (idim, growth_width, odim)
        idim = idim * 3
        self.convs = nn.ModuleList()
        conv_first = nn.Sequential(*[
            nn.Conv2d(idim, growth_width, kernel_size=[1, 1]),
            nn.BatchNorm2d(growth_width),
            nn.LeakyReLU(0.05, inplace=True),
        ])
        self.convs.append(conv_first)
        for i in range(self.num_conv - 1):
            in_channel = growth_width * (i + 1) + idim
            conv = nn.Sequential(*[
                nn.Conv2d(in_channel, growth_width, kernel_size=[1, 1], bias=True),
                nn.BatchNorm2d(growth_width),
                nn.LeakyReLU(0.05, inplace=True),
            ])
            self.convs.append(conv)
        self.conv_out = nn.Conv2d(growth_width * self.num_conv + idim, odim, kernel_size=[1, 1], bias=True)
        self._cache = BlobCache()

    def _packed(self, device):
        tensors = list(self.parameters()) + list(self.buffers())
        C, g, Co = self._cfg

        def build():
            sd = {"u." + k: v.detach() for k, v in self.state_dict().items()}
            w = packing.pack_edgeconv(sd, "u", C, g, Co, self.num_conv)
            out = {k: to_dev(v, device) for k, v in w.items()}
            if (g, Co, self.num_conv) == (32, 128, 4):
                out["wtc"] = packing.pack_edgeconv_tc(w["wd"], w["wo"]).to(device)
            return out
        return self._cache.get(tensors, device, build)

    def forward(self, x: Tensor, knn_idx: Tensor or None, is_pooling=True):
        C, g, Co = self._cfg
        if not x.is_cuda:
            raise RuntimeError("FeatureExtractUnit: puflow_b200 has no CPU fallback")
        if not is_pooling or self.num_conv != 4 or (g, Co) not in ((8, 32), (16, 64), (32, 128)):
            return self._forward_layers(x, knn_idx, is_pooling)
        B, N, _ = x.shape
        if knn_idx is None:
            if not self.is_dynamic or C != 3:
                raise ValueError("knn_idx is required (the graph is static on the PU-Flow path, interpflow.py:279)")
            knn_idx = ops.knn_topk(x.contiguous().float(), x.contiguous().float(), self.k, return_dist=False)[1]
        w = self._packed(x.device)
        pq = ops.linear(x.reshape(B * N, C).contiguous().float(), w["wpq"], w["bpq"])
        idx32 = _idx32(knn_idx)
        if "wtc" in w and idx32.shape[-1] in (8, 16, 32):
            return ops.edgeconv_tc(pq, idx32, w["wtc"], B, N, g, Co, 0.05).view(B, N, Co)     # tcgen05 path
        return ops.edgeconv(pq, idx32, w["wd"], w["wo"], B, N, g, Co, 0.05).view(B, N, Co)


    def _forward_layers(self, x: Tensor, knn_idx: Tensor, is_pooling: bool):
        """General / un-pooled form (reference interpflow.py:234-248 as written: edge features, num_conv x (conv1x1 + BN +
        LeakyReLU(0.05) + concat), conv_out [, max over K]) layer by layer through pf_linear on the per-edge rows: what
        KnnContextEncoder's stand-alone forward calls (k = 8, growth 16, 8 convs, is_pooling=False -> [B, C, N, K])."""
        if knn_idx is None:
            raise ValueError("knn_idx is required")
        x = x.contiguous().float()
        B, N, C = x.shape
        K = knn_idx.shape[-1]
        nb = ops.gather_points(x, _idx32(knn_idx).view(B, N * K).contiguous()).view(B, N, K, C) if C == 3 else \
            x[torch.arange(B, device=x.device).view(-1, 1, 1), knn_idx.long()]
        xt = x.unsqueeze(2).expand_as(nb)
        f = torch.cat([xt, nb, nb - xt], dim=-1).reshape(B * N * K, 3 * C).contiguous()
        for i in range(self.num_conv):
            f = torch.cat([f, _conv1x1(f, self.convs[i][0], self.convs[i][1], ACT_LRELU, 0.05)], dim=1).contiguous()
        f = _conv1x1(f, self.conv_out).view(B, N, K, -1)
        if is_pooling:
            return f.max(dim=2)[0]                                       # [B, N, C]
        return f.permute(0, 3, 1, 2)                                      # [B, C, N, K]


# -----------------------------------------------------------------------------------------
class FeatMergeUnit(nn.Module):
    """Linear + ReLU + Linear(no bias) (reference interpflow.py:251-258) as two pf_linear launches."""

    def __init__(self, idim: int, odim: int):
        super(FeatMergeUnit, self).__init__()
        self.conv1 = nn.Linear(idim, idim // 2, bias=True)
        self.conv2 = nn.Linear(idim // 2, odim, bias=False)
        self._cache = BlobCache()

    def forward(self, x: Tensor):
        if not x.is_cuda:
            raise RuntimeError("FeatMergeUnit: puflow_b200 has no CPU fallback")
        dev = x.device
        w1, b1, w2 = self._cache.get(
            (self.conv1.weight, self.conv1.bias, self.conv2.weight), dev,
            lambda: (to_dev(self.conv1.weight.detach().t(), dev), to_dev(self.conv1.bias.detach(), dev),
                     to_dev(self.conv2.weight.detach().t(), dev)))
        lead = x.shape[:-1]
        h = ops.linear(x.reshape(-1, x.shape[-1]).contiguous().float(), w1, b1, ACT_RELU)
        return ops.linear(h, w2).view(*lead, -1)


# -----------------------------------------------------------------------------------------
class PointInterpFlow(nn.Module):
    """The discrete PU-Flow model (reference interpflow.py:262-350)."""

    def __init__(self, pc_channel: int):
        super(PointInterpFlow, self).__init__()
        if pc_channel != 3:
            raise NotImplementedError("the B200 kernels implement the 3-channel (xyz) model the checkpoints hold")
        self.num_blocks = 6
        self.num_neighbors = 16

        self.dist = GaussianDistribution(pc_channel, mu=0.0, vars=1.0, temperature=1.0)
        self.interp = InterpolationModule(pc_channel=3, k=self.num_neighbors)

        feat_channels = [pc_channel, 32, 64] + [128] * (self.num_blocks - 2)
        growth_widths = [8, 16] + [32] * (self.num_blocks - 2)
        cond_channels = [32, 64] + [128] * (self.num_blocks - 2)

        self.feat_convs = nn.ModuleList()
        for i in range(self.num_blocks):
            self.feat_convs.append(FeatureExtractUnit(feat_channels[i], feat_channels[i + 1], self.num_neighbors,
                                                      growth_widths[i], is_dynamic=False))
        self.merge_convs = nn.ModuleList()
        for i in range(self.num_blocks):
            self.merge_convs.append(FeatMergeUnit(feat_channels[i + 1], cond_channels[i]))
        self.flow_blocks = nn.ModuleList()
        for i in range(self.num_blocks):
            self.flow_blocks.append(FlowBlock(idim=pc_channel, hdim=64, cdim=cond_channels[i], is_even=(i % 2 == 0)))
        self._engine_cache = BlobCache()
        # Precision of the tensor-core contractions: "fp32" (default: split-fp16 operands, three products per contraction,
        # fp32-class -- the parity path) or "fp16" (single pass; tolerance stated in include/puflow_b200.h).  An attribute,
        # not an environment switch: set `net.precision = "fp16"` before the next call.
        self.precision = "fp32"
        self.kernels = "tc"          # "ffma": the FP32 FFMA reference kernels (A/B comparison)
        self.cuda_graph = False      # forward() replayed from a CUDA graph per input shape

    def invalidate_packed(self):
        """Drop the packed weights (call after editing parameters through `.data`, which the cache cannot see)."""
        self._engine_cache.invalidate()
        for m in self.modules():
            c = getattr(m, "_cache", None)
            if isinstance(c, BlobCache):
                c.invalidate()

    # ---- packed-weight engine, rebuilt only when a parameter / buffer changes or the model moves ----
    def engine(self, device=None) -> PUFlowEngine:
        tensors = list(self.parameters()) + list(self.buffers())
        device = device or tensors[0].device
        if torch.device(device).type != "cuda":
            raise RuntimeError(f"PointInterpFlow parameters are on {device}: move the model to a CUDA device "
                               f"(puflow_b200 has no CPU fallback)")

        def build():
            for fb in self.flow_blocks:
                if not fb.actnorm.is_inited:
                    raise RuntimeError("ActNorm is not initialised: load a checkpoint and call "
                                       "set_to_initialized_state() (reference upsample.py:32-33)")
            return PUFlowEngine({k: v.detach() for k, v in self.state_dict().items()}, device, precision=self.precision,
                                kernels=self.kernels)
        return self._engine_cache.get(tensors, device, build, extra=(self.precision, self.kernels))

    def feat_extract(self, xyz: Tensor, knn_idx: Tensor):
        xyz = check_input(xyz, "PointInterpFlow.feat_extract")
        return self.engine(xyz.device).feat_extract(xyz, _idx32(knn_idx))

    def f(self, xyz: Tensor, cs: List[Tensor]):
        xyz = check_input(xyz, "PointInterpFlow.f")
        eng = self.engine(xyz.device)
        return eng.f(xyz, eng.conditioning(cs))

    def g(self, z: Tensor, cs: List[Tensor], upratio: int):
        """z[B,N,3,R] -> [B, N*R, 3] (row n*R + r)."""
        eng = self.engine(z.device)
        zup = torch.flatten(z.transpose(2, 3), 1, 2).contiguous().float()
        return eng.g(zup, eng.conditioning(cs), upratio)

    def set_to_initialized_state(self):
        for i in range(self.num_blocks):
            self.flow_blocks[i].actnorm.is_inited = True

    def forward(self, xyz: Tensor, upratio=4):
        xyz = check_input(xyz, "PointInterpFlow")
        eng = self.engine(xyz.device)
        # cuda_graph: replay the launch sequence from a CUDA graph captured per input shape (same result; for small
        # batches the launches of a call, not the kernels, bound the latency)
        if self.cuda_graph and not torch.is_grad_enabled():
            return eng.forward_graphed(xyz, upratio)
        # The whole network through the single C-ABI entry point pf_upsample_patches (same kernels, same results bit for
        # bit, no per-kernel Python); the kernel-by-kernel path is kept for per-launch timing from Python (ops.LaunchLog)
        # and for the FFMA reference kernels.
        if eng.flow_tc and not ops.per_launch_events_wanted() and 16 <= xyz.shape[1] <= 65535 and 1 <= upratio <= 32:
            return eng.forward_c(xyz, upratio)
        return eng.forward(xyz, upratio)

    def log_prob(self, xyz: Tensor, cs: List[Tensor]):
        xyz = check_input(xyz, "PointInterpFlow.log_prob")
        eng = self.engine(xyz.device)
        return eng.log_prob(xyz, eng.conditioning(cs))

    def sample(self, sparse: Tensor, upratio=4):
        dense, _ = self(sparse, upratio)
        return dense
# -----------------------------------------------------------------------------------------
