"""The hot path -- PointInterpFlow.forward (/root/reference/modules/discrete/interpflow.py:327-337) -- as a
fixed sequence of puflow_b200 kernel launches on the current CUDA stream.

    (1) knn_topk K=16                                     interpflow.py:328
    (2) 6 x [pf_linear (P|Q projection) -> pf_edgeconv -> pf_linear x2 (FeatMergeUnit)
             -> pf_linear (conditioning first layers)]    interpflow.py:292-300, 22-43
        pf_flow_cond                                       coupling2 nets, once per input point
    (3) pf_flow_fwd                                        interpflow.py:302-313, 339-345
    (4) pf_linear (K=3) -> pf_interp                       interpflow.py:173-186
    (5) pf_flow_inv on R x N points                        interpflow.py:315-321

`PUFlowEngine` owns the packed weights of one checkpoint on one device.  modules/discrete/interpflow.py
(the drop-in PointInterpFlow) delegates to it.
"""
import math
from typing import Dict, List, Optional

import torch

from . import ops
from ._native import ACT_NONE, ACT_RELU
from .packing import FEAT_CFG, NUM_BLOCKS, pack_state_dict

K_MAIN, K_INTERP = 16, 8          # interpflow.py:270 (num_neighbors), :167 (hard-coded k = 8)
LOG_2PI = math.log(2.0 * math.pi)


class PUFlowEngine:
    MAX_GRAPHS = 4

    def __init__(self, state_dict: Dict[str, torch.Tensor], device, precision="fp32", kernels="tc"):
        """precision: "fp32" (default; fp16 hi/lo split operands, three tensor-core products per contraction: fp32-class
        results, the path all parity numbers refer to) or "fp16" (single-pass fp16 operands, fp32 accumulation:
        PF_PREC_FP16 of include/puflow_b200.h; tolerance stated there and in tests/test_gpu_precision.py)."""
        from ._native import PRECISIONS
        if precision not in PRECISIONS:
            raise ValueError(f"precision {precision!r} not in {sorted(k for k in PRECISIONS if isinstance(k, str))}")
        self.prec = PRECISIONS[precision]
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError(f"PUFlowEngine needs a CUDA device, got {self.device} (no CPU fallback)")
        self.w = pack_state_dict(state_dict, self.device)
        # kernels="tc": the tcgen05 kernels (the product path); kernels="ffma": the FP32 FFMA kernels of round 1 everywhere
        # (edgeconv.cu, flow.cu, interp.cu, linear.cu), kept as the A/B reference the tests compare with.  A constructor
        # argument, not an environment switch: the two values are the only two paths, both tested against the fixtures.
        if kernels not in ("tc", "ffma"):
            raise ValueError(f"kernels {kernels!r} not in ('tc', 'ffma')")
        self.use_tc = kernels == "tc"
        self.flow_tc = self.use_tc
        self._graphs = {}                                # (shape, upratio) -> captured CUDA graph, see forward_graphed
        self._model_blob = None                          # packed once on first use of forward_c

    def _linear(self, a, wd, name, bias=None, act=ACT_NONE):
        """Per-point linear layer: tensor-core kernel when a packed blob exists (K in {16..128}), FFMA otherwise."""
        if self.use_tc and (name + "_tc") in wd:
            return ops.linear_tc(a, wd[name + "_tc"], wd[name].shape[1], bias, act, precision=self.prec)
        return ops.linear(a, wd[name], bias, act)

    # ---- stages (each usable on its own; the drop-in module methods call these) -------------------
    def knn(self, xyz: torch.Tensor, K: int = K_MAIN):
        return ops.knn_topk(xyz, xyz, K, return_dist=False)[1]

    def edgeconv_block(self, i: int, c: torch.Tensor, idx: torch.Tensor, Bp: int, N: int, split_in=None, split_out=None):
        """FeatureExtractUnit i on c[Bp*N, C_in] -> [Bp*N, C_out].  split_in: the same input pre-split in the operand
        layout (written by the previous block, ops.split_tiles) -- the P|Q projection then runs TMA-fed with no conversion
        pass; split_out: buffer that receives this block's output pre-split for the next block."""
        f = self.w["feat"][i]
        Cin, g, Co = FEAT_CFG[i]
        tc = self.use_tc and idx.shape[-1] in (8, 16, 32)
        if split_in is not None and "wpq_tc" in f:
            pq = ops.linear_tc_split(split_in, Bp * N, Cin, f["wpq_tc"], f["wpq"].shape[1], f["bpq"], precision=self.prec)
        else:
            pq = self._linear(c, f, "wpq", f["bpq"])
        if tc:
            return ops.edgeconv_tc(pq, idx, f["wtc"], Bp, N, g, Co, 0.05, split=split_out, precision=self.prec)
        assert split_out is None
        return ops.edgeconv(pq, idx, f["wd"], f["wo"], Bp, N, g, Co, 0.05)

    def merge(self, i: int, c: torch.Tensor):
        """FeatMergeUnit i (interpflow.py:251-258)."""
        m = self.w["merge"][i]
        h = self._linear(c, m, "w1", m["b1"], ACT_RELU)
        return self._linear(h, m, "w2")

    def feat_extract(self, xyz: torch.Tensor, idx: torch.Tensor, return_feats: bool = False):
        """PointInterpFlow.feat_extract (interpflow.py:292-300) -> list of cs[i] [Bp, N, cond_i]."""
        Bp, N, _ = xyz.shape
        c = xyz.reshape(Bp * N, 3)
        cs, feats = [], []
        for i in range(NUM_BLOCKS):
            c = self.edgeconv_block(i, c, idx, Bp, N)
            feats.append(c.view(Bp, N, -1))
            cs.append(self.merge(i, c).view(Bp, N, -1))
        return (cs, feats) if return_feats else cs

    def feat_conditioning(self, xyz: torch.Tensor, idx: torch.Tensor, want_cs: bool = False):
        """feat_extract + conditioning with the merge unit, the conditioning first layers and the coupling2 nets of each
        block fused into one kernel (pf_cond_chain_tc): cs / the 192-wide raw never reach HBM.
        -> ((raw1[6, M, 64], sb[6, M, 8]), feats, cs-or-None)."""
        Bp, N, _ = xyz.shape
        M = Bp * N
        raw1 = torch.empty((NUM_BLOCKS, M, 64), dtype=torch.float32, device=self.device)
        sb = torch.empty((NUM_BLOCKS, M, 8), dtype=torch.float32, device=self.device)
        c = xyz.reshape(M, 3)
        feats, cs = [], ([] if want_cs else None)
        presplit = idx.shape[-1] in (8, 16, 32)
        split = None
        for i in range(NUM_BLOCKS):
            nxt = ops.split_tiles(M, FEAT_CFG[i][2], self.device) if presplit and i + 1 < NUM_BLOCKS else None
            c = self.edgeconv_block(i, c, idx, Bp, N, split_in=split, split_out=nxt)
            split = nxt
            feats.append(c.view(Bp, N, -1))
            ops.cond_chain(c, self.w["chain_tc"][i], raw1[i], sb[i], precision=self.prec)
            if want_cs:
                cs.append(self.merge(i, c).view(Bp, N, -1))
        return (raw1, sb), feats, cs

    def conditioning(self, cs: List[torch.Tensor]):
        """Everything the flow needs from cs, per INPUT point: raw[6, M, 192], sb[6, M, 8]."""
        M = cs[0].shape[0] * cs[0].shape[1]
        raw = torch.empty((NUM_BLOCKS, M, 192), dtype=torch.float32, device=self.device)
        for i in range(NUM_BLOCKS):
            if self.use_tc and self.w["wc_tc"][i] is not None:
                ops.linear_tc(cs[i].reshape(M, -1), self.w["wc_tc"][i], 192, None, ACT_NONE, out=raw[i], precision=self.prec)
            else:
                ops.linear(cs[i].reshape(M, -1), self.w["wc"][i], None, ACT_NONE, out=raw[i])
        sb = ops.flow_cond(raw, self.w["cond_tc"] if self.flow_tc else self.w["cond"], precision=self.prec)
        return raw, sb

    def f(self, xyz: torch.Tensor, cond):
        """PointInterpFlow.f (interpflow.py:302-313) -> (z[Bp,N,3], log_det_J[Bp])."""
        Bp, N, _ = xyz.shape
        raw, sb = cond
        z, ssum = ops.flow_fwd(xyz.reshape(Bp * N, 3), raw, sb, self.w["flow_tc"] if self.flow_tc else self.w["flow"],
                               precision=self.prec)
        logdet = self.w["logdet_const"] * N - ssum.view(Bp, N).sum(1)
        return z.view(Bp, N, 3), logdet

    def log_prob(self, xyz: torch.Tensor, cond):
        """PointInterpFlow.log_prob (interpflow.py:339-345; standard_logp, modules/utils/probs.py:87-93)."""
        z, logdet = self.f(xyz, cond)
        logp = (-0.5 * (z * z + LOG_2PI)).flatten(1).sum(1)
        return z, -torch.mean(logp + logdet)

    def interp(self, z: torch.Tensor, xyz: torch.Tensor, upratio: int, idx: Optional[torch.Tensor] = None):
        """InterpolationModule.forward (interpflow.py:173-186) -> [Bp, N*R, 3], row n*R + r.
        `idx`: any [Bp,N,>=8] sorted kNN list over xyz (the K=16 list is re-used: same metric)."""
        Bp, N, _ = xyz.shape
        if idx is None:
            idx = self.knn(xyz, K_INTERP)
        wi = self.w["interp"]
        if self.use_tc:
            return ops.interp_tc(xyz, z, idx, wi["tc"], upratio, precision=self.prec)
        pq = ops.linear(xyz.reshape(Bp * N, 3), wi["wpq"], wi["bpq"])
        return ops.interp(xyz, z, idx, pq, wi["blob"], upratio)

    def g(self, zup: torch.Tensor, cond, upratio: int):
        """PointInterpFlow.g (interpflow.py:315-321) on zup[Bp, N*R, 3] (row n*R + r)."""
        raw, sb = cond
        Bp = zup.shape[0]
        x = ops.flow_inv(zup.reshape(-1, 3), raw, sb, self.w["flow_tc"] if self.flow_tc else self.w["flow"], upratio,
                         precision=self.prec)
        return x.view(Bp, -1, 3)

    # ---- whole path ----------------------------------------------------------------------------------
    @torch.no_grad()
    def forward(self, xyz: torch.Tensor, upratio: int = 4, return_all: bool = False):
        """xyz[Bp,N,3] fp32 CUDA -> (x[Bp, N*R, 3], logp_x scalar)."""
        if xyz.dim() != 3 or xyz.shape[2] != 3:
            raise ValueError(f"expected xyz[B,N,3], got {tuple(xyz.shape)}")
        xyz = xyz.contiguous().float()
        idx = self.knn(xyz, K_MAIN)
        if self.flow_tc:
            cond, feats, cs = self.feat_conditioning(xyz, idx, want_cs=return_all)
        else:
            cs, feats = self.feat_extract(xyz, idx, return_feats=True)
            cond = self.conditioning(cs)
        z, logp_x = self.log_prob(xyz, cond)
        zup = self.interp(z, xyz, upratio, idx)
        x = self.g(zup, cond, upratio)
        if return_all:
            return dict(knn_idx=idx, feats=feats, cs=cs, z=z, logp_x=logp_x, zup=zup, x=x)
        return x, logp_x

    @torch.no_grad()
    def forward_c(self, xyz: torch.Tensor, upratio: int = 4):
        """forward() through ONE C-ABI call (pf_upsample_patches, csrc/model.cu): same kernels, same results bit for bit
        in x; logp_x reduced in float64 on the device.  What a non-Python host would call."""
        if xyz.dim() != 3 or xyz.shape[2] != 3:
            raise ValueError(f"expected xyz[B,N,3], got {tuple(xyz.shape)}")
        return ops.upsample_patches(self.model_blob(), xyz.contiguous().float(), upratio, precision=self.prec)

    def sample(self, xyz: torch.Tensor, upratio: int = 4):
        return self.forward(xyz, upratio)[0]

    def model_blob(self):
        if self._model_blob is None:
            from .packing import pack_model_blob
            self._model_blob = pack_model_blob(self.w).to(self.device)
        return self._model_blob

    @torch.no_grad()
    def upsample_clouds(self, clouds: torch.Tensor, upratio: int, npoint: int, num_outliers: int, patch_points: int,
                        n_patch: int, out: torch.Tensor = None, graph: bool = False):
        """PatchHelper.upsample(jitter=False) [+ remove_outliers] for clouds[B,N,3] in ONE C call (pf_upsample_clouds):
        normalise -> seeds -> patches -> network -> merge FPS -> (Chamfer + outlier removal), every launch on the current
        stream.  graph=True replays the launch sequence from a CUDA graph captured per shape (static input / output /
        workspace buffers owned by the engine; the result is copied out)."""
        clouds = clouds.contiguous().float()
        if not graph:
            return ops.upsample_clouds(self.model_blob(), clouds, upratio, npoint, num_outliers, patch_points, n_patch, out=out,
                                       precision=self.prec)
        key = ("clouds", tuple(clouds.shape), int(upratio), int(npoint), int(num_outliers), int(patch_points), int(n_patch))
        ent = self._graphs.pop(key, None)
        if ent is None:
            B, N, _ = clouds.shape
            static_in = clouds.clone()
            static_out = torch.empty((B, npoint - num_outliers, 3), dtype=torch.float32, device=self.device)
            ws = torch.empty((ops.upsample_clouds_workspace_bytes(B, N, upratio, npoint, patch_points, n_patch),),
                             dtype=torch.uint8, device=self.device)
            blob = self.model_blob()
            run = lambda: ops.upsample_clouds(blob, static_in, upratio, npoint, num_outliers, patch_points, n_patch,
                                              out=static_out, workspace=ws, precision=self.prec)
            cur = torch.cuda.current_stream(self.device)
            side = torch.cuda.Stream(self.device)
            side.wait_stream(cur)
            prev = ops.set_launch_log(None)
            try:
                with torch.cuda.stream(side):
                    run()                                # lazy initialisation (smem opt-ins) before capture
                cur.wait_stream(side)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    run()
            finally:
                ops.set_launch_log(prev)
            ent = (g, static_in, static_out, ws, ops.cloud_launches(B * n_patch, num_outliers))
            while len(self._graphs) >= self.MAX_GRAPHS:
                self._graphs.pop(next(iter(self._graphs)))
        self._graphs[key] = ent
        g, static_in, static_out, _, n_launch = ent
        static_in.copy_(clouds)
        g.replay()
        ops.count_launches(n_launch)
        if out is not None:
            out.copy_(static_out)
            return out
        return static_out.clone()

    # ---- CUDA-graph replay of the whole path ---------------------------------------
    @torch.no_grad()
    def forward_graphed(self, xyz: torch.Tensor, upratio: int = 4):
        """Same result as forward(), bit for bit, with the ~30 kernel launches + tensor plumbing of one call replayed
        from a CUDA graph captured per (shape, upratio): for small batches (one 2,048-pt cloud = 32 patches) the
        launch sequence, not the kernels, bounds the latency.  Inputs are copied into the graph's static buffer and
        the outputs cloned out, so callers see ordinary tensors.  At most MAX_GRAPHS shapes are kept (LRU)."""
        xyz = xyz.contiguous().float()
        key = (tuple(xyz.shape), int(upratio))
        ent = self._graphs.pop(key, None)
        if ent is None:
            static_in = xyz.clone()
            cur = torch.cuda.current_stream(self.device)
            side = torch.cuda.Stream(self.device)
            side.wait_stream(cur)
            prev = ops.set_launch_log(None)              # no event records inside warm-up / capture
            try:
                with torch.cuda.stream(side):
                    for _ in range(2):                   # lazy initialisation (smem opt-ins, allocator) before capture
                        self.forward(static_in, upratio)
                cur.wait_stream(side)
                counter = ops.LaunchLog(events=False)
                ops.set_launch_log(counter)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    x, logp = self.forward(static_in, upratio)
            finally:
                ops.set_launch_log(prev)
            ent = (graph, static_in, x, logp, counter.count)
            while len(self._graphs) >= self.MAX_GRAPHS:
                self._graphs.pop(next(iter(self._graphs)))
        self._graphs[key] = ent                          # most recently used last
        graph, static_in, x, logp, n_launch = ent
        static_in.copy_(xyz)
        graph.replay()
        ops.count_launches(n_launch)
        return x.clone(), logp.clone()


class CNFEngine(PUFlowEngine):
    """The continuous variant, inference: PointInterpFlow of /root/reference/modules/continuous/interpflow.py:53-139.  Same
    conditioner (kNN, six EdgeConv blocks, merge units) and interpolation module as the discrete model -- the same kernels --
    and six CNF blocks integrated by pf_cnf_flow (csrc/cnf.cu).  log p is not computed (the reference estimates it with a
    random Hutchinson probe, odefunc.py:8-32; sample() never reads it)."""

    def __init__(self, state_dict: Dict[str, torch.Tensor], device, precision="fp32", atol: float = 1e-5, rtol: float = 1e-5):
        from ._native import PRECISIONS
        from .packing import pack_cnf_state_dict
        if precision not in PRECISIONS:
            raise ValueError(f"precision {precision!r} not in {sorted(k for k in PRECISIONS if isinstance(k, str))}")
        self.prec = PRECISIONS[precision]
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError(f"CNFEngine needs a CUDA device, got {self.device} (no CPU fallback)")
        self.w = pack_cnf_state_dict(state_dict, self.device)
        self.use_tc, self.flow_tc = True, False
        self.atol, self.rtol = float(atol), float(rtol)           # cnf.py:37: solver='dopri5', atol = rtol = 1e-5
        self._graphs, self._model_blob = {}, None

    def contexts(self, cs: List[torch.Tensor]):
        """The c part of every CNF layer's gate and bias, per INPUT point and block: [M, 272] each."""
        out = []
        for i in range(NUM_BLOCKS):
            b = self.w["cnf"][i]
            out.append(self._linear(cs[i].reshape(-1, cs[i].shape[-1]), b, "wctx", b["bctx"]))
        return out

    def f(self, xyz: torch.Tensor, ctxs):
        """PointInterpFlow.f (continuous/interpflow.py:96-106): blocks 0..5, each integrated over [0, T_i]."""
        y = xyz.reshape(-1, 3)
        for i in range(NUM_BLOCKS):
            b = self.w["cnf"][i]
            y = ops.cnf_flow(y, ctxs[i], b["blob"], 0.0, b["T"], 1, self.atol, self.rtol)
        return y.view(xyz.shape)

    def g(self, zup: torch.Tensor, ctxs, upratio: int):
        """PointInterpFlow.g (continuous/interpflow.py:108-115) on zup[Bp, N*R, 3] (row n*R + r): blocks 5..0 over [T_i, 0]."""
        y = zup.reshape(-1, 3)
        for i in reversed(range(NUM_BLOCKS)):
            b = self.w["cnf"][i]
            y = ops.cnf_flow(y, ctxs[i], b["blob"], b["T"], 0.0, upratio, self.atol, self.rtol)
        return y.view(zup.shape)

    @torch.no_grad()
    def forward(self, xyz: torch.Tensor, upratio: int = 4, return_all: bool = False):
        """xyz[Bp,N,3] fp32 CUDA -> (x[Bp, N*R, 3], None)  (continuous/interpflow.py:121-130; log p not computed)."""
        if xyz.dim() != 3 or xyz.shape[2] != 3:
            raise ValueError(f"expected xyz[B,N,3], got {tuple(xyz.shape)}")
        xyz = xyz.contiguous().float()
        idx = self.knn(xyz, K_MAIN)
        cs = self.feat_extract(xyz, idx)
        ctxs = self.contexts(cs)
        z = self.f(xyz, ctxs)
        zup = self.interp(z, xyz, upratio, idx)
        x = self.g(zup, ctxs, upratio)
        if return_all:
            return dict(knn_idx=idx, cs=cs, z=z, zup=zup, x=x)
        return x, None

    def forward_c(self, xyz, upratio: int = 4):
        raise NotImplementedError("the one-call C entry points cover the discrete model")

    def upsample_clouds(self, *a, **k):
        raise NotImplementedError("the one-call C entry points cover the discrete model; use modules.utils.patch.PatchHelper")

    def forward_graphed(self, xyz, upratio: int = 4):
        return self.forward(xyz, upratio)
