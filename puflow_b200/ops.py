"""torch.Tensor wrappers over the C-ABI (include/puflow_b200.h).

Each wrapper checks device / dtype / contiguity, allocates the outputs with torch (the caller-owns-
memory convention of the reference's chamfer_3D binding,
/root/reference/metric/PyTorchCD/chamfer3D/dist_chamfer_3D.py:40-52), and launches on torch's
current CUDA stream.  CPU tensors are rejected: there is no fallback path.
"""
import torch

from . import _native as N

lib = N.lib


class LaunchLog:
    """Optional launch accounting for bench.py / profiling: counts the kernels this module launches and,
    when `events` is on, brackets each C-ABI call with CUDA events on the launching stream."""

    def __init__(self, events: bool = False):
        self.events = events
        self.count = 0
        self.records = []          # (tag, start_event, end_event)

    def times_ms(self):
        out = {}
        for tag, e0, e1 in self.records:
            out.setdefault(tag, []).append(e0.elapsed_time(e1))
        return out


_log = None


def set_launch_log(log):
    """Install (or remove, with None) a LaunchLog; returns the previous one."""
    global _log
    prev, _log = _log, log
    return prev


_c_profile = False


def c_profile(on: bool):
    """Switch the library's own per-launch event timing of pf_upsample_patches / pf_upsample_clouds on / off
    (pf_profile_enable, per host thread): per-launch times then come from inside the one C call, and nothing has to go
    kernel by kernel in Python."""
    global _c_profile
    _c_profile = bool(on)
    N.check(lib.pf_profile_enable(1 if on else 0), "pf_profile_enable")


def c_profile_read():
    """{tag: [ms, ...]} of the launches of the one-call entry points recorded on THIS host thread since c_profile(True) /
    the last read (waits for them)."""
    import ctypes
    out = {}
    name, ms = ctypes.create_string_buffer(40), ctypes.c_float()
    for i in range(lib.pf_profile_intervals()):
        N.check(lib.pf_profile_interval(i, name, ctypes.byref(ms)), "pf_profile_interval")
        out.setdefault(name.value.decode(), []).append(float(ms.value))
    lib.pf_profile_clear()
    return out


def per_launch_events_wanted() -> bool:
    """True while a LaunchLog with per-launch CUDA events is installed and the library's own timing (c_profile) is off:
    callers then go kernel by kernel through the wrappers below instead of through the one-call C entry point."""
    return _log is not None and _log.events and not _c_profile


def count_launches(n: int):
    """Account for `n` kernels launched outside the wrappers below (a CUDA-graph replay)."""
    if _log is not None:
        _log.count += n


class _launch:
    def __init__(self, tag: str, n: int = 1):
        self.tag, self.n = tag, n

    def __enter__(self):
        if _log is not None:
            _log.count += self.n
            if _log.events:
                self.e0 = torch.cuda.Event(enable_timing=True)
                self.e0.record()

    def __exit__(self, *exc):
        if _log is not None and _log.events and exc[0] is None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            _log.records.append((self.tag, self.e0, e1))
        return False


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _chk(t: torch.Tensor, name: str, dtype=torch.float32):
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name}: expected a torch.Tensor")
    if not t.is_cuda:
        raise N.PuflowError(f"{name}: tensor is on {t.device}; puflow_b200 runs on CUDA only (no CPU fallback)")
    if t.dtype != dtype:
        raise TypeError(f"{name}: expected {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name}: tensor must be contiguous")
    return t


def _dev(t):
    return torch.cuda.device(t.device)


# ---- (1) kNN -------------------------------------------------------------------------------------
def knn_topk(q: torch.Tensor, p: torch.Tensor, K: int, return_dist: bool = True):
    """q[B,nq,3], p[B,np,3] -> (dist[B,nq,K] squared L2 or None, idx[B,nq,K] int32), sorted ascending by
    (distance, index).  pytorch3d.ops.knn_points semantics (interpflow.py:104,228,328)."""
    _chk(q, "q"), _chk(p, "p")
    if q.dim() != 3 or p.dim() != 3 or q.shape[2] != 3 or p.shape[2] != 3 or q.shape[0] != p.shape[0]:
        raise ValueError(f"knn_topk: expected q[B,nq,3], p[B,np,3]; got {tuple(q.shape)}, {tuple(p.shape)}")
    B, nq, _ = q.shape
    np_ = p.shape[1]
    idx = torch.empty((B, nq, K), dtype=torch.int32, device=q.device)
    dist = torch.empty((B, nq, K), dtype=torch.float32, device=q.device) if return_dist else None
    if B == 0 or nq == 0:
        return dist, idx
    with _dev(q):
        for b0 in range(0, B, 65535):
            b1 = min(B, b0 + 65535)
            with _launch(f"knn_topk_k{K}"):
                N.check(lib.pf_knn_topk(q[b0:b1].data_ptr(), p[b0:b1].data_ptr(), b1 - b0, nq, np_, K,
                                        idx[b0:b1].data_ptr(), dist[b0:b1].data_ptr() if return_dist else None,
                                        _stream()), "pf_knn_topk")
    return dist, idx


# ---- (5) Chamfer ---------------------------------------------------------------------------------
def chamfer_nn(a: torch.Tensor, b: torch.Tensor, both: bool = True):
    """a[B,n,3], b[B,m,3] -> (dist1[B,n], dist2[B,m], idx1 int32, idx2 int32); chamfer_3DDist forward
    (metric/PyTorchCD/chamfer3D/dist_chamfer_3D.py:29-81).  both=False computes the a -> b direction only
    (dist2 / idx2 are None): PatchHelper.remove_outliers reads nothing else (patch.py:201-203)."""
    _chk(a, "a"), _chk(b, "b")
    if a.dim() != 3 or b.dim() != 3 or a.shape[2] != 3 or b.shape[2] != 3 or a.shape[0] != b.shape[0]:
        raise ValueError(f"chamfer_nn: expected a[B,n,3], b[B,m,3]; got {tuple(a.shape)}, {tuple(b.shape)}")
    B, n, _ = a.shape
    m = b.shape[1]
    d1 = torch.empty((B, n), dtype=torch.float32, device=a.device)
    d2 = torch.empty((B, m), dtype=torch.float32, device=a.device) if both else None
    i1 = torch.empty((B, n), dtype=torch.int32, device=a.device)
    i2 = torch.empty((B, m), dtype=torch.int32, device=a.device) if both else None
    with _dev(a), _launch("chamfer_nn", 2 if both else 1):
        N.check(lib.pf_chamfer_nn(a.data_ptr(), n, b.data_ptr(), m, B, d1.data_ptr(), i1.data_ptr(),
                                  d2.data_ptr() if both else None, i2.data_ptr() if both else None, _stream()),
                "pf_chamfer_nn")
    return d1, d2, i1, i2


# ---- per-point linear ----------------------------------------------------------------------------
def linear(a: torch.Tensor, w: torch.Tensor, bias=None, act: int = N.ACT_NONE, slope: float = 0.0, out=None):
    """out[M,N] = act(a[M,K] @ w[K,N] + bias).  `a` may be a row-strided 2-D view (stride(1) == 1)."""
    if a.dim() != 2 or a.stride(1) != 1 or not a.is_cuda or a.dtype != torch.float32:
        raise ValueError("linear: a must be a 2-D fp32 CUDA tensor with unit inner stride")
    _chk(w, "w")
    M, K = a.shape
    if w.shape[0] != K:
        raise ValueError(f"linear: a is [M,{K}] but w is {tuple(w.shape)}")
    Nn = w.shape[1]
    if bias is not None:
        _chk(bias, "bias")
    if out is None:
        out = torch.empty((M, Nn), dtype=torch.float32, device=a.device)
    lda = a.stride(0) if M > 1 else max(K, a.stride(0))
    with _dev(a), _launch(f"linear_k{K}_n{Nn}"):
        N.check(lib.pf_linear(a.data_ptr(), lda, w.data_ptr(), bias.data_ptr() if bias is not None else None,
                              out.data_ptr(), out.stride(0), M, K, Nn, act, float(slope), _stream()), "pf_linear")
    return out


def linear_tc(a: torch.Tensor, wtc: torch.Tensor, N_out: int, bias=None, act: int = N.ACT_NONE, slope: float = 0.0,
              out=None, precision: int = 0):
    """Tensor-core (tcgen05, split-fp16) variant of `linear`: a[M,K] fp32, wtc = packing.pack_linear_tc(W[K,N]).
    K in {16,32,64,128}, N % 16 == 0."""
    if a.dim() != 2 or a.stride(1) != 1 or not a.is_cuda or a.dtype != torch.float32:
        raise ValueError("linear_tc: a must be a 2-D fp32 CUDA tensor with unit inner stride")
    _chk(wtc, "wtc", torch.uint8)
    M, K = a.shape
    if wtc.numel() != ((N_out + 127) // 128) * 128 * K * 4:
        raise ValueError(f"linear_tc: packed weight blob does not match K={K}, N={N_out}")
    if bias is not None:
        _chk(bias, "bias")
    if out is None:
        out = torch.empty((M, N_out), dtype=torch.float32, device=a.device)
    lda = a.stride(0) if M > 1 else max(K, a.stride(0))
    with _dev(a), _launch(f"linear_tc_k{K}_n{N_out}"):
        N.check(lib.pf_linear_tc(a.data_ptr(), lda, wtc.data_ptr(), bias.data_ptr() if bias is not None else None,
                                 out.data_ptr(), out.stride(0), M, K, N_out, act, float(slope), int(precision), _stream()),
                "pf_linear_tc")
    return out


# ---- (2) EdgeConv --------------------------------------------------------------------------------
def edgeconv(pq: torch.Tensor, idx: torch.Tensor, wd: torch.Tensor, wo: torch.Tensor, Bp: int, Npts: int, G: int,
             C_out: int, slope: float = 0.05):
    """pq[Bp*N, 2(4G+C_out)], idx[Bp,N,K] int32 -> out[Bp*N, C_out] (FeatureExtractUnit.forward)."""
    _chk(pq, "pq"), _chk(idx, "idx", torch.int32), _chk(wd, "wd"), _chk(wo, "wo")
    K = idx.shape[-1]
    R = 4 * G + C_out
    if pq.shape != (Bp * Npts, 2 * R) or idx.numel() != Bp * Npts * K:
        raise ValueError(f"edgeconv: pq {tuple(pq.shape)} / idx {tuple(idx.shape)} do not match Bp={Bp} N={Npts} R={R}")
    if wd.numel() != 6 * G * G or wo.numel() != 4 * G * C_out:
        raise ValueError("edgeconv: packed weight sizes do not match (G, C_out)")
    out = torch.empty((Bp * Npts, C_out), dtype=torch.float32, device=pq.device)
    with _dev(pq), _launch(f"edgeconv_g{G}_c{C_out}_k{K}"):
        N.check(lib.pf_edgeconv(pq.data_ptr(), idx.data_ptr(), wd.data_ptr(), wo.data_ptr(), out.data_ptr(), Bp, Npts,
                                K, G, C_out, float(slope), _stream()), "pf_edgeconv")
    return out


def split_tiles(M: int, K: int, device):
    """Uninitialised buffer for M rows of K channels in the pre-split operand layout (ceil(M/128) tiles of 512 K bytes)."""
    return torch.empty(((M + 127) // 128, 512 * K), dtype=torch.uint8, device=device)


def linear_tc_split(a_split: torch.Tensor, M: int, K: int, wtc: torch.Tensor, Nout: int, bias=None, act: int = N.ACT_NONE,
                    slope: float = 0.0, precision: int = 0):
    """out[M, Nout] = act(A . W + bias) for A given pre-split (split_tiles layout, written by edgeconv_tc(split=...))."""
    _chk(a_split, "a_split", torch.uint8), _chk(wtc, "wtc", torch.uint8)
    if a_split.numel() != ((M + 127) // 128) * 512 * K or K not in (32, 64, 128) or Nout % 16:
        raise ValueError("linear_tc_split: bad shapes")
    if wtc.numel() != ((Nout + 127) // 128) * 128 * K * 4:
        raise ValueError("linear_tc_split: packed weight blob has the wrong size")
    if bias is not None:
        _chk(bias, "bias")
    out = torch.empty((M, Nout), dtype=torch.float32, device=a_split.device)
    with _dev(a_split), _launch(f"linear_split_k{K}_n{Nout}"):
        N.check(lib.pf_linear_tc_split(a_split.data_ptr(), wtc.data_ptr(), bias.data_ptr() if bias is not None else None,
                                       out.data_ptr(), Nout, M, K, Nout, act, float(slope), int(precision), _stream()),
                "pf_linear_tc_split")
    return out


def edgeconv_tc(pq: torch.Tensor, idx: torch.Tensor, wtc: torch.Tensor, Bp: int, Npts: int, G: int, C_out: int,
                slope: float = 0.05, split: torch.Tensor = None, precision: int = 0):
    """Tensor-core (tcgen05, split-fp16) variant of `edgeconv` for K = 16, (G, C_out) in {(8,32), (16,64), (32,128)};
    wtc: uint8 blob from packing.pack_edgeconv_tc(wd, wo, G, C_out)."""
    _chk(pq, "pq"), _chk(idx, "idx", torch.int32), _chk(wtc, "wtc", torch.uint8)
    K = idx.shape[-1]
    R = 4 * G + C_out
    if pq.shape != (Bp * Npts, 2 * R) or idx.numel() != Bp * Npts * K:
        raise ValueError(f"edgeconv_tc: pq {tuple(pq.shape)} / idx {tuple(idx.shape)} do not match Bp={Bp} N={Npts} R={R}")
    if wtc.numel() != 24 * max(G, 16) ** 2 + 2048 * max(G, 16):
        raise ValueError("edgeconv_tc: packed weight blob has the wrong size")
    out = torch.empty((Bp * Npts, C_out), dtype=torch.float32, device=pq.device)
    if split is not None:       # also write the output pre-split (split_tiles(Bp * Npts, C_out)) for linear_tc_split
        _chk(split, "split", torch.uint8)
        if split.numel() != ((Bp * Npts + 127) // 128) * 512 * C_out:
            raise ValueError("edgeconv_tc: split buffer has the wrong size")
    with _dev(pq), _launch(f"edgeconv_tc_g{G}_c{C_out}_k{K}"):
        N.check(lib.pf_edgeconv_tc_split(pq.data_ptr(), idx.data_ptr(), wtc.data_ptr(), out.data_ptr(),
                                         split.data_ptr() if split is not None else None, Bp, Npts, K, G, C_out,
                                         float(slope), int(precision), _stream()), "pf_edgeconv_tc")
    return out


# ---- (3) flow ------------------------------------------------------------------------------------
def flow_cond(raw: torch.Tensor, cond_w: torch.Tensor, precision: int = 0):
    """raw[nb, M, 192] -> sb[nb, M, 8] = (s0,s1,s2,b0,b1,b2,0,0): coupling2 scale/bias per input point."""
    _chk(raw, "raw")
    nb, M, w = raw.shape
    sb = torch.empty((nb, M, 8), dtype=torch.float32, device=raw.device)
    if cond_w.dtype == torch.uint8:          # tensor-core variant: blob from packing.pack_cond_tc
        _chk(cond_w, "cond_w", torch.uint8)
        if w != 192 or cond_w.shape != (nb, N.PF_COND_TC_BLOCK):
            raise ValueError("flow_cond: bad shapes")
        with _dev(raw), _launch("flow_cond_tc"):
            N.check(lib.pf_flow_cond_tc(raw.data_ptr(), cond_w.data_ptr(), sb.data_ptr(), M, nb, int(precision), _stream()),
                    "pf_flow_cond_tc")
        return sb
    _chk(cond_w, "cond_w")
    if w != 192 or cond_w.shape != (nb, N.PF_COND_BLOB):
        raise ValueError("flow_cond: bad shapes")
    with _dev(raw), _launch("flow_cond"):
        N.check(lib.pf_flow_cond(raw.data_ptr(), cond_w.data_ptr(), sb.data_ptr(), M, nb, _stream()), "pf_flow_cond")
    return sb


def cond_chain(c: torch.Tensor, chain_w: torch.Tensor, raw1: torch.Tensor, sb: torch.Tensor, precision: int = 0):
    """One block of the fused conditioning chain: c[M, C] (EdgeConv output) -> raw1[M, 64], sb[M, 8] written in place
    (FeatMergeUnit interpflow.py:251-258 + first layers of the three LinearA1D nets + the coupling2 nets)."""
    _chk(c, "c"), _chk(chain_w, "chain_w", torch.uint8), _chk(raw1, "raw1"), _chk(sb, "sb")
    M, C = c.shape
    if C not in (32, 64, 128) or chain_w.numel() != N.chain_tc_bytes(C) or raw1.shape != (M, 64) or sb.shape != (M, 8):
        raise ValueError("cond_chain: bad shapes")
    with _dev(c), _launch(f"cond_chain_tc_c{C}"):
        N.check(lib.pf_cond_chain_tc(c.data_ptr(), C, chain_w.data_ptr(), raw1.data_ptr(), sb.data_ptr(), M, int(precision), _stream()),
                "pf_cond_chain_tc")
    return raw1, sb


def flow_fwd(x: torch.Tensor, raw: torch.Tensor, sb: torch.Tensor, flow_w: torch.Tensor, precision: int = 0):
    """x[M,3] -> (z[M,3], ssum[M]); FlowBlock.forward x nb (interpflow.py:66-74,302-313)."""
    tc = flow_w.dtype == torch.uint8         # tensor-core variant: blob from packing.pack_flow_tc
    _chk(x, "x"), _chk(raw, "raw"), _chk(sb, "sb"), _chk(flow_w, "flow_w", torch.uint8 if tc else torch.float32)
    nb, M, ldraw = raw.shape
    if x.shape != (M, 3) or sb.shape != (nb, M, 8) or flow_w.shape != (nb, N.PF_FLOW_TC_BLOCK if tc else N.PF_FLOW_BLOB) \
            or ldraw not in ((64, 192) if tc else (192,)):
        raise ValueError("flow_fwd: bad shapes")
    z = torch.empty_like(x)
    ssum = torch.empty((M,), dtype=torch.float32, device=x.device)
    if tc:
        with _dev(x), _launch("flow_fwd_tc"):
            N.check(lib.pf_flow_fwd_tc(x.data_ptr(), raw.data_ptr(), ldraw, sb.data_ptr(), flow_w.data_ptr(), z.data_ptr(),
                                       ssum.data_ptr(), M, nb, int(precision), _stream()), "pf_flow_fwd_tc")
        return z, ssum
    with _dev(x), _launch("flow_fwd"):
        N.check(lib.pf_flow_fwd(x.data_ptr(), raw.data_ptr(), sb.data_ptr(), flow_w.data_ptr(), z.data_ptr(),
                                ssum.data_ptr(), M, nb, _stream()), "pf_flow_fwd")
    return z, ssum


def flow_inv(zin: torch.Tensor, raw: torch.Tensor, sb: torch.Tensor, flow_w: torch.Tensor, R: int, precision: int = 0):
    """zin[M*R,3] (row n*R+r conditioned by input point n) -> x[M*R,3]; FlowBlock.inverse x nb."""
    tc = flow_w.dtype == torch.uint8
    _chk(zin, "zin"), _chk(raw, "raw"), _chk(sb, "sb"), _chk(flow_w, "flow_w", torch.uint8 if tc else torch.float32)
    nb, M, ldraw = raw.shape
    if zin.shape != (M * R, 3) or sb.shape != (nb, M, 8) or ldraw not in ((64, 192) if tc else (192,)) or \
            flow_w.shape != (nb, N.PF_FLOW_TC_BLOCK if tc else N.PF_FLOW_BLOB):
        raise ValueError("flow_inv: bad shapes")
    x = torch.empty_like(zin)
    if tc:
        with _dev(zin), _launch("flow_inv_tc"):
            N.check(lib.pf_flow_inv_tc(zin.data_ptr(), raw.data_ptr(), ldraw, sb.data_ptr(), flow_w.data_ptr(), x.data_ptr(),
                                       M, R, nb, int(precision), _stream()), "pf_flow_inv_tc")
        return x
    with _dev(zin), _launch("flow_inv"):
        N.check(lib.pf_flow_inv(zin.data_ptr(), raw.data_ptr(), sb.data_ptr(), flow_w.data_ptr(), x.data_ptr(), M, R,
                                nb, _stream()), "pf_flow_inv")
    return x


# ---- (4) interpolation ---------------------------------------------------------------------------
def interp(xyz: torch.Tensor, z: torch.Tensor, idx: torch.Tensor, pq: torch.Tensor, w: torch.Tensor, R: int):
    """xyz[Bp,N,3], z[Bp,N,3], idx[Bp,N,>=8] int32, pq[Bp*N,640] -> out[Bp, N*R, 3] (row n*R + r)."""
    _chk(xyz, "xyz"), _chk(z, "z"), _chk(idx, "idx", torch.int32), _chk(pq, "pq"), _chk(w, "w")
    Bp, Npts, _ = xyz.shape
    if z.shape != xyz.shape or pq.shape != (Bp * Npts, 640) or w.numel() != N.PF_INTERP_BLOB:
        raise ValueError("interp: bad shapes")
    if idx.shape[:2] != (Bp, Npts) or idx.shape[2] < 8:
        raise ValueError("interp: idx must be [Bp, N, >=8]")
    out = torch.empty((Bp, Npts * R, 3), dtype=torch.float32, device=xyz.device)
    with _dev(xyz), _launch("interp"):
        N.check(lib.pf_interp(xyz.data_ptr(), z.data_ptr(), idx.data_ptr(), idx.shape[2], pq.data_ptr(), w.data_ptr(),
                              out.data_ptr(), Bp, Npts, R, _stream()), "pf_interp")
    return out


def interp_tc(xyz: torch.Tensor, z: torch.Tensor, idx: torch.Tensor, wtc: torch.Tensor, R: int, precision: int = 0):
    """Tensor-core (tcgen05, split-fp16) variant of `interp`; needs no pq.  wtc: uint8[PF_INTERP_TC_BLOB] from
    packing.pack_interp_tc."""
    _chk(xyz, "xyz"), _chk(z, "z"), _chk(idx, "idx", torch.int32), _chk(wtc, "wtc", torch.uint8)
    Bp, Npts, _ = xyz.shape
    if z.shape != xyz.shape or wtc.numel() != N.PF_INTERP_TC_BLOB:
        raise ValueError("interp_tc: bad shapes")
    if idx.shape[:2] != (Bp, Npts) or idx.shape[2] < 8:
        raise ValueError("interp_tc: idx must be [Bp, N, >=8]")
    out = torch.empty((Bp, Npts * R, 3), dtype=torch.float32, device=xyz.device)
    with _dev(xyz), _launch("interp_tc"):
        N.check(lib.pf_interp_tc(xyz.data_ptr(), z.data_ptr(), idx.data_ptr(), idx.shape[2], wtc.data_ptr(),
                                 out.data_ptr(), Bp, Npts, R, int(precision), _stream()), "pf_interp_tc")
    return out


def cnf_flow(y: torch.Tensor, ctx: torch.Tensor, blob: torch.Tensor, t0: float, t1: float, rep: int = 1, atol: float = 1e-5,
             rtol: float = 1e-5, return_nfe: bool = False):
    """One CNF block of the continuous variant (pf_cnf_flow): y[M,3] at t0 -> y at t1 under dy/dt = ODEnet([t, c], y).
    ctx[ceil(M / rep), >= PF_CNF_CTX]: the c part of the gates / biases (packing.pack_cnf_block), row p // rep for point p;
    blob: fp32[PF_CNF_BLOB / 4]."""
    _chk(y, "y"), _chk(ctx, "ctx"), _chk(blob, "blob")
    M = y.shape[0]
    if y.dim() != 2 or y.shape[1] != 3 or ctx.dim() != 2 or ctx.shape[1] < N.PF_CNF_CTX or blob.numel() * 4 != N.PF_CNF_BLOB:
        raise ValueError("cnf_flow: bad shapes")
    if ctx.shape[0] * rep < M:
        raise ValueError(f"cnf_flow: {ctx.shape[0]} conditioning rows x rep {rep} < {M} points")
    out = torch.empty_like(y)
    nfe = torch.empty(M, dtype=torch.int32, device=y.device) if return_nfe else None
    with _dev(y), _launch("cnf_flow"):
        N.check(lib.pf_cnf_flow(y.data_ptr(), ctx.data_ptr(), ctx.stride(0), blob.data_ptr(), out.data_ptr(),
                                nfe.data_ptr() if return_nfe else None, M, int(rep), float(t0), float(t1), float(atol),
                                float(rtol), _stream()), "pf_cnf_flow")
    return (out, nfe) if return_nfe else out


# ---- whole network in one C call ---------------------------------------------------------------------
class _Workspaces:
    """Scratch buffers of the one-call entry points, one per (device, stream, size): two streams (or two host threads on
    two streams) never share a buffer, a handful are kept (LRU) so that alternating shapes do not re-allocate GBs."""
    KEEP = 3

    def __init__(self):
        self.buf = {}

    def get(self, device, nbytes: int):
        key = (device.index if device.index is not None else torch.cuda.current_device(), _stream(), int(nbytes))
        ws = self.buf.pop(key, None)
        if ws is None:
            while len(self.buf) >= self.KEEP:
                self.buf.pop(next(iter(self.buf)))
            ws = torch.empty((max(int(nbytes), 1),), dtype=torch.uint8, device=device)
        self.buf[key] = ws                       # most recently used last
        return ws


_workspaces = _Workspaces()


def upsample_patches(model_blob: torch.Tensor, xyz: torch.Tensor, R: int, want_logp: bool = True, workspace=None,
                     precision: int = 0):
    """pf_upsample_patches: xyz[Bp,N,3] -> (x[Bp, N*R, 3], logp scalar tensor or None) with the whole of
    PointInterpFlow.forward (interpflow.py:327-337) launched from C on the current stream.  model_blob: uint8 CUDA tensor
    from packing.pack_model_blob.  The scratch buffer is kept per (device, stream, size) unless `workspace` is given;
    24 kernel launches (22 without logp)."""
    _chk(model_blob, "model_blob", torch.uint8), _chk(xyz, "xyz")
    if xyz.dim() != 3 or xyz.shape[2] != 3:
        raise ValueError(f"upsample_patches: expected xyz[Bp,N,3], got {tuple(xyz.shape)}")
    if model_blob.numel() != lib.pf_model_bytes():
        raise ValueError("upsample_patches: model blob has the wrong size")
    Bp, Npts, _ = xyz.shape
    need = lib.pf_upsample_workspace_bytes(Bp, Npts, R)
    if need < 0:
        raise ValueError(f"upsample_patches: bad sizes Bp={Bp} N={Npts} R={R}")
    x = torch.empty((Bp, Npts * R, 3), dtype=torch.float32, device=xyz.device)
    logp = torch.empty((), dtype=torch.float32, device=xyz.device) if want_logp else None
    with _dev(xyz):
        ws = workspace if workspace is not None else _workspaces.get(xyz.device, need)
        N.check(lib.pf_upsample_patches(model_blob.data_ptr(), xyz.data_ptr(), Bp, Npts, R, ws.data_ptr(), need,
                                        x.data_ptr(), logp.data_ptr() if want_logp else None, int(precision), _stream()),
                "pf_upsample_patches")
    if Bp:
        count_launches(network_launches(Bp, want_logp))
    elif want_logp:
        logp.fill_(float("nan"))
    return x, logp


def network_launches(Bp: int, want_logp: bool) -> int:
    return ((Bp + 65534) // 65535) + 3 * 6 + 3 + (2 if want_logp else 0)


def upsample_clouds_workspace_bytes(B, Npts, R, npoint, patch_points, n_patch) -> int:
    need = lib.pf_upsample_clouds_workspace_bytes(B, Npts, R, npoint, patch_points, n_patch)
    if need < 0:
        raise ValueError(f"upsample_clouds: bad sizes B={B} N={Npts} R={R} npoint={npoint}")
    return need


def upsample_clouds(model_blob: torch.Tensor, clouds: torch.Tensor, R: int, npoint: int, num_outliers: int,
                    patch_points: int, n_patch: int, out: torch.Tensor = None, workspace=None, precision: int = 0):
    """pf_upsample_clouds: clouds[B,N,3] -> [B, npoint - num_outliers, 3]: PatchHelper.upsample(jitter=False) (+
    remove_outliers when num_outliers > 0) launched from C on the current stream (upsample.py:46-54)."""
    _chk(model_blob, "model_blob", torch.uint8), _chk(clouds, "clouds")
    if clouds.dim() != 3 or clouds.shape[2] != 3:
        raise ValueError(f"upsample_clouds: expected clouds[B,N,3], got {tuple(clouds.shape)}")
    B, Npts, _ = clouds.shape
    need = upsample_clouds_workspace_bytes(B, Npts, R, npoint, patch_points, n_patch)
    if out is None:
        out = torch.empty((B, npoint - num_outliers, 3), dtype=torch.float32, device=clouds.device)
    else:
        _chk(out, "out")
        if tuple(out.shape) != (B, npoint - num_outliers, 3):
            raise ValueError("upsample_clouds: out has the wrong shape")
    if B == 0:
        return out
    with _dev(clouds):
        ws = workspace if workspace is not None else _workspaces.get(clouds.device, need)
        N.check(lib.pf_upsample_clouds(model_blob.data_ptr(), clouds.data_ptr(), B, Npts, R, npoint, num_outliers,
                                       patch_points, n_patch, ws.data_ptr(), need, out.data_ptr(), int(precision), _stream()),
                "pf_upsample_clouds")
    count_launches(cloud_launches(B * n_patch, num_outliers))
    return out


def cloud_launches(Bp: int, num_outliers: int) -> int:
    """kernels of one pf_upsample_clouds call: 5 before the network, the network, 3 after, + Chamfer and the outlier pass."""
    return 5 + network_launches(Bp, False) + 3 + (2 if num_outliers else 0)


# ---- caller side ---------------------------------------------------------------------------------
def fps(xyz: torch.Tensor, npoint: int, mode: str = "auto", skip_near_origin: float = 0.0):
    """xyz[B,n,3] -> idx[B,npoint] int32 (modules/utils/fps.py:40-73 semantics: starts at index 0).
    skip_near_origin = t > 0: candidates with |p|^2 <= t are never selected (upstream pointnet2_ops' rule with t = 1e-3;
    pf_fps_park_near_origin pre-pass, then the same kernels).
    mode: "brute" = cluster brute-force kernel, "bucket" = exact spatially pruned kernel (one CTA per cloud),
    "cluster" = the pruned kernel with a cloud's buckets spread over a thread-block cluster (few large clouds; falls back
    to "bucket" when the clusters of all clouds do not fit the device), "queue" = the threshold-queue kernel (one warp
    draws ~100 samples per block-wide pass), "auto" = queue for n >= 4096.
    All give identical indices."""
    _chk(xyz, "xyz")
    B, n, _ = xyz.shape
    idx = torch.empty((B, npoint), dtype=torch.int32, device=xyz.device)
    if B == 0:
        return idx
    if skip_near_origin > 0.0:
        parked = torch.empty_like(xyz)
        with _dev(xyz), _launch("fps_park_near_origin"):
            N.check(lib.pf_fps_park_near_origin(xyz.data_ptr(), B, n, float(skip_near_origin), parked.data_ptr(), _stream()),
                    "pf_fps_park_near_origin")
        xyz = parked
    with _dev(xyz), _launch(f"fps_n{n}"):
        if mode == "queue" or (mode == "auto" and n >= 4096):
            work = torch.empty((lib.pf_fps_work_floats(B, n),), dtype=torch.float32, device=xyz.device)
            N.check(lib.pf_fps_queued(xyz.data_ptr(), B, n, npoint, idx.data_ptr(), work.data_ptr(), _stream()),
                    "pf_fps_queued")
        elif mode == "cluster":
            work = torch.empty((lib.pf_fps_clustered_work_floats(B, n),), dtype=torch.float32, device=xyz.device)
            N.check(lib.pf_fps_clustered(xyz.data_ptr(), B, n, npoint, idx.data_ptr(), work.data_ptr(), _stream()),
                    "pf_fps_clustered")
        elif mode == "bucket":
            work = torch.empty((lib.pf_fps_work_floats(B, n),), dtype=torch.float32, device=xyz.device)
            N.check(lib.pf_fps_bucketed(xyz.data_ptr(), B, n, npoint, idx.data_ptr(), work.data_ptr(), _stream()),
                    "pf_fps_bucketed")
        else:
            N.check(lib.pf_fps(xyz.data_ptr(), B, n, npoint, idx.data_ptr(), None, _stream()), "pf_fps")
    return idx


def knn_patch(q: torch.Tensor, p: torch.Tensor, k: int):
    """q[B,nq,3], p[B,np,3] -> idx[B,nq,k] int32 ascending by (distance, index) (knn_cuda.KNN, patch.py:107)."""
    _chk(q, "q"), _chk(p, "p")
    B, nq, _ = q.shape
    idx = torch.empty((B, nq, k), dtype=torch.int32, device=q.device)
    with _dev(q), _launch("knn_patch"):
        N.check(lib.pf_knn_patch(q.data_ptr(), p.data_ptr(), B, nq, p.shape[1], k, idx.data_ptr(), _stream()),
                "pf_knn_patch")
    return idx


# ---- caller glue as kernels (csrc/caller.cu) ------------------------------------------------------------------
def normalize_pc(pc: torch.Tensor):
    """pc[B,n,3] -> (normalised [B,n,3], centroid [B,1,3], radius [B,1,1]); PatchHelper.normalize_pc (patch.py:168-178)."""
    _chk(pc, "pc")
    if pc.dim() != 3 or pc.shape[2] != 3:
        raise ValueError(f"normalize_pc: expected pc[B,n,3], got {tuple(pc.shape)}")
    B, n, _ = pc.shape
    out = torch.empty_like(pc)
    cen = torch.empty((B, 1, 3), dtype=torch.float32, device=pc.device)
    rad = torch.empty((B, 1, 1), dtype=torch.float32, device=pc.device)
    with _dev(pc), _launch("normalize_pc"):
        N.check(lib.pf_normalize_pc(pc.data_ptr(), B, n, out.data_ptr(), cen.data_ptr(), rad.data_ptr(), _stream()),
                "pf_normalize_pc")
    return out, cen, rad


def gather_normalize(pc: torch.Tensor, idx: torch.Tensor):
    """pc[B,n,3], idx[B,items,k] int32 -> (patches [B*items,k,3] normalised, centroid [B*items,1,3], radius [B*items,1,1])."""
    _chk(pc, "pc"), _chk(idx, "idx", torch.int32)
    B, n, _ = pc.shape
    _, items, k = idx.shape
    out = torch.empty((B * items, k, 3), dtype=torch.float32, device=pc.device)
    cen = torch.empty((B * items, 1, 3), dtype=torch.float32, device=pc.device)
    rad = torch.empty((B * items, 1, 1), dtype=torch.float32, device=pc.device)
    with _dev(pc), _launch("gather_normalize"):
        N.check(lib.pf_gather_normalize(pc.data_ptr(), idx.data_ptr(), B, n, items, k, out.data_ptr(), cen.data_ptr(),
                                        rad.data_ptr(), _stream()), "pf_gather_normalize")
    return out, cen, rad


def gather_points(pc: torch.Tensor, idx: torch.Tensor, scale=None, shift=None):
    """out[b,i] = pc[b, idx[b,i]] (* scale[b] + shift[b]); pc[B,n,3], idx[B,m] int32."""
    _chk(pc, "pc"), _chk(idx, "idx", torch.int32)
    B, n, _ = pc.shape
    m = idx.shape[1]
    out = torch.empty((B, m, 3), dtype=torch.float32, device=pc.device)
    if scale is not None:
        _chk(scale, "scale"), _chk(shift, "shift")
    with _dev(pc), _launch("gather_points"):
        N.check(lib.pf_gather_points(pc.data_ptr(), idx.data_ptr(), B, n, m, scale.data_ptr() if scale is not None else None,
                                     shift.data_ptr() if shift is not None else None, out.data_ptr(), _stream()),
                "pf_gather_points")
    return out


def denorm_concat(x: torch.Tensor, patches: torch.Tensor, cen: torch.Tensor, rad: torch.Tensor):
    """cat([x, patches], 1) * rad + cen; x[I,nx,3], patches[I,np,3] -> [I, nx+np, 3] (patch.py:90-91)."""
    _chk(x, "x"), _chk(patches, "patches"), _chk(cen, "cen"), _chk(rad, "rad")
    I, nx, _ = x.shape
    np_ = patches.shape[1]
    out = torch.empty((I, nx + np_, 3), dtype=torch.float32, device=x.device)
    with _dev(x), _launch("denorm_concat"):
        N.check(lib.pf_denorm_concat(x.data_ptr(), patches.data_ptr(), cen.data_ptr(), rad.data_ptr(), I, nx, np_,
                                     out.data_ptr(), _stream()), "pf_denorm_concat")
    return out


def remove_outliers(sr: torch.Tensor, dist1: torch.Tensor, num_outliers: int):
    """sr[B,n,3], dist1[B,n] -> [B, n - num_outliers, 3]: the num_outliers largest dist1 dropped (ties: lowest index first),
    order kept (patch.py:205-214)."""
    _chk(sr, "sr"), _chk(dist1, "dist1")
    B, n, _ = sr.shape
    if num_outliers <= 0:
        return sr.clone()
    out = torch.empty((B, n - num_outliers, 3), dtype=torch.float32, device=sr.device)
    with _dev(sr), _launch("remove_outliers"):
        N.check(lib.pf_remove_outliers(sr.data_ptr(), dist1.data_ptr(), B, n, num_outliers, out.data_ptr(), _stream()),
                "pf_remove_outliers")
    return out
