"""One-time weight packing: reference checkpoint layout (408-key state_dict of PointInterpFlow,
/root/reference/modules/discrete/interpflow.py:262-290) -> the k-major fp32 blobs the
kernels consume.  All algebra is done in float64 and rounded to fp32 once.

What is folded / composed here (each is exact algebra, so only fp32 summation order changes):

* eval-mode BatchNorm2d (eps 1e-5) into the preceding 1x1 conv:  W' = W g/sqrt(v+eps),
  b' = (b-mean) g/sqrt(v+eps) + beta                                   (interpflow.py:91-98,144-151,203-221)
* the edge feature [x_i, x_j, x_j - x_i] (interpflow.py:223-232) is linear in (x_i, x_j):
  W.[x_i; x_j; x_j-x_i] = (W1-W3).x_i + (W2+W3).x_j  -> per-point projections P (own point, carries
  every bias) and Q (neighbour) for all five convs of a dense block at once;
  likewise the distance encoder's [x_i, x_j, x_i-x_j, |.|] (interpflow.py:106-114);
* DistanceEncoder.mlp.6, feat_conv.conv_out and WeightEstimationUnit.mlp.0 (+BN) have no activation
  between them (interpflow.py:98,133-134,145-146,242) and are composed into one 192->128 layer;
* InvertibleConv1x1_1D's inverse is computed once (float64) instead of torch.inverse per call
  (/root/reference/modules/flows/permutate.py:122-126); exp(+-logs) of ActNorm likewise
  (/root/reference/modules/flows/normalize.py:30-43).
"""
from typing import Dict

import torch

from ._native import chain_tc_bytes, PF_COND_BLOB, PF_COND_TC_BLOCK, PF_FLOW_BLOB, PF_FLOW_TC_BLOCK, PF_INTERP_BLOB

BN_EPS = 1e-5
NUM_BLOCKS = 6
# (C_in, growth, C_out) of feat_convs[i] and cond dim of merge_convs[i]  (interpflow.py:273-275)
FEAT_CFG = [(3, 8, 32), (32, 16, 64), (64, 32, 128), (128, 32, 128), (128, 32, 128), (128, 32, 128)]


def _f64(t):
    return t.detach().to("cpu", torch.float64)


def _conv_bn(sd, conv: str, bn: str = None):
    """(W[O,I], b[O]) of a 1x1 conv, with the following eval-mode BN folded in when `bn` is given."""
    W = _f64(sd[conv + ".weight"])
    W = W.reshape(W.shape[0], W.shape[1])
    b = _f64(sd[conv + ".bias"])
    if bn is not None:
        s = _f64(sd[bn + ".weight"]) / torch.sqrt(_f64(sd[bn + ".running_var"]) + BN_EPS)
        W = W * s[:, None]
        b = (b - _f64(sd[bn + ".running_mean"])) * s + _f64(sd[bn + ".bias"])
    return W, b


def _split_edge(W, C):
    """Columns [x_i | x_j | x_j - x_i | hidden] -> (A: coeff of x_i, B: coeff of x_j, Wh)."""
    W1, W2, W3, Wh = W[:, :C], W[:, C:2 * C], W[:, 2 * C:3 * C], W[:, 3 * C:]
    return W1 - W3, W2 + W3, Wh


def pack_edgeconv(sd, prefix: str, C: int, g: int, C_out: int, layers: int = 4):
    """FeatureExtractUnit (interpflow.py:190-248) -> dict(wpq[C,2R], bpq[2R], wd[...], wo[layers*g, C_out])."""
    A_rows, B_rows, biases, wd = [], [], [], []
    for l in range(layers):
        W, b = _conv_bn(sd, f"{prefix}.convs.{l}.0", f"{prefix}.convs.{l}.1")
        assert W.shape == (g, 3 * C + g * l), (prefix, l, W.shape)
        A, B, Wh = _split_edge(W, C)
        A_rows.append(A), B_rows.append(B), biases.append(b)
        if l > 0:
            wd.append(Wh.t().contiguous().reshape(-1))          # [g*l][g] k-major
    W, b = _conv_bn(sd, f"{prefix}.conv_out")
    assert W.shape == (C_out, 3 * C + g * layers)
    A, B, Wh = _split_edge(W, C)
    A_rows.append(A), B_rows.append(B), biases.append(b)
    A_all, B_all, b_all = torch.cat(A_rows), torch.cat(B_rows), torch.cat(biases)     # [R, C], [R]
    R = A_all.shape[0]
    return dict(
        wpq=torch.cat([A_all.t(), B_all.t()], dim=1).contiguous(),                   # [C, 2R]
        bpq=torch.cat([b_all, torch.zeros(R, dtype=torch.float64)]),
        wd=torch.cat(wd),
        wo=Wh.t().contiguous(),                                                       # [layers*g, C_out]
    )


TC_WSCALE = 1024.0      # = 1 / tc::WINV (csrc/tc.cuh)


def split_weight(mat):
    """float64/32 weights -> (hi, lo) fp16 with hi + lo ~ fp32(w) * TC_WSCALE to 22 bits.  Without the power-of-two
    factor the lo half of a typical |w| ~ 0.03 falls into the fp16 subnormals (spacing 2^-24) and the product loses
    3 bits against fp32; the kernels' epilogues multiply the accumulator by 2^-10 (exact).  |w| * 1024 must stay
    below the fp16 maximum: a checkpoint with |w| >= 60 is refused (run it with PUFlowEngine(kernels="ffma") / net.kernels = "ffma")."""
    w32 = mat.to(torch.float32)
    if w32.numel() and float(w32.abs().max()) >= 60.0:
        raise ValueError(f"weight magnitude {float(w32.abs().max()):.3g} is outside the tensor-core path's fp16 range")
    w32 = w32 * TC_WSCALE
    hi = w32.to(torch.float16)
    lo = (w32 - hi.to(torch.float32)).to(torch.float16)
    return hi, lo


def umma_kmajor_fp16(mat):
    """[rows, Kt] float64 -> (hi, lo) uint8 byte strings in the K-major no-swizzle tensor-core operand layout of
    csrc/tc.cuh: half index (r//8)*(Kt*8) + (k//8)*64 + (r%8)*8 + (k%8).  (hi, lo) = split_weight(w)."""
    rows, Kt = mat.shape
    assert rows % 8 == 0 and Kt % 16 == 0, (rows, Kt)
    hi, lo = split_weight(mat)

    def lay(x):
        return x.reshape(rows // 8, 8, Kt // 8, 8).permute(0, 2, 1, 3).contiguous().view(torch.uint8).reshape(-1)
    return lay(hi), lay(lo)


def edgeconv_tc_blob_bytes(g: int) -> int:
    gp = max(g, 16)
    return 24 * gp * gp + 2 * 128 * 4 * gp * 2


def pack_edgeconv_tc(wd, wo, g: int = 32, C_out: int = 128, layers: int = 4):
    """(wd, wo) of pack_edgeconv -> uint8[edgeconv_tc_blob_bytes(g)] for pf_edgeconv_tc, (g, C_out) in
    {(8, 32), (16, 64), (32, 128)}:  [S0 hi | S0 lo | S1 hi | S1 lo | S2 hi | S2 lo | Wo hi | Wo lo], S_m = the columns of
    K-slice m (layer m's hidden channels) of the weights of layers m+1..3 stacked along the rows ([gp (3 - m) out][gp in]):
    the kernel multiplies every new activation slice against all later layers at once.
    gp = max(g, 16) is the width of one layer's K slice in the activation operand; for g = 8 hidden channel k of layer j
    sits at K index j*gp + k and the rest of the slice is zero.  S_m in the shared-memory operand layout; Wo: plain row-major fp16 [128][layers*gp] whose row r is conv_out channel r % C_out (it is copied to tensor
    memory, one row per lane; the kernel reads channel r % C_out for edge group r // C_out from lane r)."""
    assert (g, C_out) in ((8, 32), (16, 64), (32, 128)) and layers == 4
    gp = max(g, 16)

    def pad_k(W, nl):            # [rows, nl*g] -> [rows, nl*gp]
        out = torch.zeros(W.shape[0], nl * gp, dtype=torch.float64)
        for j in range(nl):
            out[:, j * gp:j * gp + g] = W[:, j * g:(j + 1) * g]
        return out

    Ws, off = {}, 0
    for l in range(1, layers):
        Wl = wd[off:off + g * l * g].reshape(g * l, g).t().double()   # [g out, g*l in]
        off += g * l * g
        Wp = torch.zeros(gp, l * gp, dtype=torch.float64)
        Wp[:g] = pad_k(Wl, l)
        Ws[l] = Wp
    parts = []
    for m in range(layers - 1):      # K-slice m (hidden channels of layer m) against every later layer, stacked along the rows
        parts.extend(umma_kmajor_fp16(torch.cat([Ws[l][:, m * gp:(m + 1) * gp] for l in range(m + 1, layers)], 0)))
    wo_rows = pad_k(wo.t().double(), layers)                          # [C_out, layers*gp]
    wo32 = wo_rows.repeat(128 // C_out, 1).contiguous().to(torch.float32)   # row r = channel r % C_out
    wo_hi, wo_lo = split_weight(wo32)
    parts.extend([wo_hi.contiguous().view(torch.uint8).reshape(-1), wo_lo.contiguous().view(torch.uint8).reshape(-1)])
    blob = torch.cat(parts)
    assert blob.numel() == edgeconv_tc_blob_bytes(g), blob.numel()
    return blob


def linear_tc_ok(K: int, N: int) -> bool:
    return K in (16, 32, 64, 128) and N % 16 == 0


def pack_linear_tc(W):
    """W[K, N] (k-major, float64/float32) -> uint8 blob for pf_linear_tc: per 128-column chunk
    [hi (nc*K halfs) | lo (nc*K halfs) | pad to 128*K*4 bytes], each part in the csrc/tc.cuh operand layout."""
    K, Nn = W.shape
    assert linear_tc_ok(K, Nn), (K, Nn)
    chunks = []
    for n0 in range(0, Nn, 128):
        nc = min(128, Nn - n0)
        hi, lo = umma_kmajor_fp16(W[:, n0:n0 + nc].t().contiguous().double())
        pad = torch.zeros(128 * K * 4 - hi.numel() - lo.numel(), dtype=torch.uint8)
        chunks.extend([hi, lo, pad])
    return torch.cat(chunks)


def pack_interp(sd):
    """InterpolationModule (interpflow.py:85-186) -> dict(wpq[3,640], bpq[640], blob[PF_INTERP_BLOB])."""
    p = "interp.knn_context.distance_encoder.mlp"
    Wd0, bd0 = _conv_bn(sd, p + ".0", p + ".1")            # [64,10] on [x_i, x_j, x_i-x_j, dist]
    Wd3, bd3 = _conv_bn(sd, p + ".3", p + ".4")            # [64,64]
    Wd6, bd6 = _conv_bn(sd, p + ".6")                      # [128,64]
    Pd, Qd, wdist = Wd0[:, 0:3] + Wd0[:, 6:9], Wd0[:, 3:6] - Wd0[:, 6:9], Wd0[:, 9]

    e = pack_edgeconv(sd, "interp.knn_context.feat_conv", 3, 16, 128, layers=8)
    # e["wpq"]: [3, 2*256] with R = 8*16 + 128: columns [dense 128 | conv_out 128] for P then Q
    Ae, Be = e["wpq"][:, :256].t(), e["wpq"][:, 256:].t()  # [256,3]
    be = e["bpq"][:256]
    Wout_h = e["wo"].t()                                   # [128 out, 128 hidden]

    q = "interp.weight_unit.mlp"
    W0, b0 = _conv_bn(sd, q + ".0", q + ".1")              # [128,256] on [dist feats 128 | context feats 128]
    W3, b3 = _conv_bn(sd, q + ".3", q + ".4")              # [64,128]
    W6, b6 = _conv_bn(sd, q + ".6")                        # [32,64]
    W0a, W0b = W0[:, :128], W0[:, 128:]
    M1 = W0a @ Wd6                                         # [128,64]   acts on d2
    M2 = W0b @ Wout_h                                      # [128,128]  acts on the dense block's h
    Pm = W0b @ Ae[128:]                                    # [128,3]
    Qm = W0b @ Be[128:]
    bm = W0a @ bd6 + W0b @ be[128:] + b0

    P = torch.cat([Pd, Ae[:128], Pm])                      # [320,3]
    Q = torch.cat([Qd, Be[:128], Qm])
    bias = torch.cat([bd0, be[:128], bm, torch.zeros(320, dtype=torch.float64)])
    w6 = torch.zeros(64, 32, dtype=torch.float64)
    w6[:, :W6.shape[0]] = W6.t()
    b6p = torch.zeros(32, dtype=torch.float64)
    b6p[:b6.shape[0]] = b6
    blob = torch.cat([
        wdist,                                 # 0
        Wd3.t().reshape(-1),                   # 64     [64][64]
        bd3,                                   # 4160
        e["wd"],                               # 4224   layers 1..7, [16 l][16]
        W3.t().reshape(-1),                    # 11392  [128][64]
        b3,                                    # 19584
        w6.reshape(-1),                        # 19648  [64][32]
        b6p,                                   # 21696
        M1.t().reshape(-1),                    # 21728  [64][128]
        M2.t().reshape(-1),                    # 29920  [128][128]
    ])
    assert blob.numel() == PF_INTERP_BLOB, blob.numel()
    return dict(wpq=torch.cat([P.t(), Q.t()], dim=1).contiguous(), bpq=bias, blob=blob)


INTERP_TC_SMEM = 98944        # bytes of the pf_interp_tc blob that live in shared memory
INTERP_TC_BLOB = INTERP_TC_SMEM + 2 * 128 * 104 * 4


def pack_interp_tc(sd):
    """InterpolationModule (interpflow.py:85-186) -> uint8[INTERP_TC_BLOB] for pf_interp_tc.

    Every conv that reads raw coordinates takes them from a 16-wide "geo" K-slice of the activation operand,
        geo = [x_i (3) | x_j (3) | x_j - x_i (3) | |x_i - x_j| | 1 | 0 x 5],
    with the reference's UNFOLDED weights (no (W1-W3) x_i + (W2+W3) x_j cancellation) and the bias on the constant 1.
    Shared-memory part (B operands, [N][K] K-major, fp16 hi then lo each):
        Wd1g[64][16] | Wd2[64][64] | Wg[128][16] (the geo columns of the eight dense layers, layer l = rows 16l..)
        | S_m[16 (7 - m)][16] (m = 0..6: the h_m columns of layers m+1..7 -- the kernel multiplies every new slice h_m
        against all later layers at once) | W3[64][128] | W6[32][64] | fp32 bd3[64] | b3[64] | b6[32]
    Tensor-memory part (A operand of the transposed composed layer, plain row-major fp16 [128][208], hi then lo):
        M12g = [geo part | M2 (h, 128) | M1 (d2, 64)]."""
    p = "interp.knn_context.distance_encoder.mlp"
    Wd0, bd0 = _conv_bn(sd, p + ".0", p + ".1")            # [64,10] on [x_i, x_j, x_i-x_j, dist]
    Wd3, bd3 = _conv_bn(sd, p + ".3", p + ".4")            # [64,64]
    Wd6, bd6 = _conv_bn(sd, p + ".6")                      # [128,64]

    def geo(Wxi, Wxj, Wdiff, wdist, bias):
        g = torch.zeros(Wxi.shape[0], 16, dtype=torch.float64)
        g[:, 0:3], g[:, 3:6], g[:, 6:9] = Wxi, Wxj, Wdiff
        if wdist is not None:
            g[:, 9] = wdist
        g[:, 10] = bias
        return g

    parts = []

    def add(mat):
        parts.extend(umma_kmajor_fp16(mat))

    add(geo(Wd0[:, 0:3], Wd0[:, 3:6], -Wd0[:, 6:9], Wd0[:, 9], bd0))          # d1 = lrelu(Wd1g . geo)
    add(Wd3)                                                                     # d2 = lrelu(Wd3 . d1 + bd3)
    fc = "interp.knn_context.feat_conv"
    full = []                                                                    # layer l on [geo (16) | h_0 .. h_{l-1}]
    for l in range(8):
        W, b = _conv_bn(sd, f"{fc}.convs.{l}.0", f"{fc}.convs.{l}.1")           # [16, 9 + 16 l]
        full.append(torch.cat([geo(W[:, 0:3], W[:, 3:6], W[:, 6:9], None, b), W[:, 9:]], dim=1))
    add(torch.cat([w[:, 0:16] for w in full]))                                   # geo slice of all eight layers [128, 16]
    for m in range(7):                                                           # S_m: h_m's columns of layers m+1..7
        add(torch.cat([full[j][:, 16 + 16 * m:32 + 16 * m] for j in range(m + 1, 8)]))
    Wco, bco = _conv_bn(sd, f"{fc}.conv_out")                                   # [128, 9 + 128]
    q = "interp.weight_unit.mlp"
    W0, b0 = _conv_bn(sd, q + ".0", q + ".1")              # [128,256] on [dist feats 128 | context feats 128]
    W3, b3 = _conv_bn(sd, q + ".3", q + ".4")              # [64,128]
    W6, b6 = _conv_bn(sd, q + ".6")                        # [32,64]
    add(W3)
    w6 = torch.zeros(32, 64, dtype=torch.float64)
    w6[:W6.shape[0]] = W6
    add(w6)
    b6p = torch.zeros(32, dtype=torch.float64)
    b6p[:b6.shape[0]] = b6
    parts.append(torch.cat([bd3, b3, b6p]).to(torch.float32).contiguous().view(torch.uint8).reshape(-1))
    smem = torch.cat(parts)
    assert smem.numel() == INTERP_TC_SMEM, smem.numel()

    W0a, W0b = W0[:, :128], W0[:, 128:]
    G = W0b @ Wco[:, 0:9]                                  # composed geo part [128, 9]
    bm = W0a @ bd6 + W0b @ bco + b0
    M12g = torch.cat([geo(G[:, 0:3], G[:, 3:6], G[:, 6:9], None, bm), W0b @ Wco[:, 9:], W0a @ Wd6], dim=1)   # [128, 208]
    hi, lo = split_weight(M12g)
    blob = torch.cat([smem, hi.contiguous().view(torch.uint8).reshape(-1), lo.contiguous().view(torch.uint8).reshape(-1)])
    assert blob.numel() == INTERP_TC_BLOB, blob.numel()
    return blob


def identity_flow_blob():
    """A flow-step blob whose every sub-layer is the identity (float64); the set_* helpers below fill in
    the sub-layers a caller wants.  Lets the single modules/flows layer classes run on the fused kernel."""
    flow = torch.zeros(PF_FLOW_BLOB, dtype=torch.float64)
    flow[0:3], flow[6:9] = 1.0, 1.0
    flow[9:18] = flow[18:27] = torch.eye(3, dtype=torch.float64).reshape(-1)
    flow[27], flow[28] = 1.0, 0.0
    return flow


def set_actnorm(flow, logs, bias):
    """ActNorm (modules/flows/normalize.py:30-43): z = x exp(logs) + bias."""
    logs, bias = _f64(logs).reshape(3), _f64(bias).reshape(3)
    flow[0:3], flow[3:6], flow[6:9] = torch.exp(logs), bias, torch.exp(-logs)


def set_matrix(flow, W):
    """InvertibleConv1x1_1D (modules/flows/permutate.py:117-126) or any 3x3 channel mix: z_i = sum_j W[i][j] x_j."""
    W = _f64(W)
    flow[9:18], flow[18:27] = W.reshape(-1), torch.linalg.inv(W).reshape(-1)


def set_coupling1(flow, W1, W2, b2, W4, b4, tdim: int):
    """Additive coupling with LinearA1D(tdim + cdim -> 64 -> 64 -> 3 - tdim) (interpflow.py:22-43,55-58).
    Returns the conditioning part of the first layer, [cdim, 64] k-major."""
    W1, W4 = _f64(W1), _f64(W4)
    assert W1.shape[0] == 64 and W4.shape == (3 - tdim, 64), "kernels are built for hidden width 64, 3 channels"
    w1h = torch.zeros(2, 64, dtype=torch.float64)
    w1h[:tdim] = W1[:, :tdim].t()
    w4 = torch.zeros(64, 2, dtype=torch.float64)
    w4[:, :3 - tdim] = W4.t()
    b4p = torch.zeros(2, dtype=torch.float64)
    b4p[:3 - tdim] = _f64(b4)
    flow[27] = float(tdim)
    flow[32:160] = w1h.reshape(-1)
    flow[160:224] = _f64(b2)
    flow[224:352] = w4.reshape(-1)
    flow[352:354] = b4p
    flow[384:4480] = _f64(W2).t().reshape(-1)
    return W1[:, tdim:].t().contiguous()


def zero_cond_blob():
    return torch.zeros(PF_COND_BLOB, dtype=torch.float64)


def set_injector(cond, scale_layers, bias_layers):
    """Affine injector nets (coupling.py:127-151): two LinearA1D(cdim -> 64 -> 64 -> 3).  *_layers =
    (W0, W2, b2, W4, b4).  Returns the two first-layer matrices [cdim, 64] k-major (scale, bias)."""
    (Ws0, Ws2, bs2, Ws4, bs4), (Wb0, Wb2, bb2, Wb4, bb4) = scale_layers, bias_layers
    cond[0:4096] = _f64(Ws2).t().reshape(-1)
    cond[4096:8192] = _f64(Wb2).t().reshape(-1)
    cond[8192:8256], cond[8256:8320] = _f64(bs2), _f64(bb2)
    ws4, wb4 = torch.zeros(64, 4, dtype=torch.float64), torch.zeros(64, 4, dtype=torch.float64)
    ws4[:, :3], wb4[:, :3] = _f64(Ws4).t(), _f64(Wb4).t()
    cond[8320:8576], cond[8576:8832] = ws4.reshape(-1), wb4.reshape(-1)
    cond[8832:8835], cond[8836:8839] = _f64(bs4), _f64(bb4)
    return _f64(Ws0).t().contiguous(), _f64(Wb0).t().contiguous()


def pack_flow_tc(flow):
    """flow[nb, PF_FLOW_BLOB] (float64/32) -> uint8[nb, PF_FLOW_TC_BLOCK] for pf_flow_fwd_tc / pf_flow_inv_tc:
    per block W2 hi | W2 lo ([64 out][64 in], csrc/tc.cuh operand layout) | the first 384 floats as fp32."""
    out = []
    for f in flow:
        W2 = f[384:4480].reshape(64, 64).t().contiguous().double()        # [out, in]
        hi, lo = umma_kmajor_fp16(W2)
        out.append(torch.cat([hi, lo, f[:384].to(torch.float32).contiguous().view(torch.uint8)]))
    blob = torch.stack(out)
    assert blob.shape[1] == PF_FLOW_TC_BLOCK, blob.shape
    return blob


def pack_cond_tc(cond):
    """cond[nb, PF_COND_BLOB] -> uint8[nb, PF_COND_TC_BLOCK] for pf_flow_cond_tc:
    per block Ws2 hi | Ws2 lo | Wb2 hi | Wb2 lo | floats 8192.. (biases and output layers) as fp32."""
    out = []
    for c in cond:
        parts = []
        for net in range(2):
            W = c[4096 * net:4096 * (net + 1)].reshape(64, 64).t().contiguous().double()
            parts.extend(umma_kmajor_fp16(W))
        parts.append(c[8192:].to(torch.float32).contiguous().view(torch.uint8))
        out.append(torch.cat(parts))
    blob = torch.stack(out)
    assert blob.shape[1] == PF_COND_TC_BLOCK, blob.shape
    return blob


def pack_chain_tc(W1, b1, W2, wc, cond):
    """Fused conditioning chain of one block (pf_cond_chain_tc).  W1[C/2, C], b1[C/2], W2[cond_dim, C/2]: FeatMergeUnit
    conv1 / conv2 (interpflow.py:251-258); wc[cond_dim, 192]: first layers of the three LinearA1D nets (k-major);
    cond: the block's fp32 cond blob.  conv2 (linear, bias-free) is composed with wc in float64."""
    W1, W2, wc = _f64(W1), _f64(W2), _f64(wc)
    H1, C = W1.shape
    assert C in (32, 64, 128) and H1 == C // 2 and W2.shape[1] == H1 and wc.shape == (W2.shape[0], 192)
    W2c = wc.t() @ W2                                                # [192 out, C/2 in]
    parts = list(umma_kmajor_fp16(W1)) + list(umma_kmajor_fp16(W2c))
    for net in range(2):
        parts.extend(umma_kmajor_fp16(cond[4096 * net:4096 * (net + 1)].reshape(64, 64).t().contiguous().double()))
    b1p = torch.zeros(64, dtype=torch.float32)
    b1p[:H1] = _f64(b1).to(torch.float32)
    parts.append(b1p.view(torch.uint8))
    parts.append(cond[8192:].to(torch.float32).contiguous().view(torch.uint8))
    blob = torch.cat(parts)
    assert blob.numel() == chain_tc_bytes(C), (blob.numel(), C)
    return blob


def _a1d(sd, prefix):
    return (sd[prefix + ".0.weight"], sd[prefix + ".2.weight"], sd[prefix + ".2.bias"], sd[prefix + ".4.weight"],
            sd[prefix + ".4.bias"])


def pack_flow_block(sd, i_or_prefix, tdim: int = None):
    """FlowBlock (interpflow.py:46-82) -> dict(wc[cond,192], flow[PF_FLOW_BLOB], cond[PF_COND_BLOB],
    logdet_const); logdet_const = sum(logs) + log|det W| (per point).  `i_or_prefix`: block index of a
    PointInterpFlow state_dict, or the key prefix of a lone FlowBlock's state_dict ('' allowed)."""
    if isinstance(i_or_prefix, int):
        p = f"flow_blocks.{i_or_prefix}."
        tdim = 1 if i_or_prefix % 2 == 0 else 2             # interpflow.py:55 (is_even = i % 2 == 0, :283)
    else:
        p = i_or_prefix
    logs, W = _f64(sd[p + "actnorm.logs"]).reshape(3), _f64(sd[p + "permutate1.permutater.W"])
    direct = sd[p + "permutate2.permutater.direct_idx"].tolist()
    assert direct == [2, 1, 0], "the fused flow kernel implements the 'reverse' permutation (interpflow.py:53)"
    flow, cond = identity_flow_blob(), zero_cond_blob()
    set_actnorm(flow, logs, sd[p + "actnorm.bias"])
    set_matrix(flow, W)
    flow[28] = 1.0
    w1c = set_coupling1(flow, *_a1d(sd, p + "coupling1.bias_net.layers"), tdim)
    ws0, wb0 = set_injector(cond, _a1d(sd, p + "coupling2.scale_net.layers"), _a1d(sd, p + "coupling2.bias_net.layers"))
    wc = torch.cat([w1c, ws0, wb0], dim=1)
    logdet_const = float(logs.sum() + torch.linalg.slogdet(W)[1])
    return dict(wc=wc.contiguous(), flow=flow, cond=cond, logdet_const=logdet_const)


def pack_state_dict(sd: Dict[str, torch.Tensor], device="cpu") -> Dict:
    """Whole model -> nested dict of contiguous fp32 tensors on `device`."""
    out = {"feat": [], "merge": [], "wc": [], "wc_tc": [], "logdet_const": 0.0}
    flows, conds = [], []
    for i, (C, g, Co) in enumerate(FEAT_CFG):
        out["feat"].append(pack_edgeconv(sd, f"feat_convs.{i}", C, g, Co))
        out["feat"][-1]["wtc"] = pack_edgeconv_tc(out["feat"][-1]["wd"], out["feat"][-1]["wo"], g, Co)
        if linear_tc_ok(*out["feat"][-1]["wpq"].shape):
            out["feat"][-1]["wpq_tc"] = pack_linear_tc(out["feat"][-1]["wpq"])
        out["merge"].append(dict(
            w1=_f64(sd[f"merge_convs.{i}.conv1.weight"]).t().contiguous(),
            b1=_f64(sd[f"merge_convs.{i}.conv1.bias"]),
            w2=_f64(sd[f"merge_convs.{i}.conv2.weight"]).t().contiguous()))
        for nm in ("w1", "w2"):
            if linear_tc_ok(*out["merge"][-1][nm].shape):
                out["merge"][-1][nm + "_tc"] = pack_linear_tc(out["merge"][-1][nm])
        fb = pack_flow_block(sd, i)
        out["wc"].append(fb["wc"])
        out["wc_tc"].append(pack_linear_tc(fb["wc"]) if linear_tc_ok(*fb["wc"].shape) else None)
        flows.append(fb["flow"]), conds.append(fb["cond"])
        out.setdefault("chain_tc", []).append(pack_chain_tc(
            sd[f"merge_convs.{i}.conv1.weight"], sd[f"merge_convs.{i}.conv1.bias"], sd[f"merge_convs.{i}.conv2.weight"],
            fb["wc"], fb["cond"]))
        out["logdet_const"] += fb["logdet_const"]
    out["flow"] = torch.stack(flows)
    out["cond"] = torch.stack(conds)
    out["flow_tc"] = pack_flow_tc(out["flow"])
    out["cond_tc"] = pack_cond_tc(out["cond"])
    out["interp"] = pack_interp(sd)
    out["interp"]["tc"] = pack_interp_tc(sd)

    def cast(o):
        if isinstance(o, torch.Tensor):
            if o.dtype == torch.uint8:
                return o.contiguous().to(device)
            return o.to(torch.float32).contiguous().to(device)
        if isinstance(o, dict):
            return {k: cast(v) for k, v in o.items()}
        if isinstance(o, list):
            return [cast(v) for v in o]
        return o
    return cast(out)


CNF_CTX, CNF_CTX_LD, CNF_BLOB_FLOATS = 264, 272, 4880


def pack_cnf_block(sd, i: int):
    """FlowBlock i of the continuous variant (modules/continuous/interpflow.py:19-31: ODEnet with hidden (64, 64),
    ConcatSquashLinear layers, diffeq_layers.py:72-86) -> dict for pf_cnf_flow:
      wctx[cdim, 272], bctx[272]  the c part of every gate / bias pre-activation, one GEMM per input point:
                                  columns [0,64) gate0 | [64,128) bias0 | [128,192) gate1 | [192,256) bias1 | [256,259) gate2 |
                                  [259,262) bias2 | zero padding (the hyper layers' column 0 multiplies t and goes to the blob)
      blob[4880] fp32             (W0[j][0..2], b0[j]) x 64 | gate0_t[64] | bias0_t[64] | W1[64][64] | b1 | gate1_t | bias1_t |
                                  W2[3][64] | b2[3] | gate2_t[3] | bias2_t[3] | pad
      T                           sqrt_end_time ** 2 (cnf.py:74-77)."""
    p = f"flow_blocks.{i}.cnf.odefunc.diffeq.layers."
    L = [{k: _f64(sd[f"{p}{l}.{k}"]) for k in ("_layer.weight", "_layer.bias", "_hyper_bias.weight", "_hyper_gate.weight",
                                                "_hyper_gate.bias")} for l in range(3)]
    assert L[0]["_layer.weight"].shape == (64, 3) and L[1]["_layer.weight"].shape == (64, 64) and \
        L[2]["_layer.weight"].shape == (3, 64), "pf_cnf_flow is built for ODEnet hidden dims (64, 64) on 3 channels"
    cdim = L[0]["_hyper_gate.weight"].shape[1] - 1
    wctx = torch.zeros(cdim, CNF_CTX_LD, dtype=torch.float64)
    bctx = torch.zeros(CNF_CTX_LD, dtype=torch.float64)
    col = 0
    tcols = []
    for l in range(3):
        n = L[l]["_layer.weight"].shape[0]
        wctx[:, col:col + n] = L[l]["_hyper_gate.weight"][:, 1:].t()
        bctx[col:col + n] = L[l]["_hyper_gate.bias"]
        wctx[:, col + n:col + 2 * n] = L[l]["_hyper_bias.weight"][:, 1:].t()
        tcols.append((L[l]["_hyper_gate.weight"][:, 0], L[l]["_hyper_bias.weight"][:, 0]))
        col += 2 * n
    assert col == 262
    blob = torch.cat([
        torch.cat([L[0]["_layer.weight"], L[0]["_layer.bias"][:, None]], dim=1).reshape(-1), tcols[0][0], tcols[0][1],
        L[1]["_layer.weight"].reshape(-1), L[1]["_layer.bias"], tcols[1][0], tcols[1][1],
        L[2]["_layer.weight"].reshape(-1), L[2]["_layer.bias"], tcols[2][0], tcols[2][1],
        torch.zeros(CNF_BLOB_FLOATS - 4873, dtype=torch.float64)])
    assert blob.numel() == CNF_BLOB_FLOATS
    out = dict(wctx=wctx.contiguous(), bctx=bctx, blob=blob, T=float(sd[f"flow_blocks.{i}.cnf.sqrt_end_time"]) ** 2)
    if linear_tc_ok(*wctx.shape):
        out["wctx_tc"] = pack_linear_tc(wctx)
    return out


def pack_cnf_state_dict(sd: Dict[str, torch.Tensor], device="cpu") -> Dict:
    """The continuous variant (modules/continuous/interpflow.py:53-84): the discrete model's feature extractor, merge units
    and interpolation module (same keys, same kernels) + six CNF blocks."""
    out = {"feat": [], "merge": [], "cnf": []}
    for i, (C, g, Co) in enumerate(FEAT_CFG):
        out["feat"].append(pack_edgeconv(sd, f"feat_convs.{i}", C, g, Co))
        out["feat"][-1]["wtc"] = pack_edgeconv_tc(out["feat"][-1]["wd"], out["feat"][-1]["wo"], g, Co)
        if linear_tc_ok(*out["feat"][-1]["wpq"].shape):
            out["feat"][-1]["wpq_tc"] = pack_linear_tc(out["feat"][-1]["wpq"])
        out["merge"].append(dict(
            w1=_f64(sd[f"merge_convs.{i}.conv1.weight"]).t().contiguous(),
            b1=_f64(sd[f"merge_convs.{i}.conv1.bias"]),
            w2=_f64(sd[f"merge_convs.{i}.conv2.weight"]).t().contiguous()))
        for nm in ("w1", "w2"):
            if linear_tc_ok(*out["merge"][-1][nm].shape):
                out["merge"][-1][nm + "_tc"] = pack_linear_tc(out["merge"][-1][nm])
        out["cnf"].append(pack_cnf_block(sd, i))
    out["interp"] = pack_interp(sd)
    out["interp"]["tc"] = pack_interp_tc(sd)
    return _cast_tree(out, device)


def _cast_tree(o, device):
    if isinstance(o, torch.Tensor):
        if o.dtype == torch.uint8:
            return o.contiguous().to(device)
        return o.to(torch.float32).contiguous().to(device)
    if isinstance(o, dict):
        return {k: _cast_tree(v, device) for k, v in o.items()}
    if isinstance(o, list):
        return [_cast_tree(v, device) for v in o]
    return o


def pack_model_blob(w: Dict) -> torch.Tensor:
    """pack_state_dict(...) result -> the single uint8 blob pf_upsample_patches takes (layout: csrc/model.cu, enumerated
    through pf_model_entry so that C and Python cannot disagree).  CPU tensor; `.to(device)` it, or `.numpy().tofile(path)`
    for a host that is not Python."""
    from . import _native

    def piece(name):
        if name == "flow_tc":
            return w["flow_tc"]
        if name == "interp_tc":
            return w["interp"]["tc"]
        if name == "logdet_const":
            return torch.tensor([float(w["logdet_const"])], dtype=torch.float32)
        if name.startswith("chain"):
            return w["chain_tc"][int(name[5:])]
        blk, key = name.split(".")
        return w["feat"][int(blk[4:])][key]

    blob = torch.zeros(_native.lib.pf_model_bytes(), dtype=torch.uint8)
    for name, off, nbytes in _native.model_entries():
        t = piece(name).detach().cpu().contiguous()
        raw = t.view(torch.uint8).reshape(-1) if t.dtype != torch.uint8 else t.reshape(-1)
        if raw.numel() != nbytes:
            raise ValueError(f"model blob entry {name}: {raw.numel()} bytes packed, {nbytes} expected")
        blob[off:off + nbytes] = raw
    return blob
