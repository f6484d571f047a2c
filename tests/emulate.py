"""Plain-torch (CPU, fp32) evaluation of the kernels' *factored* arithmetic on the PACKED weights.

Test infrastructure only.  It lets the CPU suite validate puflow_b200/packing.py (BN folding, the
P/Q split of the edge features, the composed interpolation layer, blob layouts) against the golden
fixtures of the unmodified reference without a GPU; the GPU suite then only has to show that each
CUDA kernel computes what this file computes.
"""
import math

import torch
import torch.nn.functional as F

from puflow_b200.packing import FEAT_CFG, NUM_BLOCKS


def linear(a, w, b=None, act=None, slope=0.0):
    y = a @ w
    if b is not None:
        y = y + b
    if act == "relu":
        y = F.relu(y)
    elif act == "lrelu":
        y = F.leaky_relu(y, slope)
    return y


def _rows(idx, N):
    Bp = idx.shape[0]
    return (torch.arange(Bp)[:, None, None] * N + idx.long()).reshape(Bp * N, -1)      # [M, K] global rows


def edgeconv(pq, idx, wd, wo, N, G, Co, slope=0.05, layers=4, pool=True):
    R = layers * G + Co
    P, Q = pq[:, :R], pq[:, R:]
    j = _rows(idx, N)
    Qj = Q[j]                                                  # [M, K, R]
    h = F.leaky_relu(P[:, None, :G] + Qj[:, :, :G], slope)
    off = 0
    for l in range(1, layers):
        w = wd[off:off + l * G * G].reshape(l * G, G)
        off += l * G * G
        t = h @ w + P[:, None, l * G:(l + 1) * G] + Qj[:, :, l * G:(l + 1) * G]
        h = torch.cat([h, F.leaky_relu(t, slope)], -1)
    o = h @ wo.reshape(layers * G, Co) + Qj[:, :, layers * G:]
    if pool:
        return o.max(1)[0] + P[:, layers * G:]
    return o + P[:, None, layers * G:]


def flow_cond(raw, cond):
    nb, M, _ = raw.shape
    sb = torch.zeros(nb, M, 8)
    for i in range(nb):
        c = cond[i]
        for net in range(2):
            h = F.leaky_relu(raw[i][:, 64 + 64 * net:128 + 64 * net], 0.01)
            w2 = c[4096 * net:4096 * (net + 1)].reshape(64, 64)
            b2 = c[8192 + 64 * net:8256 + 64 * net]
            h = F.leaky_relu(h @ w2 + b2, 0.01)
            w4 = c[8320 + 256 * net:8576 + 256 * net].reshape(64, 4)
            b4 = c[8832 + 4 * net:8836 + 4 * net]
            sb[i, :, 3 * net:3 * net + 3] = (h @ w4 + b4)[:, :3]
    return sb


def _coupling(fw, x, rawrow, tdim):
    w1h = fw[32:160].reshape(2, 64)
    a1 = F.leaky_relu(rawrow[:, :64] + x[:, :tdim] @ w1h[:tdim], 0.01)
    a2 = F.leaky_relu(a1 @ fw[384:4480].reshape(64, 64) + fw[160:224], 0.01)
    return (a2 @ fw[224:352].reshape(64, 2) + fw[352:354])[:, :3 - tdim]


def flow_fwd(x, raw, sb, flow):
    nb = raw.shape[0]
    ssum = torch.zeros(x.shape[0])
    for i in range(nb):
        fw = flow[i]
        tdim = int(fw[27])
        x = x * fw[0:3] + fw[3:6]
        x = x @ fw[9:18].reshape(3, 3).t()
        o = _coupling(fw, x, raw[i], tdim)
        x = torch.cat([x[:, :tdim], x[:, tdim:] - o], 1)
        x = x.flip(1) if fw[28] != 0 else x
        s, b = sb[i][:, 0:3], sb[i][:, 3:6]
        x = (x - b) * torch.exp(-s)
        ssum = ssum + s.sum(1)
    return x, ssum


def flow_inv(z, raw, sb, flow, R):
    nb = raw.shape[0]
    for i in reversed(range(nb)):
        fw = flow[i]
        tdim = int(fw[27])
        s = sb[i][:, 0:3].repeat_interleave(R, 0)
        b = sb[i][:, 3:6].repeat_interleave(R, 0)
        z = z * torch.exp(s) + b
        z = z.flip(1) if fw[28] != 0 else z
        o = _coupling(fw, z, raw[i].repeat_interleave(R, 0), tdim)
        z = torch.cat([z[:, :tdim], z[:, tdim:] + o], 1)
        z = z @ fw[18:27].reshape(3, 3).t()
        z = (z - fw[3:6]) * fw[6:9]
    return z


def interp(xyz, z, idx, pq, blob, R):
    Bp, N, _ = xyz.shape
    M = Bp * N
    j = _rows(idx[:, :, :8], N)                                # [M, 8]
    X = xyz.reshape(M, 3)
    P, Q = pq[:, :320], pq[:, 320:]
    Qj = Q[j]
    d = X[:, None, :] - X[j]
    dist = torch.sqrt((d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2])
    wdist = blob[0:64]
    d1 = F.leaky_relu(P[:, None, :64] + Qj[:, :, :64] + dist[..., None] * wdist, 0.01)
    d2 = F.leaky_relu(d1 @ blob[64:4160].reshape(64, 64) + blob[4160:4224], 0.01)
    h = F.leaky_relu(P[:, None, 64:80] + Qj[:, :, 64:80], 0.05)
    off = 4224
    for l in range(1, 8):
        w = blob[off:off + 256 * l].reshape(16 * l, 16)
        off += 256 * l
        t = h @ w + P[:, None, 64 + 16 * l:80 + 16 * l] + Qj[:, :, 64 + 16 * l:80 + 16 * l]
        h = torch.cat([h, F.leaky_relu(t, 0.05)], -1)
    m12 = blob[21728:46304].reshape(192, 128)
    t0 = F.leaky_relu(torch.cat([d2, h], -1) @ m12 + P[:, None, 192:] + Qj[:, :, 192:], 0.01)
    t1 = F.leaky_relu(t0 @ blob[11392:19584].reshape(128, 64) + blob[19584:19648], 0.01)
    lg = t1 @ blob[19648:21696].reshape(64, 32)[:, :R] + blob[21696:21696 + R]       # [M, 8, R]
    w = torch.softmax(lg, dim=1)
    zj = z.reshape(M, 3)[j]                                                          # [M, 8, 3]
    out = torch.einsum("mkr,mkc->mrc", w, zj)                                        # [M, R, 3]
    return out.reshape(Bp, N * R, 3), w


@torch.no_grad()
def forward(w, xyz, idx16, R=4):
    """Whole path with packed weights `w` (pack_state_dict(..., 'cpu')); returns the stage dict."""
    Bp, N, _ = xyz.shape
    M = Bp * N
    c = xyz.reshape(M, 3)
    feats, cs, raw = [], [], []
    for i in range(NUM_BLOCKS):
        _, g, Co = FEAT_CFG[i]
        f = w["feat"][i]
        c = edgeconv(linear(c, f["wpq"], f["bpq"]), idx16, f["wd"], f["wo"], N, g, Co)
        feats.append(c.view(Bp, N, -1))
        m = w["merge"][i]
        ci = linear(linear(c, m["w1"], m["b1"], "relu"), m["w2"])
        cs.append(ci.view(Bp, N, -1))
        raw.append(linear(ci, w["wc"][i]))
    raw = torch.stack(raw)
    sb = flow_cond(raw, w["cond"])
    z, ssum = flow_fwd(xyz.reshape(M, 3), raw, sb, w["flow"])
    logdet = w["logdet_const"] * N - ssum.view(Bp, N).sum(1)
    logp = (-0.5 * (z * z + math.log(2 * math.pi))).view(Bp, -1).sum(1)
    logp_x = -torch.mean(logp + logdet)
    wi = w["interp"]
    zup, wts = interp(xyz, z.view(Bp, N, 3), idx16, linear(xyz.reshape(M, 3), wi["wpq"], wi["bpq"]), wi["blob"], R)
    x = flow_inv(zup.reshape(-1, 3), raw, sb, w["flow"], R).view(Bp, N * R, 3)
    return dict(feats=feats, cs=cs, raw=raw, sb=sb, z=z.view(Bp, N, 3), logdet=logdet, logp_x=logp_x, zup=zup,
                interp_w=wts, x=x)
