"""CPU oracle of the continuous (CNF) variant's flow blocks -- TEST INFRASTRUCTURE, never imported by the product path
(puflow_b200/, modules/); only tests/, bench.py's cpu_baseline leg and __graft_entry__.smoke() may use it.

What it restates (reference file:line):
  * the dynamics dy/dt = ODEnet([t, c], y): modules/continuous/odefunc.py:66-111 (ODEnet with layer_type 'concatsquash',
    nonlinearity tanh), modules/continuous/diffeq_layers.py:72-86 (ConcatSquashLinear:
    layer(x) * sigmoid(hyper_gate([t, c])) + hyper_bias([t, c])), the conditional branch of ODEfunc.forward
    (odefunc.py:136-143: context = cat([t, c]));  PINNED: tests/golden/cnf_dynamics.npz holds outputs of the reference's own
    ODEnet class on the CNF checkpoint (oracle/ref/run_reference_cnf.py);
  * the block structure: modules/continuous/interpflow.py:19-50 (FlowBlock: batch_norm=False -> identity norms; CNF with
    T = sqrt_end_time ** 2, forward integrates [0, T], inverse [T, 0]: modules/continuous/cnf.py:71-86), :96-118 (f, g);
  * the integrator: the reference calls torchdiffeq.odeint(method='dopri5', atol=rtol=1e-5) (cnf.py:3-4,99-107).
    torchdiffeq is a third-party dependency that is NOT in /root/reference and not installed here (no pinned version in the
    repository: the reference has no requirements file), so the solver is restated from the published algorithm (Dormand &
    Prince 1980 coefficients; Hairer, Norsett, Wanner, "Solving ODEs I", II.4, for the starting step; torchdiffeq's documented
    controller: safety 0.9, growth <= 10, shrink >= 0.2, exponent 1/5) and PARITY OF THE SOLVER IS UNPINNED.  Two deliberate
    differences from torchdiffeq, stated in DESIGN.md: the error norm is taken per point (every point controls its own step;
    torchdiffeq uses one RMS norm over the whole batch, which makes a point's result depend on its batch-mates), and the last
    step is clipped to end at T (torchdiffeq steps past T and interpolates).  Both solve the same ODE to the same tolerance;
    `rk4_reference` (float64, fixed fine steps) is the solver-independent truth the tests compare both with.
  * not restated: the Hutchinson divergence estimate for log p (odefunc.py:8-32: random probe vector, autograd) -- the
    inference path (sample) never reads it.
"""
import numpy as np

A = [0.0, 1 / 5, 3 / 10, 4 / 5, 8 / 9, 1.0, 1.0]
B = [[],
     [1 / 5],
     [3 / 40, 9 / 40],
     [44 / 45, -56 / 15, 32 / 9],
     [19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729],
     [9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656],
     [35 / 384, 0.0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84]]
C_SOL = [35 / 384, 0.0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84, 0.0]
C_ERR = [35 / 384 - 1951 / 21600, 0.0, 500 / 1113 - 22642 / 50085, 125 / 192 - 451 / 720, -2187 / 6784 + 12231 / 42400,
         11 / 84 - 649 / 6300, -1 / 60]


def block_weights(sd, i, dtype=np.float64):
    """The twelve tensors of flow_blocks.i.cnf.odefunc.diffeq.layers.{0,1,2} + T, as numpy."""
    p = f"flow_blocks.{i}.cnf.odefunc.diffeq.layers."
    g = lambda k: sd[p + k].detach().cpu().numpy().astype(dtype)
    layers = [dict(W=g(f"{l}._layer.weight"), b=g(f"{l}._layer.bias"), Wb=g(f"{l}._hyper_bias.weight"),
                   Wg=g(f"{l}._hyper_gate.weight"), bg=g(f"{l}._hyper_gate.bias")) for l in range(3)]
    T = float(sd[f"flow_blocks.{i}.cnf.sqrt_end_time"]) ** 2
    return layers, T


def dynamics(layers, t, y, c):
    """ODEnet.forward on context [t, c] (odefunc.py:104-111, diffeq_layers.py:79-86).  y[M,3], c[M,C], t scalar or [M]."""
    t = np.broadcast_to(np.asarray(t, dtype=y.dtype), (y.shape[0],))
    ctx = np.concatenate([t[:, None], c], axis=1)
    dx = y
    for l, w in enumerate(layers):
        gate = 1.0 / (1.0 + np.exp(-(ctx @ w["Wg"].T + w["bg"])))
        dx = (dx @ w["W"].T + w["b"]) * gate + ctx @ w["Wb"].T
        if l < 2:
            dx = np.tanh(dx)
    return dx


def rk4_reference(layers, y0, c, t0, t1, steps=400):
    """Solver-independent reference: classical RK4, float64, `steps` equal steps from t0 to t1."""
    y, c = y0.astype(np.float64), c.astype(np.float64)
    h = (t1 - t0) / steps
    f = lambda t, y: dynamics(layers, t, y, c)
    for s in range(steps):
        t = t0 + s * h
        k1 = f(t, y); k2 = f(t + h / 2, y + h / 2 * k1); k3 = f(t + h / 2, y + h / 2 * k2); k4 = f(t + h, y + h * k3)
        y = y + h / 6 * (k1 + 2 * k2 + 2 * k3 + k4)
    return y


def _norm(x):
    return np.sqrt(np.mean(x * x, axis=1))          # RMS over a point's 3 coordinates


def dopri5(layers, y0, c, t0, t1, atol=1e-5, rtol=1e-5, dtype=np.float64, max_steps=10000):
    """Adaptive Dormand-Prince 5(4), one controller per point (see the module docstring).  Returns (y(t1), evaluations per point)."""
    y = y0.astype(dtype)
    c = c.astype(dtype)
    M = y.shape[0]
    sgn = 1.0 if t1 >= t0 else -1.0
    span = abs(t1 - t0)
    f = lambda s, yy, m: sgn * dynamics(layers, t0 + sgn * s, yy, c[m])      # in the forward variable s = |t - t0|
    allm = np.arange(M)
    k1 = f(np.zeros(M, dtype), y, allm)
    # starting step (Hairer II.4, order 4 as torchdiffeq passes it)
    scale = atol + np.abs(y) * rtol
    d0, d1 = _norm(y / scale), _norm(k1 / scale)
    h0 = np.where((d0 < 1e-5) | (d1 < 1e-5), 1e-6, 0.01 * d0 / np.maximum(d1, 1e-300))
    f1 = f(h0, y + h0[:, None] * k1, allm)
    d2 = _norm((f1 - k1) / scale) / h0
    h1 = np.where((d1 <= 1e-15) & (d2 <= 1e-15), np.maximum(1e-6, h0 * 1e-3), (0.01 / np.maximum(d1, d2)) ** 0.2)
    h = np.minimum(100 * h0, h1).astype(dtype)
    s = np.zeros(M, dtype)
    nfe = np.full(M, 2)
    done = np.zeros(M, bool)
    for _ in range(max_steps):
        act = np.where(~done)[0]
        if act.size == 0:
            break
        ha = np.minimum(h[act], span - s[act])
        ya, ka = y[act], [k1[act]]
        for st in range(1, 7):
            yi = ya + ha[:, None] * sum(B[st][j] * ka[j] for j in range(st))
            ka.append(f(s[act] + A[st] * ha, yi, act))
        nfe[act] += 6
        y1 = ya + ha[:, None] * sum(C_SOL[j] * ka[j] for j in range(7))
        err = ha[:, None] * sum(C_ERR[j] * ka[j] for j in range(7))
        ratio = _norm(err / (atol + rtol * np.maximum(np.abs(ya), np.abs(y1))))
        ok = ratio <= 1.0
        acc = act[ok]
        y[acc], s[acc], k1[acc] = y1[ok], s[acc] + ha[ok], ka[6][ok]          # FSAL: stage 7 is f(t1, y1)
        with np.errstate(divide="ignore"):
            fac = np.where(ratio == 0, 10.0, np.minimum(10.0, np.maximum(0.9 / ratio ** 0.2, np.where(ratio < 1, 1.0, 0.2))))
        h[act] = ha * fac
        done[act] = s[act] >= span * (1 - 1e-12)
    assert done.all(), "dopri5 did not finish"
    return y, nfe


class CNFOracle:
    """PointInterpFlow of modules/continuous/interpflow.py:53-139 on the conditioner of oracle/puflow_oracle.py."""

    def __init__(self, state_dict):
        from oracle.puflow_oracle import PUFlowOracle
        self.cond = PUFlowOracle(state_dict)
        self.blocks = [block_weights(state_dict, i) for i in range(6)]

    def flows(self, y, cs, reverse, upratio=1, solver="dopri5"):
        order = range(5, -1, -1) if reverse else range(6)
        for i in order:
            layers, T = self.blocks[i]
            c = np.repeat(cs[i], upratio, axis=0) if upratio > 1 else cs[i]
            t0, t1 = (T, 0.0) if reverse else (0.0, T)
            y = dopri5(layers, y, c, t0, t1)[0] if solver == "dopri5" else rk4_reference(layers, y, c, t0, t1)
        return y

    def forward(self, xyz, upratio=4, solver="dopri5"):
        """xyz torch[Bp,N,3] -> dict(z, zup, x) as numpy float64 (interpflow.py:121-130)."""
        import torch
        from oracle.puflow_oracle import knn
        Bp, N, _ = xyz.shape
        xn = np.ascontiguousarray(xyz.float().numpy())
        idx = torch.from_numpy(knn(xn, xn, self.cond.K)[1])
        with torch.no_grad():
            cs = [c.reshape(Bp * N, -1).double().numpy() for c in self.cond.feat_extract(xyz.float(), idx)]
            z = self.flows(xyz.reshape(-1, 3).double().numpy(), cs, reverse=False, solver=solver)
            zt = torch.from_numpy(z).to(xyz.dtype).view(Bp, N, 3)
            fz = self.cond.interp(zt, xyz.float(), upratio, idx[:, :, : self.cond.K_INTERP].contiguous())      # [Bp,N,3,R]
        zup = fz.permute(0, 1, 3, 2).reshape(Bp, N * upratio, 3)      # row n*R + r (interpflow.py:110: flatten(transpose(2,3),1,2))
        x = self.flows(zup.reshape(-1, 3).double().numpy(), cs, reverse=True, upratio=upratio, solver=solver)
        return dict(knn_idx=idx, cs=cs, z=z.reshape(Bp, N, 3), zup=zup.numpy(), x=x.reshape(Bp, N * upratio, 3))
