"""hj_solver_oracle.py - CPU restatement (NumPy, float32) of the reference's FasTrack value-table generator.

TEST INFRASTRUCTURE ONLY (tests/, smoke, bench cpu legs) - never imported by the product package.

PARITY UNPINNED against the reference's own solver: `optimized_dp` builds its kernels with HeteroCL, which is not
installable here, and the 27 `fastrack_*_15x15.npy` outputs are missing from the checkout, so there
is no golden vector.  What is restated, with the reference's expression order and float32 arithmetic
(`hcl.config.init_dtype = hcl.Float(32)`, solver.py:93), paths under
phoenix_drone_simulation/adversarial_generation/:

* problem set-up       FasTrack_data/Fastrack.py:20-52 (UAV6D bounds, inertia), :281-320 (grid, target, tau,
                       TargetSetMode maxVWithV0, accuracy left at HJSolver's default "low")
* optimal u / d        Fastrack.py:54-128 (uMode "min", dMode "max"; a zero derivative leaves the hcl.scalar at 0)
* dynamics             Fastrack.py:133-167
* spatial derivatives  odp/spatialDerivatives/first_orderENO6D.py (all six dims alike: extrapolated ghost value
                       V + |dV| * sign(V) at both ends, no periodic wrap)
* one update           odp/computeGraphs/graph_6D.py:64-330 (Hamiltonian stage, global min/max of the one-sided
                       derivatives, Lax-Friedrichs dissipation with the 4 (u,d) corner combinations, step bound
                       0.8 / sum(max_alpha / dx), V + rate * dt, max with l0)
* time loop            odp/solver.py:152-206 (`while tNow <= tau[i] - 1e-4`)
* grid                 odp/Grid/GridProcessing.py:6-51 (periodic dims shrink max before dx is taken)

sin / cos of the grid values are tabulated once in float32 (`trig_tables`); the CUDA solver is handed the same
tables, so that it can be compared with this restatement bit for bit.
"""
from __future__ import annotations

import math

import numpy as np

F = np.float32
U_MAX = (5.3e-3, 5.3e-3, 1.43e-4)                     # Fastrack.py:22-23
IXX, IYY, IZZ = 1.65e-5, 1.65e-5, 2.92e-5             # Fastrack.py:48-50


def grid(pts: int = 15):
    """Grid(min, max, 6, [pts]*6, periodicDims=[0,1,2]) -> (points float64 [6][pts], dx float64 [6])."""
    lo = np.array([-math.pi / 2.4] * 3 + [-math.pi] * 3)
    hi = np.array([math.pi / 2.4] * 3 + [math.pi] * 3)
    for d in (0, 1, 2):
        hi[d] = lo[d] + (hi[d] - lo[d]) * (1 - 1 / pts)
    dx = (hi - lo) / (pts - 1.0)
    return np.stack([np.linspace(lo[d], hi[d], num=pts) for d in range(6)]), dx


def target(points: np.ndarray) -> np.ndarray:
    """FasTrackTarget(grid, ignore_dims=[2], center=0), Fastrack.py:296-306: float64 sum of squares -> float32."""
    pts = points.shape[1]
    data = np.zeros((pts,) * 6)
    for i in range(6):
        if i != 2:
            shape = [1] * 6
            shape[i] = pts
            data = data + np.power(points[i].reshape(shape) - 0.0, 2)
    return data.astype(F)


def trig_tables(points: np.ndarray):
    """float32 sin(phi), cos(phi), sin(theta)/cos(theta), cos(theta) of the grid values of dims 0 and 1."""
    phi, th = points[0].astype(F), points[1].astype(F)
    return np.sin(phi), np.cos(phi), (np.sin(th) / np.cos(th)).astype(F), np.cos(th)


def constants(level: float, dx: np.ndarray) -> dict:
    """Every scalar the update uses, rounded to float32 once (the same struct is handed to the CUDA solver)."""
    u_max = np.array(U_MAX)
    return dict(
        u_min=(-u_max).astype(F), u_max=u_max.astype(F),
        d_min=(level * -u_max).astype(F), d_max=(level * u_max).astype(F),        # Fastrack.py:42-44
        c=np.array([(IYY - IZZ) / IXX, (IZZ - IXX) / IYY, (IXX - IYY) / IZZ]).astype(F),
        inertia=np.array([IXX, IYY, IZZ]).astype(F), dx=dx.astype(F))


def one_sided(V: np.ndarray, axis: int, dx: np.float32):
    """first_orderENO6D.py: (left, right) derivative along `axis`."""
    V = np.moveaxis(V, axis, 0)
    L, R = np.empty_like(V), np.empty_like(V)
    d = V[1:] - V[:-1]
    L[1:] = d / dx
    R[:-1] = d / dx
    lb = V[0] + np.abs(V[1] - V[0]) * np.sign(V[0])
    L[0] = (V[0] - lb) / dx
    rb = V[-1] + np.abs(V[-1] - V[-2]) * np.sign(V[-1])
    R[-1] = (rb - V[-1]) / dx
    return np.moveaxis(L, 0, axis), np.moveaxis(R, 0, axis)


def _pick(deriv, neg_choice, pos_choice):
    """`if deriv > 0: x = pos_choice; elif deriv < 0: x = neg_choice` on an hcl.scalar initialised to 0."""
    return np.where(deriv > 0, pos_choice, np.where(deriv < 0, neg_choice, F(0))).astype(F)


def dynamics(x, trig, u, d, k):
    """UAV6D.dynamics, Fastrack.py:133-167; x = 6 broadcastable float32 arrays, trig as from trig_tables (shaped)."""
    sphi, cphi, tth, cth = trig
    p, q, r = x[3], x[4], x[5]
    f0 = (p + (r * cphi) * tth) + (q * sphi) * tth
    f1 = q * cphi - r * sphi
    f2 = (r * cphi) / cth + (q * sphi) / cth
    f3 = (k['c'][0] * r) * q + (u[0] + d[0]) / k['inertia'][0]
    f4 = (k['c'][1] * p) * r + (u[1] + d[1]) / k['inertia'][1]
    f5 = (k['c'][2] * p) * q + (u[2] + d[2]) / k['inertia'][2]
    return [f0, f1, f2, f3, f4, f5]


class Problem:
    """Grid, float32 state arrays (broadcast shapes), trig tables and constants of one FasTrack solve."""

    def __init__(self, level: float, pts: int = 15):
        self.level, self.pts = level, pts
        self.points, self.dx64 = grid(pts)
        self.k = constants(level, self.dx64)
        self.x = []
        for d in range(6):
            shape = [1] * 6
            shape[d] = pts
            self.x.append(self.points[d].astype(F).reshape(shape))
        s, c, t, ct = trig_tables(self.points)
        a0, a1 = [pts, 1, 1, 1, 1, 1], [1, pts, 1, 1, 1, 1]
        self.trig_flat = (s, c, t, ct)
        self.trig = (s.reshape(a0), c.reshape(a0), t.reshape(a1), ct.reshape(a1))
        self.l0 = target(self.points)


def step(P: Problem, V: np.ndarray, t0: np.float32, t1: np.float32):
    """One call of the compiled graph_6D kernel: returns (V_new, t0_new, info)."""
    k = P.k
    dd, dV, mn, mx = [], [], [], []
    for d in range(6):
        L, R = one_sided(V, d, k['dx'][d])
        dd.append(R - L)
        dV.append((L + R) / F(2))
        mn.append(min(F(1e9), L.min(), R.min()))
        mx.append(max(F(-1e9), L.max(), R.max()))
    u = [_pick(dV[3 + i], k['u_max'][i], k['u_min'][i]) for i in range(3)]          # uMode "min"
    dst = [_pick(dV[3 + i], k['d_max'][i], k['d_min'][i]) for i in range(3)]        # dMode "max"
    f = dynamics(P.x, P.trig, u, dst, k)
    Vn = -(((((f[0] * dV[0] + f[1] * dV[1]) + f[2] * dV[2]) + f[3] * dV[3]) + f[4] * dV[4]) + f[5] * dV[5])
    # dissipation: optimal u / d at the global minimum (L) and maximum (U) derivatives, graph_6D.py:185-300
    uL = [_pick(mn[3 + i], k['u_max'][i], k['u_min'][i]) for i in range(3)]
    uU = [_pick(mx[3 + i], k['u_max'][i], k['u_min'][i]) for i in range(3)]
    dL = [_pick(mn[3 + i], k['d_max'][i], k['d_min'][i]) for i in range(3)]
    dU = [_pick(mx[3 + i], k['d_max'][i], k['d_min'][i]) for i in range(3)]
    shape = (P.pts,) * 6
    alpha = None
    for uu, dd_ in ((uU, dL), (uL, dL), (uL, dU), (uU, dU)):
        fa = [np.broadcast_to(np.abs(v), shape) for v in dynamics(P.x, P.trig, uu, dd_, k)]
        alpha = fa if alpha is None else [np.maximum(a, b) for a, b in zip(alpha, fa)]
    diss = F(0.5) * (((((dd[0] * alpha[0] + dd[1] * alpha[1]) + dd[2] * alpha[2]) + dd[3] * alpha[3])
                      + dd[4] * alpha[4]) + dd[5] * alpha[5])
    Vn = -(Vn - diss)
    ma = [max(F(-1e9), a.max()) for a in alpha]
    dx = k['dx']
    sbi = ((((ma[0] / dx[0] + ma[1] / dx[1]) + ma[2] / dx[2]) + ma[3] / dx[3]) + ma[4] / dx[4]) + ma[5] / dx[5]
    sb = F(0.8) / sbi
    if sb > t1 - t0:
        sb = t1 - t0
    t0 = F(t0 + sb)
    Vnew = V + Vn * sb
    Vnew = np.where(Vnew < P.l0, P.l0, Vnew)                                        # maxVWithV0
    return Vnew.astype(F), t0, dict(min_deriv=np.array(mn), max_deriv=np.array(mx), max_alpha=np.array(ma), dt=sb)


def solve(level: float, pts: int = 15, lookback: float = 2.0, t_step: float = 1e-3, max_updates=None):
    """HJSolver(..., tau, {"TargetSetMode": "maxVWithV0"}) of Fastrack.py:308-317; returns (V, updates done, t)."""
    P = Problem(level, pts)
    V = P.l0.copy()
    tau = np.arange(start=0, stop=lookback + 1e-5, step=t_step)
    t_now, n = tau[0], 0
    for i in range(1, len(tau)):
        t = np.array((t_now, tau[i]), dtype=F)
        while t_now <= tau[i] - 1e-4:
            V, t[0], _ = step(P, V, t[0], t[1])
            t_now = t[0]
            n += 1
            if max_updates is not None and n >= max_updates:
                return V, n, float(t_now)
    return V, n, float(t_now)
