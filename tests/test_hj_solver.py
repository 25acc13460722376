"""HJ value-table generator.

CPU: the vectorised NumPy oracle against a scalar, cell-by-cell transliteration of the reference's loops
(graph_6D.py / first_orderENO6D.py) on a tiny grid, plus the problem struct handed to the CUDA solver.
GPU: the CUDA solver through the C ABI against the oracle - bit-exact (float32, same expression order).
Parity against optimized_dp itself is UNPINNED (HeteroCL cannot run here, no tables in the checkout).
"""
import ctypes as C

import numpy as np
import pytest

import hj_solver_oracle as H
from disturbance_crazyfile_simulation_b200 import hj_solver

F = np.float32


def _scalar_step(P, V, t0, t1):
    """One update written like the reference: nested loops over cells, hcl.scalar semantics, float32 scalars."""
    k, n = P.k, P.pts
    s, c, tth, cth = P.trig_flat
    xs = [P.x[d].ravel() for d in range(6)]

    def sign(v):
        return F(0) if v == 0 else (F(1) if v > 0 else F(-1))

    def my_abs(v):
        return v if v > 0 else -v

    def derivs(idx):
        out = []
        for d in range(6):
            i, dx = idx[d], k['dx'][d]
            lo, hi = list(idx), list(idx)
            lo[d] -= 1
            hi[d] += 1
            vc = V[tuple(idx)]
            if i == 0:
                vn = V[tuple(hi)]
                lb = vc + my_abs(vn - vc) * sign(vc)
                out.append(((vc - lb) / dx, (vn - vc) / dx))
            elif i == n - 1:
                vp = V[tuple(lo)]
                rb = vc + my_abs(vc - vp) * sign(vc)
                out.append(((vc - vp) / dx, (rb - vc) / dx))
            else:
                out.append(((vc - V[tuple(lo)]) / dx, (V[tuple(hi)] - vc) / dx))
        return out

    def pick(dv, neg, pos):
        return pos if dv > 0 else (neg if dv < 0 else F(0))

    def dyn(idx, u, d):
        p, q, r = xs[3][idx[3]], xs[4][idx[4]], xs[5][idx[5]]
        sp, cp, tt, ct = s[idx[0]], c[idx[0]], tth[idx[1]], cth[idx[1]]
        return [(p + (r * cp) * tt) + (q * sp) * tt, q * cp - r * sp, (r * cp) / ct + (q * sp) / ct,
                (k['c'][0] * r) * q + (u[0] + d[0]) / k['inertia'][0],
                (k['c'][1] * p) * r + (u[1] + d[1]) / k['inertia'][1],
                (k['c'][2] * p) * q + (u[2] + d[2]) / k['inertia'][2]]

    cells = list(np.ndindex(*V.shape))
    mn, mx = [F(1e9)] * 6, [F(-1e9)] * 6
    ham, dd = {}, {}
    for idx in cells:
        LR = derivs(idx)
        dV = [(L + R) / F(2) for L, R in LR]
        dd[idx] = [R - L for L, R in LR]
        u = [pick(dV[3 + i], k['u_max'][i], k['u_min'][i]) for i in range(3)]
        d = [pick(dV[3 + i], k['d_max'][i], k['d_min'][i]) for i in range(3)]
        f = dyn(idx, u, d)
        acc = f[0] * dV[0]
        for j in range(1, 6):
            acc = acc + f[j] * dV[j]
        ham[idx] = -acc
        for j, (L, R) in enumerate(LR):
            mn[j], mx[j] = min(mn[j], L, R), max(mx[j], L, R)
    uL = [pick(mn[3 + i], k['u_max'][i], k['u_min'][i]) for i in range(3)]
    uU = [pick(mx[3 + i], k['u_max'][i], k['u_min'][i]) for i in range(3)]
    dL = [pick(mn[3 + i], k['d_max'][i], k['d_min'][i]) for i in range(3)]
    dU = [pick(mx[3 + i], k['d_max'][i], k['d_min'][i]) for i in range(3)]
    ma, rate = [F(-1e9)] * 6, {}
    for idx in cells:
        a = [max(my_abs(p_), my_abs(q_)) for p_, q_ in zip(dyn(idx, uU, dL), dyn(idx, uL, dL))]
        a = [max(x_, my_abs(y_)) for x_, y_ in zip(a, dyn(idx, uL, dU))]
        a = [max(x_, my_abs(y_)) for x_, y_ in zip(a, dyn(idx, uU, dU))]
        acc = dd[idx][0] * a[0]
        for j in range(1, 6):
            acc = acc + dd[idx][j] * a[j]
        rate[idx] = -(ham[idx] - F(0.5) * acc)
        ma = [max(m_, x_) for m_, x_ in zip(ma, a)]
    sbi = ma[0] / k['dx'][0]
    for j in range(1, 6):
        sbi = sbi + ma[j] / k['dx'][j]
    sb = F(0.8) / sbi
    if sb > t1 - t0:
        sb = t1 - t0
    out = np.empty_like(V)
    for idx in cells:
        v = V[idx] + rate[idx] * sb
        out[idx] = P.l0[idx] if v < P.l0[idx] else v
    return out, F(t0 + sb)


def _rough_value(P, seed):
    """A value array with exact zeros, ties and sign changes, so every branch of the boundary rule runs."""
    rng = np.random.default_rng(seed)
    return (rng.integers(-3, 4, size=(P.pts,) * 6) * 0.5).astype(F)


def test_oracle_matches_scalar_transliteration():
    P = H.Problem(1.3, 4)
    for V in (P.l0.copy(), _rough_value(P, 1)):
        want, t_want = _scalar_step(P, V, F(0.0), F(1e-3))
        got, t_got, _ = H.step(P, V, F(0.0), F(1e-3))
        assert t_got == t_want
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_problem_struct_matches_oracle_constants():
    for level, pts in ((0.9, 15), (2.0, 7)):
        Q, O = hj_solver.fastrack_problem(level, pts), H.Problem(level, pts)
        assert C.sizeof(Q) == 1512 and Q.pts == pts
        for d in range(6):
            assert np.array_equal(np.array(Q.x[d][:pts], F), O.x[d].ravel())
        for got, want in ((Q.sin_phi, O.trig_flat[0]), (Q.cos_phi, O.trig_flat[1]), (Q.tan_theta, O.trig_flat[2]),
                          (Q.cos_theta, O.trig_flat[3])):
            assert np.array_equal(np.array(got[:pts], F), want)
        for name in ('dx', 'u_min', 'u_max', 'd_min', 'd_max', 'c', 'inertia'):
            assert np.array_equal(np.array(getattr(Q, name)[:], F), O.k[name]), name
        sq = np.array([Q.target_sq[d][:pts] for d in range(6)])
        l0 = (sq[0][:, None, None, None, None, None] + sq[1][None, :, None, None, None, None]
              + sq[3][None, None, None, :, None, None] + sq[4][None, None, None, None, :, None]
              + sq[5][None, None, None, None, None, :]) + np.zeros((pts,) * 6)
        assert np.array_equal(l0.astype(F), O.l0)


def test_oracle_time_loop_small():
    V, n, t = H.solve(1.5, pts=5, lookback=0.004)
    P = H.Problem(1.5, 5)
    assert n >= 4 and abs(t - 0.004) < 2e-4 and np.all(V >= P.l0) and np.isfinite(V).all()


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize('pts,level', [(6, 1.0), (7, 0.0), (15, 1.5)])
def test_gpu_update_bit_exact(pts, level):
    """One update from l0 and one from a rough array: value array, advanced time and the global reductions."""
    P = H.Problem(level, pts)
    s = hj_solver.HjSolver(hj_solver.fastrack_problem(level, pts))
    assert np.array_equal(s.get_value(), P.l0)
    for V in (P.l0.copy(), _rough_value(P, pts)):
        s.set_value(V)
        t_got, info = s.update(0.0, 1e-3)
        want, t_want, ref = H.step(P, V, F(0.0), F(1e-3))
        assert t_got == t_want and info['dt'] == ref['dt']
        for key in ('min_deriv', 'max_deriv', 'max_alpha'):
            assert np.array_equal(info[key], ref[key].astype(F)), key
        assert np.array_equal(s.get_value().view(np.uint32), want.view(np.uint32))
    s.close()


@pytest.mark.gpu
def test_gpu_time_loop_bit_exact():
    """HJSolver's loop over tau (solver.py:152-206): same number of updates, same final time, same bits."""
    for level, pts, lookback in ((1.5, 7, 0.02), (0.4, 6, 0.01)):
        want, n_want, t_want = H.solve(level, pts=pts, lookback=lookback)
        s = hj_solver.HjSolver(hj_solver.fastrack_problem(level, pts))
        n_got, t_got = s.run(lookback, 1e-3)
        assert n_got == n_want and t_got == t_want
        assert np.array_equal(s.get_value().view(np.uint32), want.view(np.uint32))
        ms, cnt = s.timing()
        assert cnt == n_got and ms > 0
        s.close()
    # capped run == the first updates of the uncapped one
    want, n_want, _ = H.solve(1.5, pts=7, lookback=0.02, max_updates=9)
    got = hj_solver.solve_fastrack(1.5, pts=7, lookback=0.02, max_updates=9)
    assert n_want == 9 and np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.gpu
def test_gpu_full_size_properties_and_env_load():
    """15^6 (the reference's size): 300 updates stay finite and above l0 (maxVWithV0), the table moves away from l0
    near the target, and the result loads into the env as the level's HJ table (cf_load_hj_table)."""
    from disturbance_crazyfile_simulation_b200 import make
    P = H.Problem(1.5, 15)
    V = hj_solver.solve_fastrack(1.5, max_updates=300)
    assert V.shape == (15,) * 6 and V.dtype == np.float32 and np.isfinite(V).all()
    assert np.all(V >= P.l0) and float((V - P.l0).max()) > 1e-3
    env = make('DroneHoverBulletFreeEnvWithAdversary-v0', num_envs=256, seed=1)
    env.load_hj_table(1.5, V)
    obs = env.reset()
    import torch
    act = torch.full((256, 4), 2.0 / 1.8 - 1.0, device='cuda')
    for _ in range(20):
        obs, rew, done, info = env.step(act)
    assert bool(torch.isfinite(obs).all())
    env.close()
