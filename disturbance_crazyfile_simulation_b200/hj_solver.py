"""FasTrack HJ value tables generated on the GPU.

The reference ships ``distur_gener`` but not the 27 ``fastrack_{level}_15x15.npy`` tables it loads; they come from
``adversarial_generation/FasTrack_data/Fastrack.py`` (``UAVSolution.get_fastrack``: 6-D grid, target
``sum_{i != 2} x_i^2``, 2 s look-back in 1 ms slices, ``maxVWithV0``) through optimized_dp's HeteroCL solver.  This
module sets that problem up (same grid rule, bounds and inertia) and runs it through the CUDA solver behind
``cf_hj_solver_*`` (csrc/cf_hj_solver.cu); ``solve_fastrack(level)`` returns the float32 ``(15,)*6`` table that
``CrazyFlieVecEnv.load_hj_table`` / ``cf_load_hj_table`` accept.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import abi

MAX_PTS = 16
U_MAX = (5.3e-3, 5.3e-3, 1.43e-4)                 # Fastrack.py:22-23
IXX, IYY, IZZ = 1.65e-5, 1.65e-5, 2.92e-5         # Fastrack.py:48-50


class HjProblem(C.Structure):
    """``cf_hj_problem`` of include/cf_abi.h."""
    _fields_ = [('pts', C.c_int32), ('reserved', C.c_int32),
                ('x', (C.c_float * MAX_PTS) * 6),
                ('sin_phi', C.c_float * MAX_PTS), ('cos_phi', C.c_float * MAX_PTS),
                ('tan_theta', C.c_float * MAX_PTS), ('cos_theta', C.c_float * MAX_PTS),
                ('target_sq', (C.c_double * MAX_PTS) * 6),
                ('dx', C.c_float * 6),
                ('u_min', C.c_float * 3), ('u_max', C.c_float * 3), ('d_min', C.c_float * 3), ('d_max', C.c_float * 3),
                ('c', C.c_float * 3), ('inertia', C.c_float * 3)]


def fastrack_problem(level: float, pts: int = 15, ignore_dims=(2,)) -> HjProblem:
    """Grid(..., periodicDims=[0,1,2]) (GridProcessing.py:6-51), UAV6D (Fastrack.py:20-52), FasTrackTarget (:296-306)."""
    assert 3 <= pts <= MAX_PTS
    lo = np.array([-math.pi / 2.4] * 3 + [-math.pi] * 3)
    hi = np.array([math.pi / 2.4] * 3 + [math.pi] * 3)
    for d in (0, 1, 2):
        hi[d] = lo[d] + (hi[d] - lo[d]) * (1 - 1 / pts)
    dx = (hi - lo) / (pts - 1.0)
    P = HjProblem()
    P.pts = pts
    f32 = np.float32
    for d in range(6):
        g = np.linspace(lo[d], hi[d], num=pts)
        for i in range(pts):
            P.x[d][i] = f32(g[i])
            P.target_sq[d][i] = 0.0 if d in ignore_dims else float(np.power(g[i] - 0.0, 2))
        P.dx[d] = f32(dx[d])
    phi = np.linspace(lo[0], hi[0], num=pts).astype(f32)
    th = np.linspace(lo[1], hi[1], num=pts).astype(f32)
    tan_th = (np.sin(th) / np.cos(th)).astype(f32)
    for i in range(pts):
        P.sin_phi[i], P.cos_phi[i] = np.sin(phi)[i], np.cos(phi)[i]
        P.tan_theta[i], P.cos_theta[i] = tan_th[i], np.cos(th)[i]
    u = np.array(U_MAX)
    for i in range(3):
        P.u_min[i], P.u_max[i] = f32(-u[i]), f32(u[i])
        P.d_min[i], P.d_max[i] = f32(level * -u[i]), f32(level * u[i])
    cs = ((IYY - IZZ) / IXX, (IZZ - IXX) / IYY, (IXX - IYY) / IZZ)
    for i in range(3):
        P.c[i], P.inertia[i] = f32(cs[i]), f32((IXX, IYY, IZZ)[i])
    return P


class HjSolver:
    """One solve on one GPU: ``update`` = one call of the reference's compiled kernel, ``run`` = its time loop."""

    def __init__(self, problem: HjProblem, device: int = 0):
        self._L = abi.lib()
        self.problem, self.pts = problem, int(problem.pts)
        self._h = C.c_void_p()
        abi.check(self._L.cf_hj_solver_create(C.byref(problem), device, C.byref(self._h)))

    def close(self) -> None:
        if self._h:
            self._L.cf_hj_solver_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def shape(self):
        return (self.pts,) * 6

    def get_value(self) -> np.ndarray:
        out = np.empty(self.shape, np.float32)
        abi.check(self._L.cf_hj_solver_get_value(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    def set_value(self, V: np.ndarray) -> None:
        V = np.ascontiguousarray(V, dtype=np.float32)
        assert V.shape == self.shape
        abi.check(self._L.cf_hj_solver_set_value(self._h, V.ctypes.data_as(C.c_void_p)))

    def update(self, t0: float, t1: float):
        t_new, info = C.c_float(), (C.c_float * 19)()
        abi.check(self._L.cf_hj_solver_update(self._h, C.c_float(t0), C.c_float(t1), C.byref(t_new), info))
        a = np.array(info[:], dtype=np.float32)
        return np.float32(t_new.value), dict(min_deriv=a[0:6], max_deriv=a[6:12], max_alpha=a[12:18], dt=a[18])

    def run(self, lookback: float = 2.0, t_step: float = 1e-3, max_updates: int = 0):
        done, t_final = C.c_int64(), C.c_double()
        abi.check(self._L.cf_hj_solver_run(self._h, lookback, t_step, int(max_updates or 0), C.byref(done), C.byref(t_final)))
        return done.value, t_final.value

    def timing(self):
        ms, n = C.c_double(), C.c_int64()
        abi.check(self._L.cf_hj_solver_timing(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value


def solve_fastrack(level: float, pts: int = 15, lookback: float = 2.0, t_step: float = 1e-3, device: int = 0,
                   max_updates: int = 0) -> np.ndarray:
    """``UAVSolution.get_fastrack`` for one disturbance level: the float32 ``(pts,)*6`` value table."""
    s = HjSolver(fastrack_problem(level, pts), device)
    try:
        s.run(lookback, t_step, max_updates)
        return s.get_value()
    finally:
        s.close()
