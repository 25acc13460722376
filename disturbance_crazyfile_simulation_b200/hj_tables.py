"""HJ value tables: names, grid constants and the synthetic stand-ins.

The reference loads ``fastrack_{level}_15x15.npy`` - a float32 ``(15,)*6`` value
function over (roll, pitch, yaw, p, q, r) - on every env step
(adversarial_generation/FasTrack_data/distur_gener.py:155).  The 27 real tables are
large blobs missing from the checkout (/root/reference/.MISSING_LARGE_BLOBS) and need
HeteroCL to regenerate, so tests and benchmarks use deterministic synthetic tables of
the same shape/dtype; real files are used when a directory holding them is given.
"""
from __future__ import annotations

import math
import os

import numpy as np

GRID_PTS = 15
GRID_DIMS = 6
TABLE_CELLS = GRID_PTS ** GRID_DIMS            # 11 390 625
# Boltzmann() support, envs/utils.py:27-39: np.around(np.arange(0.0, 2.1, 0.1), 1)
LEVELS = tuple(round(0.1 * i, 1) for i in range(21))
# distur_gener.py:152  umax;  dmax = level * umax  (distur_gener.py:160)
UMAX = (5.3e-3, 5.3e-3, 1.43e-4)


def level_index(level: float) -> int:
    """Index of a disturbance level in LEVELS (levels are multiples of 0.1)."""
    idx = int(round(float(level) * 10.0))
    if idx < 0 or abs(idx / 10.0 - float(level)) > 1e-9:
        raise ValueError(f'disturbance level {level} is not a non-negative multiple of 0.1')
    # distur_gener.py:154  assert disturbance <= 3.0
    if idx > 30:
        raise ValueError('disturbance level must be <= 3.0')
    return idx


def table_filename(level: float) -> str:
    """File name the reference formats at distur_gener.py:155."""
    return f'fastrack_{np.float64(level)}_15x15.npy'


def grid_points() -> np.ndarray:
    """The (6, 15) float64 grid coordinates exactly as odp Grid builds them
    (odp/Grid/GridProcessing.py:6-51 with the arguments of distur_gener.py:179):
    dims 0-2 periodic (upper bound shrunk by 1/15), dims 3-5 span [-pi, pi]."""
    lo = np.array([-math.pi / 2.4] * 3 + [-math.pi] * 3)
    hi = np.array([math.pi / 2.4] * 3 + [math.pi] * 3)
    pts = np.array([GRID_PTS] * 6)
    for dim in (0, 1, 2):
        hi[dim] = lo[dim] + (hi[dim] - lo[dim]) * (1 - 1 / pts[dim])
    return np.stack([np.linspace(lo[i], hi[i], num=GRID_PTS) for i in range(6)])


def synthetic_table(level: float, seed: int = 15) -> np.ndarray:
    """Deterministic float32 (15,)*6 stand-in: 50/50 blend of seeded white noise and the
    analytic FasTrack target  sum_{i != 2} x_i^2  (Fastrack.py:296-306), so the
    gradient has structure and sign changes are frequent."""
    li = level_index(level)
    rng = np.random.default_rng(seed * 1000 + li)
    noise = rng.standard_normal((GRID_PTS,) * 6, dtype=np.float32)
    g = grid_points()
    target = np.zeros((GRID_PTS,) * 6, dtype=np.float64)
    for dim in (0, 1, 3, 4, 5):
        shape = [1] * 6
        shape[dim] = GRID_PTS
        target = target + (g[dim] ** 2).reshape(shape)
    return (0.5 * noise + 0.5 * target.astype(np.float32)).astype(np.float32)


def tie_table(seed: int = 3) -> np.ndarray:
    """Adversarial table: few distinct small-integer values, so exact zeros, exact ties
    (zero gradients) and sign(V)=0 boundary extrapolations all occur."""
    rng = np.random.default_rng(seed)
    return rng.integers(-1, 2, size=(GRID_PTS,) * 6).astype(np.float32)


def _read_table(path: str) -> np.ndarray:
    V = np.load(path)
    if V.shape != (GRID_PTS,) * 6:
        raise ValueError(f'{path}: expected shape {(GRID_PTS,) * 6}, got {V.shape}')
    return np.ascontiguousarray(V, dtype=np.float32)


def default_cache_dir() -> str:
    return os.environ.get('CF_HJ_TABLE_CACHE') or os.path.join(os.path.expanduser('~'), '.cache',
                                                               'disturbance_crazyfile_simulation_b200', 'hj_tables')


def load_table(level: float, table_dir: str | None = None, seed: int | None = None, source: str | None = None,
               device: int = 0) -> np.ndarray:
    """The float32 ``(15,)*6`` value table of one disturbance level.

    ``table_dir`` given: ``fastrack_{level}_15x15.npy`` from there, FileNotFoundError when absent - what ``np.load`` at
    distur_gener.py:155 does.  Otherwise ``source`` (default: ``$CF_HJ_TABLES``, else ``'solve'``):
      'solve'      the real table, generated once with the GPU solver (``hj_solver.solve_fastrack``, the reference's
                   Fastrack.py problem) and cached under ``default_cache_dir()`` in the reference's file naming;
      'synthetic'  the seeded stand-in of ``synthetic_table`` - an explicit opt-in for tests and benchmarks, never a
                   silent fallback (a policy trained against it sees a random-sign torque, not the HJ-optimal one).
    A ``seed`` without a ``source`` selects 'synthetic' (the fixtures name their table by seed)."""
    if table_dir is not None:
        path = os.path.join(table_dir, table_filename(level))
        if not os.path.exists(path):
            raise FileNotFoundError(f'{path}: no HJ value table for disturbance level {level} (distur_gener.py:155); generate '
                                    f'the tables with tools/make_hj_tables.py or pass hj_tables="solve" / "synthetic"')
        return _read_table(path)
    if source is None:
        source = 'synthetic' if seed is not None else (os.environ.get('CF_HJ_TABLES') or 'solve')
    if source == 'synthetic':
        return synthetic_table(level, 15 if seed is None else seed)
    if source != 'solve':
        raise ValueError(f'hj_tables must be "solve", "synthetic" or a directory via hj_table_dir, got {source!r}')
    cache = default_cache_dir()
    path = os.path.join(cache, table_filename(level))
    if os.path.exists(path):
        return _read_table(path)
    import sys
    from . import hj_solver
    print(f'[cf_b200] generating the HJ value table of disturbance level {level} on GPU {device} (a few seconds, cached in '
          f'{cache})', file=sys.stderr)
    V = hj_solver.solve_fastrack(level, device=device)
    os.makedirs(cache, exist_ok=True)
    tmp = f'{path}.{os.getpid()}.tmp.npy'
    np.save(tmp, V)
    os.replace(tmp, path)
    return V
