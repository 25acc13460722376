"""oracle-A, part 2: run the reference's OWN Python env stack on the fake Bullet client.

TEST INFRASTRUCTURE ONLY (see fake_bullet.py).  Usable only where ``/root/reference``
exists (this container); it generates the golden fixtures under ``tests/golden/`` via
``oracle/make_golden.py`` and validates the travelling C restatement
(``oracle/cf_oracle.c``).  Nothing here runs on the GPU box.

What is stubbed (absent, uninstallable packages) and how:
  gym            20-line real stub: Env, spaces.Box, envs.registration.register, make()
                 with gym-0.19 TimeLimit semantics (done=True once elapsed >= max steps).
  pybullet*      oracle/fake_bullet.py
  heterocl, matplotlib, plotly, imp, mpi4py, imageio   MagicMock (never executed on
                 the step path; distur_gener.py:1-15 imports them at module top)
Everything else - envs/base.py, hover_free.py, hover.py, agents.py, physics.py,
control.py, sensors.py, utils.py, distur_gener.py, odp/Grid/GridProcessing.py,
algs/core.py (Buffer / discount_cumsum), utils/online_mean_std.py - is the reference's
own code, imported from /root/reference and executed verbatim.
"""
from __future__ import annotations

import contextlib
import os
import sys
import tempfile
import types
from unittest import mock

import numpy as np

REF_ROOT = os.environ.get('CF_REFERENCE_ROOT', '/root/reference')
_HERE = os.path.dirname(os.path.abspath(__file__))


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, 'phoenix_drone_simulation'))


# ----------------------------------------------------------------------------- gym stub
def _build_gym_stub():
    gym = types.ModuleType('gym')
    spaces = types.ModuleType('gym.spaces')
    envs = types.ModuleType('gym.envs')
    registration = types.ModuleType('gym.envs.registration')

    class Env:
        metadata = {}

        def seed(self, seed=None):  # gym 0.19 Env.seed is a no-op returning None
            return

        @property
        def unwrapped(self):
            return self

    class Box:
        def __init__(self, low, high, shape=None, dtype=np.float32):
            self.low = np.asarray(low, dtype=dtype)
            self.high = np.asarray(high, dtype=dtype)
            self.shape = self.low.shape
            self.dtype = np.dtype(dtype)

        def sample(self):
            # gym uses a per-space RandomState; routed through np.random here so that the
            # harness recorder sees the draw (DroneHover*WithRandomAdversary only)
            return np.random.uniform(self.low, self.high).astype(self.dtype)

    class Discrete:
        def __init__(self, n):
            self.n = n

    class TimeLimit:
        def __init__(self, env, max_episode_steps):
            self.env = env
            self._max_episode_steps = max_episode_steps
            self._elapsed_steps = None

        def __getattr__(self, name):
            return getattr(self.env, name)

        @property
        def unwrapped(self):
            return self.env

        def step(self, action):
            obs, rew, done, info = self.env.step(action)
            self._elapsed_steps += 1
            if self._elapsed_steps >= self._max_episode_steps:
                info['TimeLimit.truncated'] = not done
                done = True
            return obs, rew, done, info

        def reset(self, **kwargs):
            self._elapsed_steps = 0
            return self.env.reset(**kwargs)

    registry = {}

    def register(id, entry_point, max_episode_steps=None, **kwargs):
        registry[id] = (entry_point, max_episode_steps, kwargs)

    def make(id, **kwargs):
        import importlib
        entry_point, max_steps, _ = registry[id]
        mod_name, cls_name = entry_point.split(':')
        cls = getattr(importlib.import_module(mod_name), cls_name)
        env = cls(**kwargs)
        return TimeLimit(env, max_steps) if max_steps else env

    spaces.Box, spaces.Discrete = Box, Discrete
    registration.register, registration.registry = register, registry
    envs.registration = registration
    gym.Env, gym.spaces, gym.envs, gym.make = Env, spaces, envs, make
    gym.wrappers = types.SimpleNamespace(TimeLimit=TimeLimit)
    return {'gym': gym, 'gym.spaces': spaces, 'gym.envs': envs,
            'gym.envs.registration': registration}


_installed = False


def install():
    """Put the stubs in sys.modules and the reference on sys.path (idempotent)."""
    global _installed
    if _installed:
        return
    if not reference_available():
        raise RuntimeError(f'reference checkout not found at {REF_ROOT}')
    if _HERE not in sys.path:
        sys.path.insert(0, _HERE)
    import fake_bullet

    sys.modules.update(_build_gym_stub())
    sys.modules['pybullet'] = fake_bullet
    pbd = types.ModuleType('pybullet_data')
    pbd.getDataPath = lambda: '/nonexistent/pybullet_data'
    sys.modules['pybullet_data'] = pbd
    pbu = types.ModuleType('pybullet_utils')
    pbu.bullet_client = fake_bullet
    sys.modules['pybullet_utils'] = pbu
    sys.modules['pybullet_utils.bullet_client'] = fake_bullet
    for name in ('heterocl', 'matplotlib', 'matplotlib.pyplot', 'plotly',
                 'plotly.graph_objects', 'imp', 'mpi4py', 'mpi4py.MPI', 'imageio'):
        sys.modules.setdefault(name, mock.MagicMock())
    sys.path.insert(0, REF_ROOT)
    sys.path.insert(0, os.path.join(REF_ROOT, 'phoenix_drone_simulation', 'adversarial_generation'))
    import phoenix_drone_simulation  # noqa: F401  (runs the 15 register() calls)
    _installed = True


TABLE_REL = 'phoenix_drone_simulation/adversarial_generation/FasTrack_data'


@contextlib.contextmanager
def table_dir(tables: dict):
    """chdir into a temp dir that holds ``fastrack_{level}_15x15.npy`` at the CWD-relative
    path distur_gener.py:155 loads from.  ``tables`` maps level (float) -> float32 (15,)*6."""
    old = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        d = os.path.join(tmp, TABLE_REL)
        os.makedirs(d)
        for level, V in tables.items():
            np.save(os.path.join(d, f'fastrack_{np.float64(level)}_15x15.npy'), np.asarray(V, np.float32))
        os.chdir(tmp)
        try:
            yield tmp
        finally:
            os.chdir(old)


class StdDrawRecorder:
    """Replaces the global np.random entry points the env stack uses (normal, uniform, randn,
    choice) by wrappers that draw the STANDARD variate from the same legacy MT19937 stream and
    apply NumPy's own affine formula (legacy_normal: loc + scale*gauss; legacy uniform:
    low + (high-low)*double; legacy choice(p=): cdf.searchsorted(random_sample(), 'right')).
    Results and stream order are bit-identical to the unwrapped calls (checked by
    tests/test_oracle_reference.py); the log holds (kind, standard draws) in call order."""

    def __init__(self):
        self.log = []
        self._orig = {}

    def _normal(self, loc=0.0, scale=1.0, size=None):
        z = np.random.standard_normal(size)
        self.log.append(('normal', np.array(z, dtype=np.float64).reshape(-1)))
        return loc + scale * z

    def _randn(self, *shape):
        z = np.random.standard_normal(shape if shape else None)
        self.log.append(('normal', np.array(z, dtype=np.float64).reshape(-1)))
        return z

    def _uniform(self, low=0.0, high=1.0, size=None):
        if size is None:
            size = np.broadcast(np.asarray(low), np.asarray(high)).shape or None
        u = np.random.random_sample(size)
        self.log.append(('uniform', np.array(u, dtype=np.float64).reshape(-1)))
        return low + (np.asarray(high) - np.asarray(low)) * u

    def _choice(self, a, size=None, replace=True, p=None):
        assert size is None and p is not None
        a = np.asarray(a)
        cdf = np.asarray(p, dtype=np.float64).cumsum()
        cdf /= cdf[-1]
        u = np.random.random_sample()
        self.log.append(('choice', np.array([u], dtype=np.float64)))
        return a[cdf.searchsorted(u, side='right')]

    def __enter__(self):
        for name in ('normal', 'uniform', 'randn', 'choice'):
            self._orig[name] = getattr(np.random, name)
            setattr(np.random, name, getattr(self, '_' + name))
        return self

    def __exit__(self, *exc):
        for name, orig in self._orig.items():
            setattr(np.random, name, orig)

    def take(self):
        out, self.log = self.log, []
        return out


def make_ref_env(env_id: str, **kwargs):
    install()
    import gym
    return gym.make(env_id, **kwargs)


def env_snapshot(env) -> dict:
    """Everything that determines the future of a reference env, as float64 arrays."""
    e = env.unwrapped
    d, bc = e.drone, e.bc
    ring = np.array(d.action_buffer, dtype=np.float64)
    snap = dict(
        pos=bc.pos.copy(), quat=bc.quat.copy(), vel=bc.vel.copy(), omega_world=bc.omega.copy(),
        motor_x=np.array(d.x, dtype=np.float64), ring=ring, ring_idx=np.array(d.action_idx),
        ou=np.array(d.thrust_noise.state, dtype=np.float64),
        gyro_bias=np.array(e.sensor_noise.gyro_bias, dtype=np.float64),
        lpf=np.broadcast_to(np.asarray(e.gyro_lpf._x, dtype=np.float64), (3,)).copy(),
        obs_hist=np.array([np.asarray(o, dtype=np.float64) for o in e.observation_history]),
        act_hist=np.array([np.asarray(a, dtype=np.float64) for a in e.action_history]),
        last_action=np.array(d.last_action, dtype=np.float64),
        rpy_dot=np.array(d.rpy_dot, dtype=np.float64), rpy=np.array(d.rpy, dtype=np.float64),
        dt=np.array(bc.dt), mass=np.array(bc.mass), inertia=bc.inertia.copy(),
        ftf1=np.array(d.force_torque_factor_1),
        motor_A=np.broadcast_to(np.asarray(d.A, dtype=np.float64), (4,)).copy(),
        motor_B=np.broadcast_to(np.asarray(d.B, dtype=np.float64), (4,)).copy(),
        motor_K=np.broadcast_to(np.asarray(d.K, dtype=np.float64), (4,)).copy(),
        level=np.array(float(e.disturbance_level)),
        iteration=np.array(e.iteration),
    )
    return snap
