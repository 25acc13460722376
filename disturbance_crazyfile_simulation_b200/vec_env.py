"""Batched CrazyFlie hover environment: the Python host side above the C ABI.

Mirrors the gym interface of the reference envs (``reset`` / ``step`` / ``observation_space``
/ ``action_space`` / ``disturbance_level`` / ``_max_episode_steps`` / ``seed`` / ``close``;
envs/base.py:139-148, 420-507 and the step overrides of envs/hover_free.py, envs/hover.py)
for ``num_envs`` independent envs whose state lives in HBM.  PyTorch is used for device
memory and streams only; all arithmetic is in ``csrc/`` (hand-written sm_100a CUDA).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from . import abi
from .hj_tables import LEVELS, level_index, load_table

#: ids registered by phoenix_drone_simulation/__init__.py that this package implements
ENV_IDS = (
    'DroneHoverBulletFreeEnvWithoutAdversary-v0',
    'DroneHoverBulletFreeEnvWithAdversary-v0',
    'DroneHoverBulletFreeEnvWithRandomHJAdversary-v0',
    'DroneHoverBulletEnv-v0',
    'DroneHoverBulletEnvWithoutAdversary-v0',
    'DroneHoverBulletEnvWithAdversary-v0',
    'DroneHoverBulletEnvWithRandomHJAdversary-v0',
    'DroneHoverBulletEnvWithCurriculumHJAdversary-v0',
    'DroneHoverBulletEnvWithRandomAdversary-v0',
    'DroneHoverBulletEnvWithAdversaryInitial-v0',
)

# constructor kwargs of the reference env classes accepted by make() (base.py:26-47,
# hover_free.py:18-39) -> cf_config field
_KWARGS = {
    'observation_noise': 'observation_noise', 'domain_randomization': 'domain_randomization',
    'enable_reset_distribution': 'enable_reset_distribution', 'aggregate_phy_steps': 'aggregate_phy_steps',
    'latency': 'latency', 'motor_time_constant': 'motor_time_constant', 'motor_thrust_noise': 'motor_thrust_noise',
    'penalty_action': 'penalty_action', 'penalty_angle': 'penalty_angle', 'penalty_spin': 'penalty_spin',
    'penalty_terminal': 'penalty_terminal', 'penalty_velocity': 'penalty_velocity', 'penalty_z': 'penalty_z',
    'disturbance_level': 'disturbance_level', 'max_episode_steps': 'max_episode_steps',
    # switches of this implementation
    'use_ground_effect': 'use_ground_effect', 'gyroscopic': 'gyroscopic', 'hj_mode': 'hj_mode',
    'rng_mode': 'rng_mode', 'auto_reset': 'auto_reset', 'linear_damping': 'linear_damping',
    'angular_damping': 'angular_damping',
}


@dataclass
class Box:
    """Minimal stand-in for gym.spaces.Box (gym is not a dependency)."""
    low: np.ndarray
    high: np.ndarray
    shape: tuple
    dtype: type = np.float32

    def sample(self) -> np.ndarray:
        return np.random.uniform(self.low, self.high).astype(self.dtype)

    def contains(self, x) -> bool:
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


class CrazyFlieVecEnv:
    """``num_envs`` reference hover envs stepped by one CUDA kernel launch.

    ``reset() -> obs[N, obs_dim]``; ``step(act[N,4]) -> (obs, rew[N], done[N] bool, info)`` with
    device tensors.  Done envs are reset inside the step (``auto_reset``); ``info`` carries
    ``cost``, ``disturbance_level`` (lazy), ``truncated`` (``TimeLimit.truncated``) and
    ``terminal_observation`` for the envs that finished.
    """

    def __init__(self, env_id: str, num_envs: int = 1, device: Optional[int] = None, seed: int = 0,
                 env_id_base: int = 0, hj_table_dir: Optional[str] = None, hj_table_seed: Optional[int] = None,
                 hj_tables: Optional[str] = None, return_terminal_obs: bool = False, share_tables_with: Optional['CrazyFlieVecEnv'] = None, **kwargs):
        L = abi.lib()
        if not torch.cuda.is_available():
            raise RuntimeError('CrazyFlieVecEnv needs a CUDA device (no CPU fallback)')
        self.id = env_id
        self.num_envs = int(num_envs)
        self.device_index = torch.cuda.current_device() if device is None else int(device)
        self.device = torch.device('cuda', self.device_index)
        cfg = abi.config_default(env_id)
        for k, v in kwargs.items():
            if k not in _KWARGS:
                raise TypeError(f'unsupported env kwarg {k!r}')
            field = _KWARGS[k]
            ftype = dict(abi.CfConfig._fields_)[field]
            setattr(cfg, field, int(v) if ftype is C.c_int32 else float(v))
        self.cfg = cfg
        self._kwargs = dict(kwargs)                      # so that a validation env can be built with the same knobs
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            abi.check(L.cf_create(C.byref(cfg), self.num_envs, self.device_index, int(seed) & (2 ** 64 - 1),
                                  int(env_id_base), C.byref(self._h)))
        self._L = L
        self.obs_dim = L.cf_obs_dim(self._h)
        # spaces as in envs/base.py:139-148
        self.observation_space = Box(-1000.0 * np.ones(self.obs_dim, np.float32), 1000.0 * np.ones(self.obs_dim, np.float32),
                                     (self.obs_dim,))
        self.action_space = Box(-np.ones(4, np.float32), np.ones(4, np.float32), (4,))
        self._max_episode_steps = cfg.max_episode_steps
        n, od = self.num_envs, self.obs_dim
        self._obs = torch.empty((n, od), dtype=torch.float32, device=self.device)
        self._rew = torch.empty((n,), dtype=torch.float32, device=self.device)
        self._done = torch.empty((n,), dtype=torch.uint8, device=self.device)
        self._cost = torch.empty((n,), dtype=torch.float32, device=self.device)
        self._trunc = torch.empty((n,), dtype=torch.uint8, device=self.device)
        self._term = torch.zeros((n, od), dtype=torch.float32, device=self.device) if return_terminal_obs else None
        self._raster = None                          # frames: created on first use (capture_image / render)
        self._table_owner = share_tables_with        # keeps the owner alive
        if share_tables_with is not None:
            abi.check(L.cf_share_hj_tables(self._h, share_tables_with._h))
        else:
            self._load_tables(hj_table_dir, hj_table_seed, hj_tables)

    # ------------------------------------------------------------------ tables
    def _needed_levels(self):
        kind = self.cfg.disturbance_kind
        if kind == abi.CF_DSTB_HJ_FIXED:
            li = level_index(self.cfg.disturbance_level)
            return [li] if li > 0 else []
        if kind == abi.CF_DSTB_HJ_BOLTZMANN:
            return [level_index(l) for l in LEVELS if l > 0]
        return []

    def _load_tables(self, table_dir, seed, source):
        """Value tables of the levels this env can draw (hj_tables.load_table): files from ``hj_table_dir`` (missing file
        = FileNotFoundError, as np.load at distur_gener.py:155), else ``hj_tables='solve'`` (default: real tables from the
        GPU solver, cached) or the explicit opt-in ``hj_tables='synthetic'`` / ``hj_table_seed=...`` of tests and benchmarks."""
        for li in self._needed_levels():
            self.load_hj_table(li / 10.0, load_table(li / 10.0, table_dir, seed, source, self.device_index))

    def load_hj_table(self, level: float, V: np.ndarray) -> None:
        """np.load(fastrack_{level}_15x15.npy) of distur_gener.py:155, hoisted out of the step."""
        V = np.ascontiguousarray(V, dtype=np.float32)
        if V.shape != (15,) * 6:
            raise ValueError(f'value table must have shape (15,)*6, got {V.shape}')
        abi.check(self._L.cf_load_hj_table(self._h, level_index(level), V.ctypes.data_as(C.c_void_p)))

    # ------------------------------------------------------------------ gym surface
    def seed(self, seed=None):
        """gym 0.19 Env.seed is a no-op in the reference (iwpg.py:138); the RNG key is fixed at make()."""
        return

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def reset(self, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        m = None
        if mask is not None:
            m = mask.to(device=self.device, dtype=torch.uint8).contiguous()
        abi.check(self._L.cf_reset_dev(self._h, _ptr(m), _ptr(self._obs), self._stream()))
        return self._obs

    def step(self, action: torch.Tensor):
        act = action.to(device=self.device, dtype=torch.float32).contiguous()
        if act.shape != (self.num_envs, 4):
            raise ValueError(f'action must have shape ({self.num_envs}, 4), got {tuple(act.shape)}')
        abi.check(self._L.cf_step_dev(self._h, _ptr(act), _ptr(self._obs), _ptr(self._rew), _ptr(self._done),
                                      _ptr(self._cost), _ptr(self._term), _ptr(self._trunc), self._stream()))
        info = {'cost': self._cost, 'truncated': self._trunc.bool()}
        if self._term is not None:
            info['terminal_observation'] = self._term
        return self._obs, self._rew, self._done.bool(), info

    def step_into(self, action: torch.Tensor, obs: torch.Tensor, rew: torch.Tensor, done: torch.Tensor,
                  cost: Optional[torch.Tensor] = None, term_obs: Optional[torch.Tensor] = None,
                  truncated: Optional[torch.Tensor] = None) -> None:
        """``step`` writing straight into caller-owned tensors (rows of a rollout buffer): obs [N, obs_dim], rew [N] float32,
        done / truncated [N] uint8, cost [N]; ``term_obs`` [N, obs_dim] receives the last observation of the envs that finished.
        No copies, nothing returned; asynchronous on the current stream."""
        n, od = self.num_envs, self.obs_dim
        for t, shape, dt in ((action, (n, 4), torch.float32), (obs, (n, od), torch.float32), (rew, (n,), torch.float32),
                             (done, (n,), torch.uint8)):
            if t.shape != shape or t.dtype != dt or not t.is_contiguous() or t.device != self.device:
                raise ValueError(f'step_into: expected contiguous {dt} {shape} on {self.device}, got {t.dtype} {tuple(t.shape)}')
        abi.check(self._L.cf_step_dev(self._h, _ptr(action), _ptr(obs), _ptr(rew), _ptr(done), _ptr(cost), _ptr(term_obs),
                                      _ptr(truncated), self._stream()))

    def step_k(self, actions: torch.Tensor, obs: Optional[torch.Tensor] = None, rew: Optional[torch.Tensor] = None,
               done: Optional[torch.Tensor] = None, cost: Optional[torch.Tensor] = None):
        """K control steps in one launch (the tile's state stays on chip): actions [K, N, 4] ->
        obs [K, N, obs_dim], rew / done(uint8) / cost [K, N].  Open-loop / throughput path: there are no ``truncated`` /
        ``terminal_observation`` outputs, so with auto-reset a time-limit cut cannot be told from a true terminal - training
        loops that bootstrap from V(s_T) use :meth:`step` / :meth:`step_into`."""
        act = actions.to(device=self.device, dtype=torch.float32).contiguous()
        k = act.shape[0]
        if act.shape != (k, self.num_envs, 4):
            raise ValueError('actions must have shape (K, num_envs, 4)')
        n, od = self.num_envs, self.obs_dim
        obs = torch.empty((k, n, od), dtype=torch.float32, device=self.device) if obs is None else obs
        rew = torch.empty((k, n), dtype=torch.float32, device=self.device) if rew is None else rew
        done = torch.empty((k, n), dtype=torch.uint8, device=self.device) if done is None else done
        cost = torch.empty((k, n), dtype=torch.float32, device=self.device) if cost is None else cost
        abi.check(self._L.cf_step_k_dev(self._h, k, _ptr(act), _ptr(obs), _ptr(rew), _ptr(done), _ptr(cost), self._stream()))
        return obs, rew, done, cost

    @property
    def disturbance_level(self) -> np.ndarray:
        out = np.empty(self.num_envs, np.float32)
        abi.check(self._L.cf_get_levels(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    # ------------------------------------------------------------------ frames
    def set_visual_mesh(self, tris: np.ndarray, rgb: np.ndarray, bound_radius: Optional[float] = None) -> None:
        """Body-frame triangle soup the frames show (default: raster.proxy_mesh()); e.g. raster.mesh_with_stl_props(dir)."""
        from . import raster
        r = float(np.linalg.norm(np.asarray(tris).reshape(-1, 3), axis=1).max() * 1.001) if bound_radius is None else bound_radius
        if self._raster is not None:
            self._raster.close()
        self._raster = raster.Rasteriser(self.device_index, (tris, rgb, r))

    def _rasteriser(self):
        from . import raster
        if self._raster is None:
            self._raster = raster.Rasteriser(self.device_index)
        return self._raster

    def poses(self, indices: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Device tensor [F, 7]: position xyz + quaternion xyzw of the envs ``indices`` (int32 CUDA tensor; default all)."""
        dev = self.device
        if indices is None:
            idx, n = None, self.num_envs
        else:
            idx = indices.to(device=dev, dtype=torch.int32).contiguous()
            n = idx.numel()
        out = torch.empty((n, 7), dtype=torch.float32, device=dev)
        abi.check(self._L.cf_get_poses_dev(self._h, _ptr(idx), n, _ptr(out), self._stream()))
        return out

    def capture_images(self, indices: Optional[torch.Tensor] = None, camera=None) -> torch.Tensor:
        """uint8 CUDA tensor [F, 1080, 1920, 3]: one ``capture_image()`` frame per env in ``indices``, one launch."""
        with torch.cuda.device(self.device):
            return self._rasteriser().render(self.poses(indices), camera=camera)

    def capture_image(self, index: int = 0) -> np.ndarray:
        """``capture_image()`` of the reference (envs/base.py:520-544): [1080, 1920, 3] uint8, camera following the drone."""
        idx = torch.tensor([index], dtype=torch.int32, device=self.device)
        return self.capture_images(idx)[0].cpu().numpy()

    def render(self, mode='human', index: int = 0):
        """``render('rgb_array')`` (envs/base.py:379-416): the fixed camera at the origin.  'human' opens the PyBullet
        GUI in the reference; graphics is unavailable on this box."""
        if mode == 'rgb_array':
            from . import raster
            idx = torch.tensor([index], dtype=torch.int32, device=self.device)
            return self.capture_images(idx, camera=raster.render_camera())[0].cpu().numpy()
        if mode == 'human':
            raise NotImplementedError('render("human") is the PyBullet GUI (envs/base.py:364-377): unavailable on this box')
        raise NotImplementedError(mode)

    # ------------------------------------------------------------------ state / parity hooks
    def get_state(self) -> np.ndarray:
        out = np.empty((self.num_envs, abi.CF_STATE_WORDS), np.float32)
        abi.check(self._L.cf_get_state(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    def set_state(self, rec: np.ndarray) -> None:
        rec = np.ascontiguousarray(rec, dtype=np.float32)
        assert rec.shape == (self.num_envs, abi.CF_STATE_WORDS)
        abi.check(self._L.cf_set_state(self._h, rec.ctypes.data_as(C.c_void_p)))

    def set_noise_override(self, step_z=None, step_u=None, reset_z=None, reset_u=None) -> None:
        arrs = []
        for a, slots in ((step_z, abi.CF_STEP_NORMALS), (step_u, abi.CF_STEP_UNIFORMS),
                         (reset_z, abi.CF_RESET_NORMALS), (reset_u, abi.CF_RESET_UNIFORMS)):
            if a is None:
                arrs.append(None)
                continue
            a = np.ascontiguousarray(a, dtype=np.float32)
            assert a.shape == (slots, self.num_envs), (a.shape, slots)
            arrs.append(a)
        abi.check(self._L.cf_set_noise_override(
            self._h, *[None if a is None else a.ctypes.data_as(C.c_void_p) for a in arrs]))

    def hj_lookup(self, level: float, states: np.ndarray):
        """Grid indices and (+1/-1) signs of the optimal roll/pitch disturbance for float64 states [n, 6]."""
        s = np.ascontiguousarray(states, dtype=np.float64)
        n = s.shape[0]
        idx = np.empty((n, 6), np.int32)
        sign = np.empty((n, 2), np.float32)
        abi.check(self._L.cf_hj_lookup_host(self._h, level_index(level), s.ctypes.data_as(C.c_void_p), n,
                                            idx.ctypes.data_as(C.c_void_p), sign.ctypes.data_as(C.c_void_p)))
        return idx, sign

    def hj_lookup_f32(self, level: float, states: np.ndarray):
        """The same lookup for float32 states through the step kernel's own float32 threshold rule."""
        s = np.ascontiguousarray(states, dtype=np.float32)
        n = s.shape[0]
        idx = np.empty((n, 6), np.int32)
        sign = np.empty((n, 2), np.float32)
        abi.check(self._L.cf_hj_lookup_f32_host(self._h, level_index(level), s.ctypes.data_as(C.c_void_p), n,
                                                idx.ctypes.data_as(C.c_void_p), sign.ctypes.data_as(C.c_void_p)))
        return idx, sign

    def episode_stats(self, reset: bool = False) -> dict:
        out = (C.c_double * 4)()
        abi.check(self._L.cf_episode_stats(self._h, int(reset), out, None, None))
        n = out[0]
        return {'episodes': int(n), 'return_sum': out[1], 'length_sum': out[2],
                'mean_return': out[1] / n if n else float('nan'), 'mean_length': out[2] / n if n else float('nan')}

    def last_episodes(self):
        """(return [N], length [N], disturbance level [N]) of each env's last finished episode (length 0: none yet) - the
        columns of the reference's episodic_data.csv (utils/loggers.py:612-618)."""
        n = self.num_envs
        ret, ln, lv = np.empty(n, np.float32), np.empty(n, np.int32), np.empty(n, np.float32)
        abi.check(self._L.cf_last_episodes(self._h, ret.ctypes.data_as(C.c_void_p), ln.ctypes.data_as(C.c_void_p),
                                           lv.ctypes.data_as(C.c_void_p)))
        return ret, ln, lv

    def set_prefetch_next(self, nxt: Optional['CrazyFlieVecEnv'], next_act: Optional[torch.Tensor] = None) -> None:
        """Shard pipelining hint: while this env is stepped, prefetch the state of ``nxt`` (the env stepped next
        on this GPU) and its action rows into the L2.  Results never depend on it; ``None`` switches it off."""
        self._pf_keep = (nxt, next_act)              # keep both alive while the hint is set
        abi.check(self._L.cf_set_prefetch_next(self._h, None if nxt is None else nxt._h, _ptr(next_act)))

    def kernel_timing(self, enable: bool, reset: bool = False):
        ms, cnt = C.c_double(), C.c_int64()
        abi.check(self._L.cf_kernel_timing(self._h, int(enable), int(reset), C.byref(ms), C.byref(cnt)))
        return ms.value, cnt.value

    @property
    def launch_count(self) -> int:
        return int(self._L.cf_launch_count(self._h))

    # host-pointer API (NumPy in / NumPy out), what a gym-style caller uses
    def reset_host(self, obs: np.ndarray, mask: Optional[np.ndarray] = None) -> np.ndarray:
        abi.check(self._L.cf_reset_host(self._h, None if mask is None else mask.ctypes.data_as(C.c_void_p),
                                        obs.ctypes.data_as(C.c_void_p)))
        return obs

    def step_host(self, act, obs, rew, done, cost=None, term_obs=None, truncated=None):
        p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
        abi.check(self._L.cf_step_host(self._h, p(act), p(obs), p(rew), p(done), p(cost), p(term_obs), p(truncated)))

    def step_host_async(self, act, obs, rew, done, cost=None, term_obs=None, truncated=None):
        """step_host without the wait: pinned buffers stay untouched until :meth:`host_sync` returns."""
        p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
        abi.check(self._L.cf_step_host_async(self._h, p(act), p(obs), p(rew), p(done), p(cost), p(term_obs), p(truncated)))

    def host_sync(self):
        abi.check(self._L.cf_host_sync(self._h))

    def host_buffers(self, pinned: bool = True):
        """(act, obs, rew, done, cost) NumPy arrays for :meth:`step_host` / :meth:`step_host_async`, the four outputs ADJACENT in
        one block in the order the library downloads with a single copy (obs | rew | cost | done, include/cf_abi.h); ``pinned``
        takes the block from ``cf_host_alloc`` (true asynchronous DMA).  Keep the returned ``owner`` alive (it frees the pinned
        memory when collected): returns ``(arrays, owner)``."""
        n, od = self.num_envs, self.obs_dim
        nbytes = n * (4 * od + 9)

        class _Owner:
            def __init__(o, L, sizes):
                o._L, o.ptrs = L, []
                for sz in sizes:
                    q = L.cf_host_alloc(sz)
                    if not q:
                        raise MemoryError(L.cf_last_error().decode())
                    o.ptrs.append(q)

            def __del__(o):
                for q in o.ptrs:
                    o._L.cf_host_free(q)
                o.ptrs = []

        if pinned:
            owner = _Owner(self._L, (16 * n, nbytes))
            raw = lambda q, sz: np.ctypeslib.as_array((C.c_uint8 * sz).from_address(q))
            act, blk = raw(owner.ptrs[0], 16 * n).view(np.float32).reshape(n, 4), raw(owner.ptrs[1], nbytes)
        else:
            owner, act, blk = None, np.zeros((n, 4), np.float32), np.zeros(nbytes, np.uint8)
        obs = blk[:4 * n * od].view(np.float32).reshape(n, od)
        rew = blk[4 * n * od:4 * n * (od + 1)].view(np.float32)
        cost = blk[4 * n * (od + 1):4 * n * (od + 2)].view(np.float32)
        done = blk[4 * n * (od + 2):]
        return (act, obs, rew, done, cost), owner

    # ------------------------------------------------------------------ penalty terms (what the reference logs per step)
    def penalty_terms(self, action, physical_done, state: Optional[np.ndarray] = None) -> dict:
        """The terms ``compute_reward`` keeps for logging (hover_free.py:206-235: ``penalty_crash_log``, ``penalty_rpy_log``,
        ``penalty_rpy_dot_log``, ``penalty_z_log``, ``penalty_velocity_log``, ``penalty_log``; iwpg.py:531-532 writes them to
        penalty_data.csv per validation step), recomputed on the host from the post-step state record in float64.  The step
        kernel folds them into the reward; this is the logging path of small validation batches (one ``get_state`` per call).
        ``action`` [N, 4], ``physical_done`` [N] (done without the time-limit cut)."""
        F, cfg = abi.STATE_FIELDS, self.cfg
        st = (self.get_state() if state is None else state).astype(np.float64)
        g = lambda name: st[:, F[name][0]:F[name][0] + F[name][1]]
        pos, q, vel, ow = g('pos'), g('quat'), g('vel'), g('omega_w')
        x, y, z, w = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
        R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                      2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                      2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], axis=1).reshape(-1, 3, 3)
        wb = np.einsum('nji,nj->ni', R, ow)                 # R^T w (agents.py:452-453)
        sarg = np.clip(-2 * (x * z - w * y), -1, 1)
        rpy = np.stack([np.arctan2(2 * (y * z + w * x), w * w - x * x - y * y + z * z), np.arcsin(sarg),
                        np.arctan2(2 * (x * y + w * z), w * w + x * x - y * y - z * z)], axis=1)
        a = np.asarray(action, np.float64).reshape(-1, 4)
        out = dict(penalty_rpy=cfg.penalty_angle * np.linalg.norm(rpy, axis=1),
                   penalty_rpy_dot=cfg.penalty_spin * np.linalg.norm(wb, axis=1),
                   penalty_velocity=cfg.penalty_velocity * np.linalg.norm(vel, axis=1),
                   penalty_crash=cfg.penalty_terminal * np.asarray(physical_done, np.float64),
                   penalty_z=cfg.penalty_z * np.abs(pos[:, 2] - 1.0))
        p_action = cfg.penalty_action * np.linalg.norm(0.5 * (np.clip(a, -1, 1) + 1), axis=1)
        out['penalty'] = out['penalty_rpy'] + out['penalty_rpy_dot'] + out['penalty_velocity'] + p_action + out['penalty_crash']
        return out

    def close(self):
        if getattr(self, '_raster', None) is not None:
            self._raster.close()
            self._raster = None
        if getattr(self, '_h', None) is not None and self._h.value:
            self._L.cf_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class GymAdapter:
    """Single-env adapter with the reference's gym signature (NumPy in/out), so that the rollout
    loop of algs/iwpg/iwpg.py:372-410 runs unmodified: ``reset() -> ndarray[obs_dim]``,
    ``step(a) -> (obs, float, bool, dict)``; like gym's TimeLimit it does not auto-reset."""

    def __init__(self, env_id: str, **kwargs):
        kwargs = dict(kwargs)
        kwargs['auto_reset'] = 0
        self.vec = CrazyFlieVecEnv(env_id, num_envs=1, **kwargs)
        self.id = env_id
        self.observation_space, self.action_space = self.vec.observation_space, self.vec.action_space
        self._max_episode_steps = self.vec._max_episode_steps
        od = self.vec.obs_dim
        self._obs, self._act = np.zeros((1, od), np.float32), np.zeros((1, 4), np.float32)
        self._rew, self._done = np.zeros(1, np.float32), np.zeros(1, np.uint8)
        self._cost, self._trunc = np.zeros(1, np.float32), np.zeros(1, np.uint8)

    @property
    def unwrapped(self):
        return self

    @property
    def disturbance_level(self) -> float:
        return float(self.vec.disturbance_level[0])

    def seed(self, seed=None):
        return

    def reset(self) -> np.ndarray:
        self.vec.reset_host(self._obs)
        return self._obs[0].astype(np.float64)

    def step(self, action):
        self._act[0] = np.asarray(action, dtype=np.float32)
        self.vec.step_host(self._act, self._obs, self._rew, self._done, self._cost, None, self._trunc)
        info = {'cost': float(self._cost[0]), 'disturbance_level': self.disturbance_level}
        if self._trunc[0]:
            info['TimeLimit.truncated'] = True
        self._pen = None
        self._pen_args = (self._act.copy(), np.array([bool(self._done[0]) and not bool(self._trunc[0])]))
        return self._obs[0].astype(np.float64), float(self._rew[0]), bool(self._done[0]), info

    def _penalty(self, key: str) -> float:
        """penalty_*_log attributes of the reference env (hover_free.py:229-233), evaluated lazily after a step."""
        if getattr(self, '_pen', None) is None:
            self._pen = self.vec.penalty_terms(*self._pen_args)
        return float(self._pen[key][0])

    penalty_crash_log = property(lambda self: self._penalty('penalty_crash'))
    penalty_rpy_log = property(lambda self: self._penalty('penalty_rpy'))
    penalty_rpy_dot_log = property(lambda self: self._penalty('penalty_rpy_dot'))
    penalty_z_log = property(lambda self: self._penalty('penalty_z'))
    penalty_velocity_log = property(lambda self: self._penalty('penalty_velocity'))
    penalty_log = property(lambda self: self._penalty('penalty'))

    def render(self, mode='human'):
        return self.vec.render(mode)

    def capture_image(self) -> np.ndarray:
        return self.vec.capture_image(0)

    def close(self):
        self.vec.close()


def make(env_id: str, num_envs: Optional[int] = None, **kwargs):
    """``gym.make(id, **kwargs)`` of the reference (iwpg.py:76-77).  With ``num_envs`` it returns
    the batched :class:`CrazyFlieVecEnv`; without, the single-env :class:`GymAdapter`."""
    if env_id not in ENV_IDS:
        raise KeyError(f'unknown env id {env_id!r}; implemented: {ENV_IDS}')
    if num_envs is None:
        return GymAdapter(env_id, **kwargs)
    return CrazyFlieVecEnv(env_id, num_envs=num_envs, **kwargs)


def register_with_gym() -> int:
    """Register the ids with gym / gymnasium when one is importable (neither is a dependency)."""
    count = 0
    for modname in ('gym', 'gymnasium'):
        try:
            mod = __import__(modname)
            for env_id in ENV_IDS:
                mod.register(id=env_id, entry_point=lambda _id=env_id, **kw: GymAdapter(_id, **kw), max_episode_steps=500)
                count += 1
        except Exception:
            continue
    return count
