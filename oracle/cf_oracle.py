"""ctypes front-end of oracle/libcf_oracle.so (oracle-B, the float64 C restatement).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU
baseline / reference arm - never by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
if _HERE not in sys.path:
    sys.path.insert(0, _HERE)

# Mirror of `cf_config` and the draw / record sizes of include/cf_abi.h.  Defined here (not imported from the product
# package) so that bench.py's reference arm runs without loading anything of the product; tests/test_abi_cpu.py checks
# that the two struct layouts are the same.
CF_STEP_NORMALS, CF_STEP_UNIFORMS, CF_RESET_NORMALS, CF_RESET_UNIFORMS = 44, 8, 48, 42
CF_STATE_WORDS = 76


class CfConfig(C.Structure):
    _fields_ = (
        [(n, C.c_int32) for n in (
            'abi_version', 'reward_kind', 'disturbance_kind', 'observation_noise',
            'enable_reset_distribution', 'aggregate_phy_steps', 'max_episode_steps', 'gyroscopic',
            'use_ground_effect', 'hj_mode', 'rng_mode', 'auto_reset')]
        + [(n, C.c_double) for n in (
            'domain_randomization', 'disturbance_level',
            'penalty_action', 'penalty_angle', 'penalty_spin', 'penalty_terminal', 'penalty_velocity', 'penalty_z',
            'done_roll_pitch_rad', 'done_rate_deg',
            'init_pos_lim', 'init_angle', 'init_yaw', 'init_vel_lim', 'init_rate_lim', 'init_yaw_rate_lim',
            'cost_rate_limit', 'sim_time_step', 'motor_time_constant', 'motor_thrust_noise', 'latency',
            'linear_damping', 'angular_damping')])

    def copy(self):
        out = CfConfig()
        C.memmove(C.byref(out), C.byref(self), C.sizeof(CfConfig))
        return out


def _cfg_copy(cfg):
    """a private copy of a cf_config given as this module's CfConfig or the product binding's (same layout)"""
    assert C.sizeof(cfg) == C.sizeof(CfConfig), 'cf_config layout mismatch'
    out = CfConfig()
    C.memmove(C.byref(out), C.byref(cfg), C.sizeof(CfConfig))
    return out


import build as _build  # noqa: E402  (oracle/build.py)

_lib = None
_P = C.c_void_p


def lib():
    global _lib
    if _lib is None:
        path = _build.build()
        L = C.CDLL(path)
        L.cfo_sizeof_env.restype = C.c_int
        L.cfo_num_threads.restype = C.c_int
        L.cfo_obs_dim.restype = C.c_int
        L.cfo_obs_dim.argtypes = [_P]
        L.cfo_init_batch.argtypes = [_P, C.c_int, _P, C.c_uint64, C.c_int64]
        L.cfo_reset_batch.argtypes = [_P, C.c_int, _P, _P, _P, _P, _P]
        L.cfo_step_batch.argtypes = [_P, C.c_int, _P] + [_P] * 11
        L.cfo_get_state_batch.argtypes = [_P, C.c_int, _P]
        L.cfo_set_state_batch.argtypes = [_P, C.c_int, _P]
        L.cfo_set_table.argtypes = [C.c_int, _P]
        L.cfo_hj_index.argtypes = [_P, _P]
        L.cfo_hj_index_batch.argtypes = [_P, C.c_int, _P]
        L.cfo_hj_signbits_batch.argtypes = [_P, _P, C.c_int, _P]
        L.cfo_hj_signs_all.argtypes = [_P, _P]
        L.cfo_hj_signbits.restype = C.c_int
        L.cfo_hj_signbits.argtypes = [_P, _P]
        L.cfo_grid_points.argtypes = [_P]
        L.cfo_boltzmann_cdf.argtypes = [_P]
        L.cfo_gae.argtypes = [_P, _P, _P, _P, _P, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, _P, _P, _P]
        L.cfo_set_motor_A.argtypes = [_P, _P]
        L.cfo_config_default.restype = C.c_int
        L.cfo_config_default.argtypes = [C.c_char_p, _P]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(_P)


def _f64(a, shape=None):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        assert a.shape == shape, (a.shape, shape)
    return a


_tables = {}   # keep the float32 arrays alive: the C side stores raw pointers


def set_table(level_idx: int, V: np.ndarray) -> None:
    V = np.ascontiguousarray(V, dtype=np.float32)
    assert V.shape == (15,) * 6
    _tables[level_idx] = V
    lib().cfo_set_table(level_idx, _p(V))


def config_default(env_id: str) -> CfConfig:
    """Constructor defaults of a registered env id (cfo_config_default), as the oracle's own CfConfig."""
    cfg = CfConfig()
    if lib().cfo_config_default(env_id.encode(), C.byref(cfg)) != 0:
        raise KeyError(env_id)
    return cfg


def synthetic_table(level: float, seed: int = 15) -> np.ndarray:
    """The float32 (15,)*6 stand-in value table of the tests and benchmarks (the real blobs are missing from the reference
    checkout): 50/50 blend of seeded white noise and the FasTrack target sum_{i != 2} x_i^2 (Fastrack.py:296-306).
    Same numbers as the product's hj_tables.synthetic_table (tests/test_abi_cpu.py checks it)."""
    li = int(round(float(level) * 10.0))
    rng = np.random.default_rng(seed * 1000 + li)
    noise = rng.standard_normal((15,) * 6, dtype=np.float32)
    g = grid_points()
    target = np.zeros((15,) * 6, dtype=np.float64)
    for dim in (0, 1, 3, 4, 5):
        shape = [1] * 6
        shape[dim] = 15
        target = target + (g[dim] ** 2).reshape(shape)
    return (0.5 * noise + 0.5 * target.astype(np.float32)).astype(np.float32)


def grid_points() -> np.ndarray:
    out = np.empty((6, 15), np.float64)
    lib().cfo_grid_points(_p(out))
    return out


def boltzmann_cdf() -> np.ndarray:
    out = np.empty(21, np.float64)
    lib().cfo_boltzmann_cdf(_p(out))
    return out


def hj_index(states: np.ndarray) -> np.ndarray:
    """Grid.get_index for float64 states [n, 6] -> int32 [n, 6]."""
    s = _f64(states)
    out = np.empty(s.shape, np.int32)
    lib().cfo_hj_index_batch(_p(s), s.shape[0], _p(out))
    return out


def hj_signbits(V: np.ndarray, idx: np.ndarray) -> np.ndarray:
    """bit0: dV/dx3 > 0, bit1: dV/dx4 > 0 at the given cells."""
    V = np.ascontiguousarray(V, dtype=np.float32)
    idx = np.ascontiguousarray(idx, dtype=np.int32)
    out = np.empty(idx.shape[0], np.int32)
    lib().cfo_hj_signbits_batch(_p(V), _p(idx), idx.shape[0], _p(out))
    return out


def hj_signs_all(V: np.ndarray) -> np.ndarray:
    """Packed 2-bit sign table of a whole value table (uint8, 15**6/4 + 1 bytes)."""
    V = np.ascontiguousarray(V, dtype=np.float32)
    out = np.empty((15 ** 6 + 3) // 4, np.uint8)
    lib().cfo_hj_signs_all(_p(V), _p(out))
    return out


class OracleEnvs:
    """n float64 reference-restatement envs (cfo_env array) behind the batch drivers."""

    def __init__(self, cfg: CfConfig, n: int, seed: int = 0, gid_base: int = 0):
        self.L = lib()
        self.cfg = _cfg_copy(cfg)
        self.n = n
        self._buf = np.zeros(n * self.L.cfo_sizeof_env(), np.uint8)
        self.L.cfo_init_batch(_p(self._buf), n, C.byref(self.cfg), seed & (2 ** 64 - 1), gid_base)
        self.obs_dim = self.L.cfo_obs_dim(C.byref(self.cfg))

    def reset(self, mask=None, nz=None, un=None) -> np.ndarray:
        obs = np.zeros((self.n, self.obs_dim), np.float64)
        m = None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)
        nz, un = _f64(nz, (CF_RESET_NORMALS, self.n)), _f64(un, (CF_RESET_UNIFORMS, self.n))
        self.L.cfo_reset_batch(_p(self._buf), self.n, C.byref(self.cfg), _p(m), _p(nz), _p(un), _p(obs))
        return obs

    def step(self, act, nz=None, un=None, rst_nz=None, rst_un=None):
        n, od = self.n, self.obs_dim
        act = _f64(act, (n, 4))
        nz, un = _f64(nz, (CF_STEP_NORMALS, n)), _f64(un, (CF_STEP_UNIFORMS, n))
        rst_nz, rst_un = _f64(rst_nz, (CF_RESET_NORMALS, n)), _f64(rst_un, (CF_RESET_UNIFORMS, n))
        obs, term = np.zeros((n, od)), np.zeros((n, od))
        rew, cost = np.zeros(n), np.zeros(n)
        done, trunc = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
        self.L.cfo_step_batch(_p(self._buf), n, C.byref(self.cfg), _p(act), _p(nz), _p(un), _p(rst_nz), _p(rst_un),
                              _p(obs), _p(rew), _p(done), _p(cost), _p(trunc), _p(term))
        return obs, rew, done, cost, trunc, term

    def get_state(self) -> np.ndarray:
        out = np.zeros((self.n, CF_STATE_WORDS), np.float64)
        self.L.cfo_get_state_batch(_p(self._buf), self.n, _p(out))
        return out

    def set_state(self, rec) -> None:
        rec = _f64(rec, (self.n, CF_STATE_WORDS))
        self.L.cfo_set_state_batch(_p(self._buf), self.n, _p(rec))

    def set_motor_A(self, i: int, A) -> None:
        A = _f64(A, (4,))
        sz = self.L.cfo_sizeof_env()
        self.L.cfo_set_motor_A(C.c_void_p(self._buf.ctypes.data + i * sz), _p(A))


def gae(rew, val, done, end_val, last_val, gamma, lam, rew_scale=0.0):
    """Buffer.finish_path over a time-major [T, N] rollout (algs/core.py:498-535)."""
    rew = np.ascontiguousarray(rew, np.float32)
    val = np.ascontiguousarray(val, np.float32)
    done = np.ascontiguousarray(done, np.uint8)
    end_val = None if end_val is None else np.ascontiguousarray(end_val, np.float32)
    last_val = np.ascontiguousarray(last_val, np.float32)
    T, N = rew.shape
    adv, tgt, ret = (np.empty((T, N), np.float32) for _ in range(3))
    lib().cfo_gae(_p(rew), _p(val), _p(done), _p(end_val), _p(last_val), gamma, lam, rew_scale, T, N,
                  _p(adv), _p(tgt), _p(ret))
    return adv, tgt, ret


def state_record_from_f32(rec32: np.ndarray) -> np.ndarray:
    """float32 host record of the CUDA library (integer words as raw bits) -> float64 oracle record."""
    rec32 = np.ascontiguousarray(rec32, np.float32)
    out = rec32.astype(np.float64)
    bits = rec32.view(np.uint32)
    for w in (57, 58, 59):
        out[:, w] = bits[:, w].astype(np.float64)
    return out


def state_record_to_f32(rec64: np.ndarray) -> np.ndarray:
    """float64 oracle record -> float32 record with integer words stored as raw bits."""
    rec64 = np.asarray(rec64, np.float64)
    out = rec64.astype(np.float32)
    bits = out.view(np.uint32)
    for w in (57, 58, 59):
        bits[:, w] = rec64[:, w].astype(np.uint64).astype(np.uint32)
    return out
