"""Shared helpers of the test-suite (fixtures, oracle replays, tolerances)."""
import ctypes as C
import glob
import json
import os

import numpy as np

from disturbance_crazyfile_simulation_b200 import abi, hj_tables
from disturbance_crazyfile_simulation_b200.vec_env import _KWARGS

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
GOLDEN_NAMES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, '*.npz')))

# north_star tolerance: per-step state and obs within 1e-5 relative (fp32 kernel vs the float64 reference)
RTOL = 1e-5


def load_fixture(name):
    fix = dict(np.load(os.path.join(GOLDEN_DIR, name + '.npz')))
    fix['meta'] = json.loads(str(fix['meta']))
    return fix


def config_for(env_id, kwargs):
    cfg = abi.config_default(env_id)
    for k, v in kwargs.items():
        field = _KWARGS[k]
        ftype = dict(abi.CfConfig._fields_)[field]
        setattr(cfg, field, int(v) if ftype is C.c_int32 else float(v))
    return cfg


def fixture_tables(fix):
    """{level_idx: float32 table} used by a fixture (synthetic, regenerated from the seed)."""
    meta = fix['meta']
    if meta['table_seed'] is None:
        return {}
    cfg = config_for(meta['env_id'], meta['kwargs'])
    boltz = cfg.disturbance_kind == abi.CF_DSTB_HJ_BOLTZMANN
    levels = hj_tables.LEVELS if boltz else (cfg.disturbance_level,)
    return {hj_tables.level_index(l): hj_tables.synthetic_table(l, meta['table_seed']) for l in levels if l > 0}


def rel_err(a, b, floor=1.0):
    """max |a-b| / max(|b|, floor) - relative to the magnitude of the quantity, absolute below `floor`."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor))) if a.size else 0.0


# per-word scale floors of the state record for relative comparison (natural magnitudes)
def state_scale():
    F = abi.STATE_FIELDS
    s = np.ones(abi.CF_STATE_WORDS)
    for name, v in dict(pos=1.0, ep_ret=1.0, quat=1.0, vel=1.0, omega_w=1.0, lpf=1.0, gyro_bias=1e-2, motor_x=1.0,
                        ou=0.1, ring=1.0, act_prev1=1.0, act_prev2=1.0, obs_prev=1.0, dt=5e-3, mass=3e-2,
                        inertia=1e-5, ftf1=6e-3, motor_B=6e-2, motor_K=0.1).items():
        o, n = F[name]
        s[o:o + n] = v
    return s
