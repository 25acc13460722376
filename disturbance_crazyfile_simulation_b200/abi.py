"""ctypes binding of the C ABI in ``include/cf_abi.h`` (``csrc/libcf_b200.so``).

The library is the product: there is no Python/CPU fallback.  If it has not been built
(``python -m disturbance_crazyfile_simulation_b200.build`` or ``__graft_entry__.build()``)
importing :func:`lib` raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('CF_LIB_PATH') or os.path.join(HERE, 'csrc', 'libcf_b200.so')   # override: experimental builds (tools/)

CF_ABI_VERSION = 1
CF_OK, CF_ERR_INVALID, CF_ERR_CUDA, CF_ERR_NO_TABLE, CF_ERR_ALLOC = 0, -1, -2, -3, -4
CF_DSTB_NONE, CF_DSTB_HJ_FIXED, CF_DSTB_HJ_BOLTZMANN, CF_DSTB_RANDOM_BOX = 0, 1, 2, 3
CF_REWARD_FREE, CF_REWARD_TARGET = 0, 1
CF_RNG_PHILOX, CF_RNG_INJECTED = 0, 1
CF_HJ_VALUE_TABLE, CF_HJ_SIGN_TABLE = 0, 1
CF_HJ_CELLS = 15 ** 6
CF_MAX_LEVELS = 31
CF_STEP_NORMALS, CF_STEP_UNIFORMS, CF_RESET_NORMALS, CF_RESET_UNIFORMS = 44, 8, 48, 42
CF_STATE_WORDS = 76

# word offsets of the host-visible state record (enum in cf_abi.h)
STATE_FIELDS = dict(
    pos=(0, 3), ep_ret=(3, 1), quat=(4, 4), vel=(8, 3), omega_w=(11, 3), lpf=(14, 3), gyro_bias=(17, 3),
    motor_x=(20, 4), ou=(24, 4), ring=(28, 8), act_prev1=(36, 4), act_prev2=(40, 4), obs_prev=(44, 13),
    ep_len=(57, 1), rng_ctr=(58, 1), misc=(59, 1), dt=(60, 1), mass=(61, 1), inertia=(62, 3), ftf1=(65, 1),
    motor_B=(66, 4), motor_K=(70, 4))
INT_WORDS = (57, 58, 59)


class CfConfig(C.Structure):
    """``cf_config`` of cf_abi.h: the reference env constructors' kwargs, flattened."""
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

    def as_dict(self) -> dict:
        return {n: getattr(self, n) for n, _ in self._fields_}

    def copy(self) -> 'CfConfig':
        out = CfConfig()
        C.memmove(C.byref(out), C.byref(self), C.sizeof(CfConfig))
        return out


class CfError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f'cf_abi error {code}: {msg}')
        self.code = code


_P = C.c_void_p


class CfPolicyWeights(C.Structure):
    """``cf_policy_weights`` of include/cf_abi.h (device pointers)."""
    _fields_ = [('obs_mean', C.c_void_p), ('obs_std', C.c_void_p),
                ('pi_w', C.c_void_p * 3), ('pi_b', C.c_void_p * 3), ('v_w', C.c_void_p * 3), ('v_b', C.c_void_p * 3),
                ('log_std', C.c_void_p), ('pi_hidden', C.c_int32), ('v_hidden', C.c_int32), ('obs_eps', C.c_float),
                ('reserved', C.c_int32)]
_SIGS = {
    'cf_abi_version': (C.c_int, []),
    'cf_last_error': (C.c_char_p, []),
    'cf_config_default': (C.c_int, [C.c_char_p, C.POINTER(CfConfig)]),
    'cf_create': (C.c_int, [C.POINTER(CfConfig), C.c_int32, C.c_int32, C.c_uint64, C.c_int64, C.POINTER(_P)]),
    'cf_destroy': (C.c_int, [_P]),
    'cf_obs_dim': (C.c_int, [_P]),
    'cf_num_envs': (C.c_int, [_P]),
    'cf_load_hj_table': (C.c_int, [_P, C.c_int32, _P]),
    'cf_share_hj_tables': (C.c_int, [_P, _P]),
    'cf_get_hj_signs': (C.c_int, [_P, C.c_int32, _P]),
    'cf_reset_dev': (C.c_int, [_P, _P, _P, _P]),
    'cf_step_dev': (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _P]),
    'cf_step_k_dev': (C.c_int, [_P, C.c_int32, _P, _P, _P, _P, _P, _P]),
    'cf_reset_host': (C.c_int, [_P, _P, _P]),
    'cf_step_host': (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _P]),
    'cf_step_host_async': (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _P]),
    'cf_host_sync': (C.c_int, [_P]),
    'cf_get_state': (C.c_int, [_P, _P]),
    'cf_set_state': (C.c_int, [_P, _P]),
    'cf_get_levels': (C.c_int, [_P, _P]),
    'cf_set_noise_override': (C.c_int, [_P, _P, _P, _P, _P]),
    'cf_hj_lookup_host': (C.c_int, [_P, C.c_int32, _P, C.c_int32, _P, _P]),
    'cf_hj_lookup_f32_host': (C.c_int, [_P, C.c_int32, _P, C.c_int32, _P, _P]),
    'cf_episode_stats': (C.c_int, [_P, C.c_int32, _P, _P, _P]),
    'cf_last_episodes': (C.c_int, [_P, _P, _P, _P]),
    'cf_gae_dev': (C.c_int, [_P, _P, _P, _P, _P, C.c_float, C.c_float, C.c_float, C.c_int32, C.c_int32, _P, _P, _P, _P]),
    'cf_gae_host': (C.c_int, [_P, _P, _P, _P, _P, C.c_float, C.c_float, C.c_float, C.c_int32, C.c_int32, _P, _P, _P]),
    'cf_set_prefetch_next': (C.c_int, [_P, _P, _P]),
    'cf_set_trace_buffer': (C.c_int, [_P, _P]),
    'cf_num_tiles': (C.c_int, [_P]),
    'cf_hj_solver_create': (C.c_int, [_P, C.c_int32, _P]),
    'cf_hj_solver_destroy': (C.c_int, [_P]),
    'cf_hj_solver_set_value': (C.c_int, [_P, _P]),
    'cf_hj_solver_get_value': (C.c_int, [_P, _P]),
    'cf_hj_solver_update': (C.c_int, [_P, C.c_float, C.c_float, _P, _P]),
    'cf_hj_solver_run': (C.c_int, [_P, C.c_double, C.c_double, C.c_int64, _P, _P]),
    'cf_hj_solver_timing': (C.c_int, [_P, _P, _P]),
    'cf_policy_step_dev': (C.c_int, [_P, _P, _P, C.c_int32, _P, _P, _P, _P]),
    'cf_ppo_workspace_floats': (C.c_int, []),
    'cf_ppo_pi_grad_dev': (C.c_int, [_P, C.c_int32, _P, _P, _P, _P, C.c_int64, C.c_float, _P, _P, _P, _P, _P]),
    'cf_ppo_v_grad_dev': (C.c_int, [_P, C.c_int32, _P, _P, _P, C.c_int64, _P, _P, _P, _P, _P]),
    'cf_ppo_static_max_samples': (C.c_int64, []),
    'cf_ppo_pi_grad_static_dev': (C.c_int, [_P, C.c_int32, _P, _P, _P, _P, C.c_int64, C.c_float, _P, _P, _P, _P]),
    'cf_pd_actions_dev': (C.c_int, [_P, C.c_int32, C.c_int32, _P, _P, _P]),
    'cf_raster_create': (C.c_int, [_P, C.c_int32, _P, _P, C.c_int32, C.c_float]),
    'cf_raster_destroy': (None, [_P]),
    'cf_raster_render_dev': (C.c_int, [_P, _P, _P, C.c_int32, _P, _P]),
    'cf_raster_render_host': (C.c_int, [_P, _P, _P, C.c_int32, _P]),
    'cf_get_poses_dev': (C.c_int, [_P, _P, C.c_int32, _P, _P]),
    'cf_debug_draws': (C.c_int, [_P, C.c_int32, C.c_int32, C.c_uint32, C.c_uint32, C.c_int32, C.c_int32, _P, _P]),
    'cf_debug_box_muller': (C.c_int, [C.c_int32, _P, C.c_int64, _P, _P]),
    'cf_debug_euler': (C.c_int, [C.c_int32, _P, C.c_int64, _P, _P]),
    'cf_launch_count': (C.c_int64, [_P]),
    'cf_kernel_timing': (C.c_int, [_P, C.c_int32, C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    'cf_host_alloc': (_P, [C.c_uint64]),
    'cf_host_free': (None, [_P]),
}
EXPORTS = tuple(_SIGS)

_lib = None


def lib() -> C.CDLL:
    """The loaded C-ABI library.  Raises when it is missing: there is no fallback path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f'{LIB_PATH} not found - build the CUDA extension first '
                f'(python -m disturbance_crazyfile_simulation_b200.build); there is no CPU fallback')
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        if L.cf_abi_version() != CF_ABI_VERSION:
            raise ImportError('libcf_b200.so was built from a different cf_abi.h; rebuild it')
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != CF_OK:
        raise CfError(rc, lib().cf_last_error().decode())


def config_default(env_id: str) -> CfConfig:
    cfg = CfConfig()
    check(lib().cf_config_default(env_id.encode(), C.byref(cfg)))
    return cfg
