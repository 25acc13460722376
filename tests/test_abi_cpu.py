"""CPU checks of the drop-in boundary: the C-ABI library loads, exports every symbol that include/cf_abi.h
declares, and its host-only entry points (config defaults, error reporting) behave - no compute calls."""
import ctypes as C
import os
import re

import pytest

from disturbance_crazyfile_simulation_b200 import ENV_IDS, abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, 'include', 'cf_abi.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(cf_[a-z0-9_]+)\s*\(', src)))


def test_library_exports_every_declared_symbol():
    names = _declared()
    assert len(names) >= 30 and 'cf_step_dev' in names and 'cf_gae_dev' in names and 'cf_set_prefetch_next' in names
    L = abi.lib()
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert sorted(abi.EXPORTS) == names                     # the ctypes binding covers the whole header
    assert L.cf_abi_version() == abi.CF_ABI_VERSION


def test_config_defaults_follow_the_reference_constructors():
    for env_id in ENV_IDS:
        cfg = abi.config_default(env_id)
        assert cfg.aggregate_phy_steps == 2 and cfg.max_episode_steps == 500 and cfg.observation_noise == 1
        assert abs(cfg.domain_randomization - 0.1) < 1e-12 and abs(cfg.sim_time_step - 0.005) < 1e-15
        free = 'Free' in env_id
        assert cfg.reward_kind == (abi.CF_REWARD_FREE if free else abi.CF_REWARD_TARGET)
        assert cfg.penalty_terminal == 1000.0
    c = abi.config_default('DroneHoverBulletFreeEnvWithAdversary-v0')
    assert c.disturbance_kind == abi.CF_DSTB_HJ_FIXED and abs(c.disturbance_level - 1.5) < 1e-12   # hover_free.py:319
    c = abi.config_default('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0')
    assert c.disturbance_kind == abi.CF_DSTB_HJ_BOLTZMANN
    c = abi.config_default('DroneHoverBulletEnvWithAdversaryInitial-v0')
    assert abs(c.init_angle - 3.141592653589793 / 4) < 1e-12                                        # hover.py:1315-1317


def test_errors_are_codes_with_messages_not_exceptions():
    L = abi.lib()
    cfg = abi.CfConfig()
    assert L.cf_config_default(b'NoSuchEnv-v0', C.byref(cfg)) == abi.CF_ERR_INVALID
    assert b'NoSuchEnv-v0' in L.cf_last_error()
    with pytest.raises(abi.CfError):
        abi.config_default('NoSuchEnv-v0')
    assert L.cf_step_dev(None, None, None, None, None, None, None, None, None) == abi.CF_ERR_INVALID
    assert L.cf_gae_dev(None, None, None, None, None, 0.99, 0.95, 0.0, 1, 1, None, None, None, None) == abi.CF_ERR_INVALID
    cfg = abi.config_default('DroneHoverBulletFreeEnvWithoutAdversary-v0')
    cfg.aggregate_phy_steps = 3
    h = C.c_void_p()
    assert L.cf_create(C.byref(cfg), 4, 0, 0, 0, C.byref(h)) == abi.CF_ERR_INVALID and not h.value
