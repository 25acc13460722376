"""GPU: the PRODUCTION RNG path (rng_mode = CF_RNG_PHILOX - what every benchmark number runs) against oracle-B.

Philox4x32-10 is integer work: the raw 32-bit words and the u01 bits must be IDENTICAL to the oracle's; the Box-Muller
normals (MUFU lg2 / sqrt / sin / cos on the GPU, libm float64 in the oracle) agree to a stated absolute bound, checked
over every one of the 2^23 u01 values.  Then the whole step: with both sides on the same stream, 60 control steps with
in-place resets agree step by step (oracle re-synchronised to the GPU state after every step, so a single misplaced draw
would show as an O(1) difference), and a free-running copy stays within the measured drift bound.
"""
import ctypes as C

import numpy as np
import pytest

import cf_oracle as co
from disturbance_crazyfile_simulation_b200 import abi, hj_tables
from disturbance_crazyfile_simulation_b200.vec_env import make
from helpers import config_for, rel_err

pytestmark = pytest.mark.gpu

# measured on B200 (profiles/r2_rng_bits.json): worst |z_gpu - z_f64| over all 2^23 radius words x 64 angles and all 2^23
# angle words x 64 radii
NORMAL_ABS_BOUND = 4e-6


def _debug_draws(env, first, count, ctr, tag, nzb, nub):
    words = np.zeros((count, nzb + nub, 4), np.uint32)
    draws = np.zeros((count, nzb + nub, 4), np.float32)
    abi.check(env._L.cf_debug_draws(env._h, first, count, ctr, tag, nzb, nub, words.ctypes.data_as(C.c_void_p),
                                    draws.ctypes.data_as(C.c_void_p)))
    return words, draws


def _oracle_words(seed, gid, ctr, tag, block):
    key = (C.c_uint32 * 2)(seed & 0xffffffff, (seed >> 32) & 0xffffffff)
    c = (C.c_uint32 * 4)(gid & 0xffffffff, ctr, block, (tag ^ (((gid >> 32) & 0xffffffff) << 8)) & 0xffffffff)
    out = (C.c_uint32 * 4)()
    co.lib().cfo_philox(c, key, out)
    return np.array(out[:], dtype=np.uint32)


def _u01_f64(w):
    return ((w >> 9).astype(np.float64) * 2.0 + 1.0) / 16777216.0


def _box_muller_f64(w):
    """[..., 4] uint32 -> [..., 4] float64 normals, the oracle's formula (oracle/cf_oracle.c draw_event)."""
    u = _u01_f64(w)
    r0, r1 = np.sqrt(-2.0 * np.log(u[..., 0])), np.sqrt(-2.0 * np.log(u[..., 2]))
    t0, t1 = np.pi * (2.0 * u[..., 1] - 1.0), np.pi * (2.0 * u[..., 3] - 1.0)
    return np.stack([r0 * np.cos(t0), r0 * np.sin(t0), r1 * np.cos(t1), r1 * np.sin(t1)], axis=-1)


@pytest.mark.parametrize('seed,base', [(123, 1000), (0xDEADBEEFCAFE1234, (7 << 32) + 4294967290)])
def test_philox_words_and_u01_bits_identical_to_oracle(seed, base):
    """All blocks of a step event (11 normal + 2 uniform), of both reset events (12 + 11) and of the constructor draw,
    for env ids that straddle a 2^32 boundary (the high id word goes into counter word 3) and a 64-bit seed."""
    n = 16
    env = make('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', num_envs=n, seed=seed, env_id_base=base)
    for tag, nzb, nub in ((0, 11, 2), (1, 12, 11), (2, 12, 11), (3, 1, 0)):
        for ctr in (0, 1, 77, 2 ** 31 - 1):
            words, draws = _debug_draws(env, 0, n, ctr, tag, nzb, nub)
            ref = np.zeros_like(words)
            for j in range(n):
                for b in range(nzb + nub):
                    ref[j, b] = _oracle_words(seed, base + j, ctr, tag, b if b < nzb else 16 + (b - nzb))
            assert np.array_equal(words, ref), f'raw Philox words differ (tag {tag}, counter {ctr})'
            u_ref = _u01_f64(ref[:, nzb:]).astype(np.float32)            # (2k+1)/2^24: exact in float32
            assert np.array_equal(draws[:, nzb:].view(np.uint32), u_ref.view(np.uint32)), 'u01 bits differ'
            z_ref = _box_muller_f64(ref[:, :nzb])
            assert np.max(np.abs(draws[:, :nzb] - z_ref)) < NORMAL_ABS_BOUND
    # the oracle's own event generator agrees with this decomposition (ties the words to the slots the step consumes)
    env.close()


def test_box_muller_every_u01_value():
    """Every 23-bit u01 mantissa through the kernel's Box-Muller: as the radius word (x 64 angle words) and as the angle
    word (x 64 radius words).  No NaN / inf anywhere (u01 never reaches 0 or 1), absolute error against float64 below
    NORMAL_ABS_BOUND, and the u01 values exact."""
    rng = np.random.default_rng(5)
    L = abi.lib()
    worst = 0.0
    others = rng.integers(0, 2 ** 32, 64, dtype=np.uint64).astype(np.uint32)
    others[:4] = [0, 0xffffffff, 0x80000000, 0x7fffffff]              # angle -pi, +pi, 0; radius extremes
    sweep = (np.arange(2 ** 23, dtype=np.uint64) << 9).astype(np.uint32)
    for k in range(64):
        w = np.empty((2 ** 23, 4), np.uint32)
        w[:, 0] = sweep; w[:, 1] = others[k]                              # pair 0: radius sweeps, angle fixed
        w[:, 2] = others[k]; w[:, 3] = sweep                              # pair 1: radius fixed, angle sweeps
        z = np.empty((2 ** 23, 4), np.float32)
        u = np.empty((2 ** 23, 4), np.float32)
        abi.check(L.cf_debug_box_muller(0, w.ctypes.data_as(C.c_void_p), w.shape[0], z.ctypes.data_as(C.c_void_p),
                                        u.ctypes.data_as(C.c_void_p)))
        assert np.isfinite(z).all(), 'Box-Muller produced a non-finite normal'
        if k == 0:
            assert np.array_equal(u, _u01_f64(w).astype(np.float32))
            assert u.min() > 0.0 and u.max() < 1.0
        worst = max(worst, float(np.max(np.abs(z - _box_muller_f64(w)))))
    print(f'worst |z_gpu - z_f64| over 2 x 64 x 2^23 draws: {worst:.3e}')
    assert worst < NORMAL_ABS_BOUND


# free-running 60-step drift of the Philox path (relative, worst over 2048 envs that have not diverged in their done flags);
# measured on B200: see profiles/r2_rng_bits.json
DRIFT_BOUND_60 = 2e-2
STEP_BOUND = 5e-5


def test_philox_stream_60_steps_with_resets():
    import torch
    n = 2048
    env_id = 'DroneHoverBulletFreeEnvWithRandomHJAdversary-v0'
    cfg = config_for(env_id, {})
    for l in hj_tables.LEVELS:
        if l > 0:
            co.set_table(hj_tables.level_index(l), hj_tables.synthetic_table(l, 15))
    free = co.OracleEnvs(cfg, n, seed=123, gid_base=1000)       # never re-synchronised
    sync = co.OracleEnvs(cfg, n, seed=123, gid_base=1000)       # set to the GPU's state after every step
    env = make(env_id, num_envs=n, seed=123, env_id_base=1000, hj_table_seed=15)
    o = env.reset().cpu().numpy()
    assert rel_err(o, free.reset()) < STEP_BOUND
    sync.reset()
    sync.set_state(co.state_record_from_f32(env.get_state()))
    rng = np.random.default_rng(0)
    hover = 2.0 / 1.8 - 1.0
    agree = np.ones(n, bool)
    worst_step = worst_free = 0.0
    resets = flag_mismatch = 0
    for t in range(60):
        # a quarter of the envs get rough actions so that episodes end and in-place resets (tag 1 events) happen
        a = (hover + 0.05 * rng.standard_normal((n, 4))).astype(np.float32)
        a[: n // 4] = rng.uniform(-1, 1, (n // 4, 4)).astype(np.float32)
        o, r, d, info = env.step(torch.tensor(a, device='cuda'))
        o, r, d = o.cpu().numpy(), r.cpu().numpy(), d.cpu().numpy().astype(np.uint8)
        o_s, r_s, d_s, c_s, _, _ = sync.step(a)
        same = d == d_s
        flag_mismatch += int((~same).sum())
        worst_step = max(worst_step, rel_err(o[same], o_s[same]), rel_err(r[same], r_s[same]))
        resets += int(d.sum())
        st = co.state_record_from_f32(env.get_state())
        ctr_w, misc_w = abi.STATE_FIELDS['rng_ctr'][0], abi.STATE_FIELDS['misc'][0]
        st_s = sync.get_state()
        assert np.array_equal(st[same, ctr_w], st_s[same, ctr_w]), 'event counters diverged'
        assert np.array_equal(st[same, misc_w] % 256, st_s[same, misc_w] % 256), 'Boltzmann level draws differ'
        sync.set_state(st)
        o_f, r_f, d_f, _, _, _ = free.step(a)
        agree &= d == d_f
        worst_free = max(worst_free, rel_err(o[agree], o_f[agree]))
    print(f'philox 60 steps: per-step worst {worst_step:.3e}, free-running worst {worst_free:.3e}, resets {resets}, '
          f'done-flag mismatches (knife-edge) {flag_mismatch}, envs still flag-identical after 60 steps {int(agree.sum())}/{n}')
    assert resets > 200
    assert worst_step < STEP_BOUND
    assert flag_mismatch <= 2
    assert agree.mean() > 0.99 and worst_free < DRIFT_BOUND_60
    env.close()


def test_noise_state_moments_philox():
    """OU thrust-noise state, gyro bias walk and gyro low-pass output in Philox mode against the closed forms of
    envs/utils.py:112-134 (OU: theta 0.15, sigma 0.2 * 0.05), envs/sensors.py:121-134 (rotorS gyro model at dt = 1/200) and
    envs/utils.py:85-105 (low-pass y <- 0.5 y + 0.5 u)."""
    import torch
    n, T = 65536, 40
    env = make('DroneHoverBulletFreeEnvWithoutAdversary-v0', num_envs=n, seed=77, domain_randomization=-1, auto_reset=0,
               max_episode_steps=0)
    env.reset()
    hover = torch.full((n, 4), 2.0 / 1.8 - 1.0, device='cuda')
    for _ in range(T):
        env.step(hover)
    st = co.state_record_from_f32(env.get_state())
    F = abi.STATE_FIELDS
    ou = st[:, F['ou'][0]:F['ou'][0] + 4]
    # x <- 0.85 x + 0.01 z, two updates per control step from x = 0: var = 1e-4 (1 - 0.85^(4T)) / (1 - 0.85^2)
    var_ou = 1e-4 * (1 - 0.85 ** (4 * T)) / (1 - 0.85 ** 2)
    assert abs(ou.mean()) < 3e-4 and abs(ou.var() / var_ou - 1) < 0.02, (ou.mean(), ou.var(), var_ou)
    # b <- pi b + sigma_b z: one constructor-time draw, two per reset, three per control step (a12)
    dt, corr = 1.0 / 200.0, 1000.0
    sg = 0.000175 / dt ** 0.5
    sigma_b = (-(sg ** 2) * (corr / 2) * (np.exp(-2 * dt / corr) - 1)) ** 0.5
    pi_g = np.exp(-dt / corr)
    k = 1 + 2 + 3 * T
    var_b = sigma_b ** 2 * (1 - pi_g ** (2 * k)) / (1 - pi_g ** 2)
    bias = st[:, F['gyro_bias'][0]:F['gyro_bias'][0] + 3]
    assert abs(bias.mean()) < 5 * np.sqrt(var_b / bias.size) and abs(bias.var() / var_b - 1) < 0.02, (bias.var(), var_b)
    # low-pass of (w + bias + 0.0105 z + deg2rad(5) z'): the white part has stationary variance s^2 / 3; the true body rate in
    # steady hover (constant hover action, no disturbance) contributes its own small variance, bounded from the state itself
    lpf = st[:, F['lpf'][0]:F['lpf'][0] + 3]
    s2 = 0.0105 ** 2 + np.deg2rad(5.0) ** 2
    ow = st[:, F['omega_w'][0]:F['omega_w'][0] + 3]
    resid = lpf.var() - var_b - ow.var()
    assert abs(resid / (s2 / 3.0) - 1) < 0.05, (resid, s2 / 3.0, ow.var())
    env.close()
