"""CPU: the restated Bullet step (oracle/fake_bullet.py) against the reference's own numpy dynamics,
``SimplePhysics.step_forward`` (envs/physics.py:127-200) - the one piece of rigid-body arithmetic the reference ships.

The fixture (oracle/make_simple_physics_golden.py, run where /root/reference exists) holds, for 2 000 random states,
the post-state of ONE substep taken by the reference's PyBulletPhysics.step_forward on the fake client (its own
agents.py force hand-off: LINK_FRAME thrust on the four prop links, yaw torque on the centre-of-mass link) and by the
reference's SimplePhysics from the same pre-state, with what SimplePhysics lacks switched off on the Bullet side
(damping B3, drag) - and a 1 000-substep rollout of both under one action sequence.

What this pins on the restated integrator (SURVEY section 7, B-list): B1 semi-implicit Euler ordering, B2 the
gyroscopic term (sign and presence), B4 forces cleared per step, B5/B6 the LINK_FRAME wrench (thrust direction in the
world, torque arms and signs, yaw torque passed to the base), and that the quaternion is advanced by the WORLD rate
on the left.  What it cannot pin: B3 (the 0.04 multibody damping; SimplePhysics has none) - shown below as the one
switch whose default moves the result away from SimplePhysics.
"""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), 'oracle'))
import fake_bullet as fb  # noqa: E402

FIX = os.path.join(HERE, 'golden', 'xcheck', 'simple_physics_crosscheck.npz')
POS, QUAT, VEL, OMW, WB, RPY = slice(0, 3), slice(3, 7), slice(7, 10), slice(10, 13), slice(13, 16), slice(16, 19)


@pytest.fixture(scope='module')
def fix():
    return dict(np.load(FIX))


def _wrap(a):
    return (a + np.pi) % (2 * np.pi) - np.pi


def _euler_rate_matrix(rpy):
    """rpy_dot = W(rpy) w_body for the ZYX (yaw-pitch-roll) convention of getEulerFromQuaternion."""
    sr, cr, tp, cp = np.sin(rpy[0]), np.cos(rpy[0]), np.tan(rpy[1]), np.cos(rpy[1])
    return np.array([[1.0, sr * tp, cr * tp], [0.0, cr, -sr], [0.0, sr / cp, cr / cp]])


def test_single_step_translation_and_body_rate_agree_to_rounding(fix):
    b, s = fix['ss_bullet'], fix['ss_simple']
    assert np.max(np.abs(b[:, POS] - s[:, POS])) < 1e-13          # B1, B5 (thrust direction), gravity, mass
    assert np.max(np.abs(b[:, VEL] - s[:, VEL])) < 1e-13
    assert np.max(np.abs(b[:, WB] - s[:, WB])) < 1e-11            # B2, B5/B6 (arms, signs), inverse inertia


def test_single_step_attitude_differs_by_the_euler_rate_kinematics_only(fix):
    """SimplePhysics: rpy' = rpy + dt w_body (exactly, by its own definition).  Bullet: exponential map of the world rate
    = rpy + dt W(rpy) w_body + O(dt^2 |w|^2).  So the two differ by dt (W - I) w_body to second order."""
    pre, b, s = fix['ss_pre'], fix['ss_bullet'], fix['ss_simple']
    dt = float(fix['ss_dt'])
    worst_simple = worst_bullet = worst_gap = 0.0
    for k in range(len(pre)):
        wb = s[k, WB]                                             # the new body rate, identical on both sides
        worst_simple = max(worst_simple, np.max(np.abs(s[k, RPY] - (pre[k, RPY] + dt * wb))))
        W = _euler_rate_matrix(pre[k, RPY])
        pred = pre[k, RPY] + dt * (W @ wb)
        worst_bullet = max(worst_bullet, np.max(np.abs(_wrap(b[k, RPY] - pred))))
        worst_gap = max(worst_gap, np.max(np.abs(_wrap(b[k, RPY] - s[k, RPY]))))
    assert worst_simple < 1e-14
    # second-order remainder: |w| <= 3 sqrt(3) + dt alpha, dt = 5 ms -> (dt |w|)^2 ~ 1e-3 at the very most; measured 2.2e-4 (first-order gap to SimplePhysics: 1.1e-2)
    assert worst_bullet < 5e-4, worst_bullet
    assert worst_gap > 20 * worst_bullet                          # the first-order term is what separates the two models


def _replay(fix, k, model):
    """One stepSimulation of fake_bullet from sample k's pre-state and recorded motor forces (the hand-off of
    agents.py:300-337 written out: prop links 0-3 LINK_FRAME force (0,0,f), link 4 LINK_FRAME torque (0,0,tz))."""
    bc = fb.BulletClient()
    bc.model = model
    bc.drone_id = 1
    bc.link_pos = np.array([[0.028, -0.028, 0.0108], [-0.028, -0.028, 0.0108], [-0.028, 0.028, 0.0108],
                            [0.028, 0.028, 0.0108], [0.0, 0.0, 0.0]])
    bc.mass, bc.inertia, bc.dt = float(fix['ss_mass']), fix['ss_J'].copy(), float(fix['ss_dt'])
    pre = fix['ss_pre'][k]
    bc.pos, bc.quat, bc.vel, bc.omega = pre[POS].copy(), pre[QUAT].copy(), pre[VEL].copy(), pre[OMW].copy()
    for i in range(4):
        bc.applyExternalForce(1, i, [0.0, 0.0, fix['ss_forces'][k, i]], [0.0, 0.0, 0.0], fb.LINK_FRAME)
    bc.applyExternalTorque(1, 4, [0.0, 0.0, fix['ss_ztorque'][k]], fb.LINK_FRAME)
    bc.stepSimulation()
    R = bc._R()
    return bc.pos, bc.vel, R.T @ bc.omega


def test_fixture_is_what_fake_bullet_computes_and_switches_are_visible(fix):
    """(1) the recorded Bullet-side post-states are reproduced by the committed fake_bullet (guards the fixture against
    drift of the restatement); (2) sensitivity: without the gyroscopic term the body rates no longer match
    SimplePhysics, and with Bullet's default damping neither do the velocities - the cross-check can see B2, and B3 is
    the one assumption it leaves open."""
    s = fix['ss_simple']
    off = fb.BulletModel(gyroscopic=True, linear_damping=0.0, angular_damping=0.0)
    nogyro = fb.BulletModel(gyroscopic=False, linear_damping=0.0, angular_damping=0.0)
    damped = fb.BulletModel()
    e_ok = e_gyro = e_damp_v = e_damp_w = 0.0
    for k in range(0, 400):
        p, v, w = _replay(fix, k, off)
        e_ok = max(e_ok, np.max(np.abs(p - fix['ss_bullet'][k, POS])), np.max(np.abs(v - fix['ss_bullet'][k, VEL])),
                   np.max(np.abs(w - fix['ss_bullet'][k, WB])))
        _, _, w2 = _replay(fix, k, nogyro)
        e_gyro = max(e_gyro, np.max(np.abs(w2 - s[k, WB])))
        _, v3, w3 = _replay(fix, k, damped)
        e_damp_v = max(e_damp_v, np.max(np.abs(v3 - s[k, VEL])))
        e_damp_w = max(e_damp_w, np.max(np.abs(w3 - s[k, WB])))
    assert e_ok < 1e-12
    assert e_gyro > 1e-3                       # dt |w x Jw| / J ~ 0.005 * 9 * (2.64 - 1.33) / 1.33
    assert 1e-4 < e_damp_v < 2e-3 and 1e-4 < e_damp_w < 5e-3    # dt * 0.04 (1 + |v|) |v|: the size of what B3 asserts


def test_rollout_1000_substeps(fix):
    """Same 1 000 actions (attitude PD with +-0.2 rad set-point steps) through both models: the body rates stay equal to
    rounding for the whole 5 s (their dynamics do not involve the attitude), attitude / velocity / position drift apart at
    the rate of SimplePhysics' Euler-rate approximation.  Bounds = 2x the measured values."""
    b, s = fix['ro_bullet'], fix['ro_simple']
    assert np.max(np.abs(b[:, WB] - s[:, WB])) < 1e-12
    d = np.abs(b - s)
    assert d[:100, POS].max() < 2.1e-3 and d[:100, VEL].max() < 1.6e-2 and np.abs(_wrap(b[:100, RPY] - s[:100, RPY])).max() < 4.5e-2
    assert d[:, POS].max() < 0.45 and d[:, VEL].max() < 0.33 and np.abs(_wrap(b[:, RPY] - s[:, RPY])).max() < 0.25


@pytest.mark.skipif(not os.path.isdir('/root/reference/phoenix_drone_simulation'), reason='reference checkout not present')
def test_fixture_regenerates_from_the_reference(fix):
    """Where the reference exists (the build container): the first single-step samples regenerate bit for bit."""
    import make_simple_physics_golden as gen
    out = gen.single_steps(n=20, seed=0)
    for key in ('ss_pre', 'ss_bullet', 'ss_simple', 'ss_forces'):
        assert np.array_equal(out[key], fix[key][:20]), key
