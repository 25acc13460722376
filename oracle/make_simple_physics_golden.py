"""Cross-check of the restated Bullet step (oracle/fake_bullet.py, B1/B2/B5/B6) against the ONE piece of
rigid-body arithmetic the reference ships itself: ``SimplePhysics.step_forward`` (envs/physics.py:127-200).

TEST INFRASTRUCTURE ONLY.  Runs in the build container (needs /root/reference); writes
``tests/golden/xcheck/simple_physics_crosscheck.npz`` which travels.  Usage: python oracle/make_simple_physics_golden.py

Both sides are the reference's OWN Python objects: one ``DroneHoverBulletEnv-v0`` (cf21x_bullet agent: PT1 motor
model, latency ring, the force hand-off of agents.py:300-337) whose ``physics`` is either the registered
``PyBulletPhysics`` on the fake client or a ``SimplePhysics`` instance built on the same drone.  To make the two
models state the same physics, what SimplePhysics does not have is switched off on the Bullet side (multibody
damping B3 = 0, DRAG_COEFF = 0) and the arm is made equal (``drone.L / sqrt(2) := 0.028``, the URDF prop offset;
the reference's L = 0.0397 gives 0.02807).  What remains different is the integrator: SimplePhysics adds
``dt * rpy_dot`` (BODY rates) to the Euler angles and integrates the body rates in the body frame; Bullet advances
the quaternion by the exponential map of the WORLD rate and integrates the world rate.  For one step from the same
state the two agree to rounding in position, velocity AND body rate (R'^T w' = R^T w' because R' differs from R by a
rotation about w' itself) and differ in attitude by exactly the Euler-rate kinematics SimplePhysics leaves out,
dt (W(rpy) - I) w_body + O(dt^2): that is what tests/test_simple_physics_crosscheck.py asserts, term by term.

Part 1 "single steps": 2 000 random states, one substep each from identical pre-states.
Part 2 "rollout": 1 000 substeps (5 s) of one action sequence (attitude PD on the SimplePhysics copy, fed to both).
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_harness as rh  # noqa: E402

OUT = os.path.join(ROOT, 'tests', 'golden', 'xcheck', 'simple_physics_crosscheck.npz')
KW = dict(observation_noise=-1, domain_randomization=-1, enable_reset_distribution=False, motor_thrust_noise=0.0)


def make_pair():
    """(env, bullet physics, simple physics) on one reference env object."""
    import fake_bullet
    env = rh.make_ref_env('DroneHoverBulletEnv-v0', **KW)
    e = env.unwrapped
    from phoenix_drone_simulation.envs.physics import SimplePhysics
    assert isinstance(e.bc, fake_bullet.BulletClient)
    e.bc.model.linear_damping = 0.0
    e.bc.model.angular_damping = 0.0
    e.drone.DRAG_COEFF = np.zeros(3)
    e.drone.L = 0.028 * np.sqrt(2.0)
    env.reset()
    simple = SimplePhysics(e.drone, e.bc, time_step=e.time_step)
    return env, e.physics, simple


def set_state(e, pos, quat, vel, omega_w, x):
    e.bc.resetBasePositionAndOrientation(e.drone.body_unique_id, pos, quat)
    e.bc.resetBaseVelocity(e.drone.body_unique_id, vel, omega_w)
    e.drone.update_information()
    e.drone.x = np.array(x, dtype=np.float64)
    e.drone.action_buffer[:] = 0.0
    e.drone.action_idx = 0


def get_state(e):
    d = e.drone
    d.update_information()
    return np.concatenate([d.xyz, d.quaternion, d.xyz_dot, np.asarray(e.bc.omega), d.rpy_dot, d.rpy])


def single_steps(n=2000, seed=0):
    env, bullet, simple = make_pair()
    e = env.unwrapped
    rng = np.random.default_rng(seed)
    pre, post_b, post_s, acts, xs, forces, ztq = [], [], [], [], [], [], []
    for _ in range(n):
        rpy = rng.uniform(-0.5, 0.5, 3) * np.array([1.0, 1.0, 6.0])
        quat = np.array(e.bc.getQuaternionFromEuler(rpy))
        pos = np.array([0.0, 0.0, 1.0]) + rng.uniform(-0.3, 0.3, 3)
        vel = rng.uniform(-1.0, 1.0, 3)
        om = rng.uniform(-3.0, 3.0, 3)
        x = rng.uniform(0.5, 0.95, 4)
        a = rng.uniform(-1.0, 1.0, 4)
        for phys, out in ((bullet, post_b), (simple, post_s)):
            set_state(e, pos, quat, vel, om, x)
            e.drone.action_buffer[:] = a                # the delayed action leaving the ring == a
            phys.step_forward(a)
            if phys is simple:
                # SimplePhysics writes drone.rpy_dot itself and hands R_old^T w to Bullet (physics.py:190-199);
                # its own state variables are what it integrates
                out.append(np.concatenate([e.drone.xyz, e.drone.quaternion, e.drone.xyz_dot, np.zeros(3), e.drone.rpy_dot,
                                           e.drone.rpy]))
            else:
                out.append(get_state(e))
        f = np.array(e.drone.y, dtype=np.float64)          # motor forces of the step just taken (agents.py:291)
        tq = e.drone.force_torque_factor_1 * f + e.drone.force_torque_factor_0
        forces.append(f); ztq.append(-tq[0] + tq[1] - tq[2] + tq[3])
        set_state(e, pos, quat, vel, om, x)
        pre.append(get_state(e)); acts.append(a); xs.append(x)
    return dict(ss_pre=np.array(pre), ss_bullet=np.array(post_b), ss_simple=np.array(post_s), ss_act=np.array(acts),
                ss_x=np.array(xs), ss_forces=np.array(forces), ss_ztorque=np.array(ztq), ss_dt=np.array(e.time_step), ss_mass=np.array(e.drone.m), ss_J=np.diag(e.drone.J).copy(),
                ss_K=np.array(e.drone.K), ss_B=np.broadcast_to(e.drone.B, (4,)).copy(), ss_ftf=np.array(
                    [e.drone.force_torque_factor_0, e.drone.force_torque_factor_1]))


def rollout(steps=1000):
    env_s, _, simple = make_pair()
    env_b, bullet, _ = make_pair()
    es, eb = env_s.unwrapped, env_b.unwrapped
    hover = es.drone.HOVER_ACTION
    traj_s, traj_b, acts = [], [], []
    for t in range(steps):
        d = es.drone
        # attitude PD with a roll / pitch set-point schedule (+-0.2 rad square waves), mixed like control.py:33-50
        sp = np.array([0.2 * np.sign(np.sin(2 * np.pi * t / 400.0)), 0.15 * np.sign(np.sin(2 * np.pi * t / 300.0 + 1.0)), 0.0])
        u = 0.6 * (6.0 * (sp - d.rpy) - d.rpy_dot) * np.array([0.05, 0.05, 0.02])
        T = hover + 0.3 * (1.0 - d.xyz[2]) - 0.2 * d.xyz_dot[2]
        r, p, y = u
        a = np.clip(np.array([T - r - p - y, T - r + p + y, T + r + p - y, T + r - p + y]), -1, 1)
        simple.step_forward(a)
        bullet.step_forward(a)
        traj_s.append(np.concatenate([d.xyz, d.quaternion, d.xyz_dot, np.zeros(3), d.rpy_dot, d.rpy]))
        traj_b.append(get_state(eb))
        acts.append(a)
    return dict(ro_simple=np.array(traj_s), ro_bullet=np.array(traj_b), ro_act=np.array(acts))


def main():
    out = single_steps()
    out.update(rollout())
    np.savez_compressed(OUT, **out)
    print('wrote', OUT, os.path.getsize(OUT) // 1024, 'KiB')
    d = np.abs(out['ss_bullet'] - out['ss_simple'])
    names = dict(pos=slice(0, 3), quat=slice(3, 7), vel=slice(7, 10), rate_body=slice(13, 16), rpy=slice(16, 19))
    for k, s in names.items():
        print(f'single step  {k:10s} max |bullet - simple| = {d[:, s].max():.3e}')
    r = np.abs(out['ro_bullet'] - out['ro_simple'])
    for k, s in names.items():
        print(f'rollout      {k:10s} max over 1000 substeps  = {r[:, s].max():.3e}   at 100: {r[:100, s].max():.3e}')


if __name__ == '__main__':
    main()
