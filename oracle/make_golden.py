"""Generate tests/golden/*.npz by running the reference's OWN Python env stack
(/root/reference, on oracle/fake_bullet.py) with every np.random draw recorded.

TEST INFRASTRUCTURE ONLY.  Runs in the build container (needs /root/reference); the fixtures it
writes travel to the GPU box.  Usage:  python oracle/make_golden.py [--check-only]

Each fixture is one single-env trajectory of events (kind 0 = reset, 1 = step) holding, per
event: the action, the standard draws in the slot layout of include/cf_abi.h / DESIGN.md, the
reference's outputs (obs, reward, done, cost, TimeLimit.truncated) and the full env state record
(CF_S_* layout, float64) after the event, plus the record before the first event.
After writing, every fixture is replayed through oracle-B (oracle/cf_oracle.c) and the maximum
deviation is printed; tests/test_oracle_golden.py asserts the same bound.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_harness as rh  # noqa: E402
from disturbance_crazyfile_simulation_b200 import abi, hj_tables  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, 'tests', 'golden')
NZ_MAX, NU_MAX = abi.CF_RESET_NORMALS, abi.CF_RESET_UNIFORMS

# name -> (env id, constructor kwargs, steps, action source, np.random seed, table seed)
CASES = {
    # BASELINE.json configs[0]: 1 env, 1000 steps of fixed-seed random actions, noise/DR/reset-dist off
    'free_without_clean': ('DroneHoverBulletFreeEnvWithoutAdversary-v0',
                           dict(observation_noise=-1, domain_randomization=-1, enable_reset_distribution=False),
                           1000, 'uniform', 0, None),
    'free_without_full': ('DroneHoverBulletFreeEnvWithoutAdversary-v0', {}, 300, 'hover', 1, None),
    'free_adv15_full': ('DroneHoverBulletFreeEnvWithAdversary-v0', {}, 400, 'hover', 2, 15),
    'free_randomhj_full': ('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', {}, 400, 'mixed', 3, 15),
    'initial_adv_full': ('DroneHoverBulletEnvWithAdversaryInitial-v0', {}, 300, 'hover', 4, 15),
    'random_box_full': ('DroneHoverBulletEnvWithRandomAdversary-v0', {}, 200, 'hover', 5, None),
    'bullet_base_full': ('DroneHoverBulletEnv-v0', {}, 200, 'hover', 6, None),
    'target_randomhj_nodr': ('DroneHoverBulletEnvWithRandomHJAdversary-v0', dict(domain_randomization=-1), 200, 'mixed', 7, 15),
    # BASELINE.json configs[3] with the ground-effect switch ON (physics.py:27-58, 118-120, 244-246; the reference never
    # enables it): the reference object's own `physics.use_ground_effect` is set after construction; the 'descend'
    # actions let the drone sink from ~1 m to the z < 0.2 m termination again and again, so the term grows from 4e-4 to
    # 1e-2 of the thrust over the pre-states of the fixture
    'initial_adv_ground': ('DroneHoverBulletEnvWithAdversaryInitial-v0', dict(use_ground_effect=1), 400, 'descend', 8, 15),
    # non-default agent kwargs of DroneBaseEnv.__init__ (envs/base.py:26-47): latency, motor time constant, thrust noise
    'free_adv15_agentkw': ('DroneHoverBulletFreeEnvWithAdversary-v0',
                           dict(latency=0.012, motor_time_constant=0.12, motor_thrust_noise=0.10), 200, 'hover', 9, 15),
}
# kwargs of ours that the reference constructor does not take: applied to the reference object after construction
POST_KWARGS = {'use_ground_effect': lambda env, v: setattr(env.unwrapped.physics, 'use_ground_effect', bool(v))}


def actions_for(kind: str, steps: int, seed: int) -> np.ndarray:
    rng = np.random.default_rng(seed)
    if kind == 'uniform':
        return rng.uniform(-1, 1, (steps, 4)).astype(np.float32)
    hover = 2.0 / 1.8 - 1.0
    a = (hover + 0.05 * rng.standard_normal((steps, 4))).astype(np.float32)
    if kind == 'descend':          # below hover thrust: sink to the ground-proximity termination, with a little attitude excitation
        a = (hover - 0.12 + 0.05 * rng.standard_normal((steps, 4))).astype(np.float32)
    if kind == 'mixed':            # some out-of-range and wild actions to hit the clip / done paths
        wild = rng.uniform(-1.5, 1.5, (steps, 4)).astype(np.float32)
        pick = rng.uniform(size=steps) < 0.15
        a[pick] = wild[pick]
    return a


def record_of(env, ep_ret: float) -> np.ndarray:
    """Reference env -> CF_S_* state record (float64; integer words as exact doubles)."""
    s = rh.env_snapshot(env)
    r = np.zeros(abi.CF_STATE_WORDS)
    F = abi.STATE_FIELDS

    def put(name, v):
        o, n = F[name]
        r[o:o + n] = np.asarray(v, dtype=np.float64).reshape(-1)

    put('pos', s['pos']); put('ep_ret', ep_ret); put('quat', s['quat']); put('vel', s['vel'])
    put('omega_w', s['omega_world']); put('lpf', s['lpf']); put('gyro_bias', s['gyro_bias'])
    put('motor_x', s['motor_x']); put('ou', s['ou']); put('ring', s['ring'])
    if len(s['act_hist']):      # the history deques are empty until the first reset
        put('act_prev1', s['act_hist'][-1][:4]); put('act_prev2', s['act_hist'][0][:4])
        put('obs_prev', s['obs_hist'][-1][:13])
    elapsed = getattr(env, '_elapsed_steps', None)
    put('ep_len', 0 if elapsed is None else elapsed)
    level_idx = int(round(float(s['level']) * 10))
    put('misc', level_idx | (int(s['ring_idx']) << 8) | ((int(s['iteration']) & 1) << 12))
    put('dt', s['dt']); put('mass', s['mass']); put('inertia', s['inertia']); put('ftf1', s['ftf1'])
    put('motor_B', s['motor_B']); put('motor_K', s['motor_K'])
    return r


class SlotParser:
    """Turns the recorder log of one event into the slot-major standard-draw arrays."""

    def __init__(self, log):
        self.log = list(log)

    def pop(self, kind, n):
        k, v = self.log.pop(0)
        assert k == kind and len(v) == n, (k, len(v), kind, n)
        return v

    def obs_full(self, nz, un, zo, uo):
        """compute_observation 100 Hz branch: N3 pos,U3,N3 vel,U3(dead),N3 bias,N3 rw,N3 turn-on,N3 rot,U3,N3,N3 (acc, dead)"""
        nz[zo:zo + 3] = self.pop('normal', 3); un[uo:uo + 3] = self.pop('uniform', 3)
        nz[zo + 3:zo + 6] = self.pop('normal', 3); self.pop('uniform', 3)
        for k in range(3):
            nz[zo + 6 + 3 * k:zo + 9 + 3 * k] = self.pop('normal', 3)
        nz[zo + 15:zo + 18] = self.pop('normal', 3); un[uo + 3:uo + 6] = self.pop('uniform', 3)
        self.pop('normal', 3); self.pop('normal', 3)

    def gyro_only(self, nz, zo):
        for k in range(3):
            nz[zo + 3 * k:zo + 3 + 3 * k] = self.pop('normal', 3)

    def done(self):
        assert not self.log, self.log


def parse_reset(log, noise, dr, reset_dist, boltz):
    nz, un = np.zeros(NZ_MAX), np.full(NU_MAX, 0.5)
    p = SlotParser(log)
    if reset_dist:
        un[0:3] = p.pop('uniform', 3); un[3:6] = p.pop('uniform', 3); un[6] = p.pop('uniform', 1)[0]
        un[7:10] = p.pop('uniform', 3); un[10:13] = p.pop('uniform', 3); un[13] = p.pop('uniform', 1)[0]
        nz[0:4] = p.pop('normal', 4); nz[4:12] = p.pop('normal', 8)
    if dr:
        un[14] = p.pop('uniform', 1)[0]; un[15] = p.pop('uniform', 1)[0]; un[16:19] = p.pop('uniform', 3)
        un[19] = p.pop('uniform', 1)[0]; un[20] = p.pop('uniform', 1)[0]
        un[21:25] = p.pop('uniform', 4); un[25:29] = p.pop('uniform', 4)
    if noise:
        p.obs_full(nz, un, 12, 29)
        p.obs_full(nz, un, 30, 35)
    if boltz:
        un[41] = p.pop('choice', 1)[0]
    p.done()
    return nz, un


def parse_step(log, noise, random_box):
    nz, un = np.zeros(NZ_MAX), np.full(NU_MAX, 0.5)
    p = SlotParser(log)
    if random_box:
        u = p.pop('uniform', 3)
        un[6:8] = u[:2]
    scratch_z, scratch_u = np.zeros(18), np.zeros(6)
    nz[0:4] = p.pop('normal', 4)
    if noise:   # full call whose pos/vel/rot part is dead: keep only its gyro normals
        p.obs_full(scratch_z, scratch_u, 0, 0)
        nz[4:13] = scratch_z[6:15]
    nz[13:17] = p.pop('normal', 4)
    if noise:
        p.gyro_only(nz, 17)
        p.obs_full(nz, un, 26, 0)
    p.done()
    return nz, un


def run_case(name):
    env_id, kwargs, steps, act_kind, seed, table_seed = CASES[name]
    cfg = abi_config(env_id, kwargs)
    noise, dr = cfg.observation_noise > 0, cfg.domain_randomization > 0
    reset_dist = bool(cfg.enable_reset_distribution)
    boltz = cfg.disturbance_kind == abi.CF_DSTB_HJ_BOLTZMANN
    hj = cfg.disturbance_kind in (abi.CF_DSTB_HJ_FIXED, abi.CF_DSTB_HJ_BOLTZMANN)
    random_box = cfg.disturbance_kind == abi.CF_DSTB_RANDOM_BOX
    tables = {}
    if hj:
        levels = hj_tables.LEVELS if boltz else (cfg.disturbance_level,)
        tables = {l: hj_tables.synthetic_table(l, table_seed) for l in levels}
    acts = actions_for(act_kind, steps, seed + 100)
    ev = dict(kind=[], act=[], nz=[], un=[], obs=[], rew=[], done=[], cost=[], trunc=[], state=[])

    def push(kind, a, nz, un, obs, rew, done, cost, trunc, state):
        for k, v in zip(ev, (kind, a, nz, un, obs, rew, done, cost, trunc, state)):
            ev[k].append(v)

    np.random.seed(seed)
    with rh.table_dir(tables), rh.StdDrawRecorder() as rec:
        env = rh.make_ref_env(env_id, **{k: v for k, v in kwargs.items() if k not in POST_KWARGS})
        for k, v in kwargs.items():
            if k in POST_KWARGS:
                POST_KWARGS[k](env, v)
        rec.take()
        state0 = record_of(env, 0.0)
        ep_ret = 0.0
        obs = env.reset()
        nz, un = parse_reset(rec.take(), noise, dr, reset_dist, boltz)
        push(0, np.zeros(4), nz, un, np.asarray(obs, np.float64), 0.0, 0, 0.0, 0, record_of(env, 0.0))
        for t in range(steps):
            a = acts[t]
            obs, r, done, info = env.step(a)
            ep_ret += r
            nz, un = parse_step(rec.take(), noise, random_box)
            push(1, a.astype(np.float64), nz, un, np.asarray(obs, np.float64), float(r), int(bool(done)),
                 float(info['cost']), int(bool(info.get('TimeLimit.truncated', False))), record_of(env, ep_ret))
            if done:
                ep_ret = 0.0
                obs = env.reset()
                nz, un = parse_reset(rec.take(), noise, dr, reset_dist, boltz)
                push(0, np.zeros(4), nz, un, np.asarray(obs, np.float64), 0.0, 0, 0.0, 0, record_of(env, 0.0))
    out = {k: np.array(v) for k, v in ev.items()}
    out['kind'] = out['kind'].astype(np.int8)
    out['done'] = out['done'].astype(np.uint8)
    out['trunc'] = out['trunc'].astype(np.uint8)
    out['state0'] = state0
    out['meta'] = np.array(json.dumps(dict(env_id=env_id, kwargs={k: (bool(v) if isinstance(v, bool) else v) for k, v in kwargs.items()},
                                           table_seed=table_seed, np_seed=seed, actions=act_kind)))
    return out


def abi_config(env_id, kwargs):
    """cf_config for a reference constructor call (same mapping as vec_env.make, without a GPU)."""
    from disturbance_crazyfile_simulation_b200.vec_env import _KWARGS
    import ctypes as C
    cfg = abi.config_default(env_id)
    for k, v in kwargs.items():
        field = _KWARGS[k]
        ftype = dict(abi.CfConfig._fields_)[field]
        setattr(cfg, field, int(v) if ftype is C.c_int32 else float(v))
    return cfg


def replay_oracle(fix, cfg=None):
    """Replay a fixture through oracle-B; returns dict of max deviations."""
    import cf_oracle as co
    meta = json.loads(str(fix['meta']))
    cfg = abi_config(meta['env_id'], meta['kwargs']) if cfg is None else cfg
    cfg.rng_mode, cfg.auto_reset = abi.CF_RNG_INJECTED, 0
    if meta['table_seed'] is not None:
        boltz = cfg.disturbance_kind == abi.CF_DSTB_HJ_BOLTZMANN
        for l in (hj_tables.LEVELS if boltz else (cfg.disturbance_level,)):
            co.set_table(hj_tables.level_index(l), hj_tables.synthetic_table(l, meta['table_seed']))
    orc = co.OracleEnvs(cfg, 1)
    orc.set_state(fix['state0'][None])
    dev = dict(obs=0.0, rew=0.0, state=0.0, flags=0)
    for i, kind in enumerate(fix['kind']):
        nz, un = fix['nz'][i], fix['un'][i]
        if kind == 0:
            obs = orc.reset(None, nz[:, None], un[:, None])[0]
        else:
            obs, rew, done, cost, trunc, _ = orc.step(fix['act'][i][None], nz[:abi.CF_STEP_NORMALS, None],
                                                      un[:abi.CF_STEP_UNIFORMS, None])
            obs = obs[0]
            dev['rew'] = max(dev['rew'], abs(rew[0] - fix['rew'][i]) / max(1.0, abs(fix['rew'][i])))
            dev['flags'] += int(done[0] != fix['done'][i]) + int(cost[0] != fix['cost'][i]) + int(trunc[0] != fix['trunc'][i])
        dev['obs'] = max(dev['obs'], float(np.max(np.abs(obs - fix['obs'][i]))))
        st, ref = orc.get_state()[0], fix['state'][i].copy()
        skip = np.zeros(abi.CF_STATE_WORDS, bool)
        skip[abi.STATE_FIELDS['rng_ctr'][0]] = True
        o, n = abi.STATE_FIELDS['act_prev2']
        if ref[abi.STATE_FIELDS['ep_len'][0]] == 1:
            skip[o:o + n] = True        # aliased view of the action ring: value unused
        dev['state'] = max(dev['state'], float(np.max(np.abs(st - ref)[~skip])))
    return dev


def main():
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    check_only = '--check-only' in sys.argv
    names = [a for a in sys.argv[1:] if not a.startswith('--')] or list(CASES)
    for name in names:
        path = os.path.join(GOLDEN_DIR, name + '.npz')
        if not check_only:
            fix = run_case(name)
            np.savez_compressed(path, **fix)
        fix = dict(np.load(path))
        dev = replay_oracle(fix)
        n_reset = int((fix['kind'] == 0).sum())
        print(f'{name:24s} events={len(fix["kind"])} resets={n_reset} size={os.path.getsize(path) / 1024:.0f} KiB  oracle-B dev: {dev}')


if __name__ == '__main__':
    main()
