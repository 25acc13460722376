"""Closed-loop statistical acceptance against the reference's OWN logged runs.

tests/golden/ref_runs/* holds checkpoints and logs the reference wrote while training on real PyBullet with its real HJ
tables (copied by oracle/copy_ref_runs.py).  Here the same policies fly in the CUDA env - restated Bullet step (a8/a10)
and tables from the CUDA FasTrack solver (f3) - as large ensembles of the deterministic-policy validation episodes of
``IWPGAlgorithm.val_once`` (algs/iwpg/iwpg.py:513-536), and every logged (level, steps, return) row is placed inside the
ensemble's distribution for its level.  Shared by tests/test_closed_loop_reference_runs.py and tools/closed_loop_report.py.
"""
import csv
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
RUN_DIR = os.path.join(HERE, 'golden', 'ref_runs')
RUNS = {
    'free_randomhj_61267': ('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', (0, 50, 100, 150)),
    'free_adv15_15431': ('DroneHoverBulletFreeEnvWithAdversary-v0', (0, 50, 100, 150, 200)),
    'free_without_61305': ('DroneHoverBulletFreeEnvWithoutAdversary-v0', (0, 50, 100, 150)),
}


def validation_rows(run):
    """[(epoch, level, steps, return)] of the reference's validation_results.csv."""
    rows = []
    with open(os.path.join(RUN_DIR, run, 'validation_results.csv')) as f:
        for r in csv.DictReader(f):
            rows.append((int(r['Training_Epoch']), round(float(r['disturbance_level']), 1), float(r['Steps']), float(r['Returns'])))
    return rows


def episodic_rows(run):
    """[(level, steps, return)] of the first training episodes (epoch 0) of episodic_data.csv."""
    rows = []
    with open(os.path.join(RUN_DIR, run, 'episodic_data_head.csv')) as f:
        for r in csv.DictReader(f):
            rows.append((round(float(r['disturbance_level']), 1), float(r['Steps']), float(r['Returns'])))
    return rows


def load_actor_critic(run, epoch, device):
    import torch
    from disturbance_crazyfile_simulation_b200.trainer import ActorCritic
    sd = torch.load(os.path.join(RUN_DIR, run, f'model{epoch}.pt'), map_location='cpu', weights_only=True)
    ac = ActorCritic(34, 4, use_scaled_rewards=True)
    missing, unexpected = ac.load_state_dict(sd, strict=True), None   # pi.*, v.*, obs_oms.*, ret_oms.*: every key must match
    return ac.to(device).eval()


def fly(env_id, ac, n, seed=1, deterministic=True, identity_obs_norm=False, max_steps=500, **env_kw):
    """``n`` validation episodes side by side (no auto-reset): returns per-episode (level, steps, return, cost)."""
    import torch
    from disturbance_crazyfile_simulation_b200.vec_env import CrazyFlieVecEnv
    env = CrazyFlieVecEnv(env_id, num_envs=n, seed=seed, auto_reset=0, hj_tables='solve', **env_kw)
    dev = env.device
    if identity_obs_norm:
        import copy
        ac = copy.deepcopy(ac)
        ac.obs_oms.mean.data.zero_(); ac.obs_oms.std.data.fill_(1.0)
    torch.manual_seed(seed)
    o = env.reset()
    levels = np.round(env.disturbance_level.astype(np.float64), 1)
    alive = torch.ones(n, dtype=torch.bool, device=dev)
    ret = torch.zeros(n, device=dev, dtype=torch.float64)
    cost = torch.zeros(n, device=dev)
    length = torch.zeros(n, dtype=torch.int32, device=dev)
    with torch.no_grad():
        for t in range(max_steps):
            a, _, _ = ac.step(o, deterministic=deterministic)
            o, r, d, info = env.step(a)
            ret += torch.where(alive, r.double(), torch.zeros_like(ret))
            cost += torch.where(alive, info['cost'], torch.zeros_like(cost))
            length += alive.to(torch.int32)
            alive &= ~d
            if t % 50 == 49 and not bool(alive.any()):
                break
    out = dict(level=levels, steps=length.cpu().numpy().astype(np.float64), ret=ret.cpu().numpy(), cost=cost.cpu().numpy())
    env.close()
    return out


def tail_prob(sample, x):
    """Two-sided tail probability of x inside an empirical sample (ties count on both sides)."""
    sample = np.asarray(sample)
    lo = float(np.mean(sample <= x))
    hi = float(np.mean(sample >= x))
    return min(1.0, 2.0 * min(lo, hi))


def place_rows(rows, ens):
    """For every logged (epoch, level, steps, return) row: tail probabilities inside the ensemble of its level."""
    out = []
    for (_, level, steps, ret) in rows:
        sel = np.abs(ens['level'] - level) < 0.05
        s, r = ens['steps'][sel], ens['ret'][sel]
        # returns are compared among episodes of the same fate (full length vs ended early): the -1000 terminal penalty
        # makes the unconditional distribution bimodal
        full = steps >= 500
        same = (s >= 500) if full else (s < 500)
        p_fate = float(np.mean(same)) if len(s) else 0.0
        p_steps = 1.0 if full else (tail_prob(s[same], steps) if same.any() else 0.0)
        p_ret = tail_prob(r[same] / np.maximum(s[same], 1.0), ret / max(steps, 1.0)) if same.any() else 0.0
        out.append(dict(level=level, steps=steps, ret=ret, n=int(sel.sum()), p_fate=p_fate, p_steps=p_steps, p_ret=p_ret,
                        ens_survive=float(np.mean(s >= 500)) if len(s) else float('nan'),
                        ens_ret_per_step_q=[float(q) for q in np.quantile(r[same] / np.maximum(s[same], 1.0), [0.01, 0.5, 0.99])]
                        if same.any() else None))
    return out


def summarize(placed, alpha=0.01):
    ok = [p['p_fate'] >= alpha and p['p_steps'] >= alpha and p['p_ret'] >= alpha for p in placed]
    return dict(rows=len(placed), consistent=int(sum(ok)))
