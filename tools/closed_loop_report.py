"""GPU: fly the reference's own checkpoints in the CUDA env and place every logged validation row inside the ensemble
(tests/closed_loop.py); repeat with the Bullet-model switches B2 / B3 changed.  Writes one JSON report.
Usage: python tools/closed_loop_report.py OUT.json [--n 16384] [--switches]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import closed_loop as cl  # noqa: E402


def main():
    out_path = sys.argv[1]
    n = int(sys.argv[sys.argv.index('--n') + 1]) if '--n' in sys.argv else 16384
    settings = {'default (gyroscopic on, damping 0.04)': {}}
    if '--switches' in sys.argv:
        settings['damping 0'] = dict(linear_damping=0.0, angular_damping=0.0)
        settings['gyroscopic off'] = dict(gyroscopic=0)
        settings['damping 0.08'] = dict(linear_damping=0.08, angular_damping=0.08)
    report = {}
    t0 = time.time()
    for sname, kw in settings.items():
        rep = report[sname] = {}
        for run, (env_id, epochs) in cl.RUNS.items():
            rows = cl.validation_rows(run)
            for ep in epochs:
                ac = cl.load_actor_critic(run, ep, torch.device('cuda'))
                nn = n * 8 if 'RandomHJ' in env_id else n
                ens = cl.fly(env_id, ac, nn, seed=100 + ep, **kw)
                placed = cl.place_rows([r for r in rows if r[0] == ep], ens)
                s = cl.summarize(placed)
                rep[f'{run}/epoch{ep}'] = dict(summary=s, ens_mean_steps=float(ens['steps'].mean()),
                                               ens_survive=float((ens['steps'] >= 500).mean()),
                                               ens_mean_ret=float(ens['ret'].mean()),
                                               ref_mean_steps=float(np.mean([r[2] for r in rows if r[0] == ep])),
                                               ref_mean_ret=float(np.mean([r[3] for r in rows if r[0] == ep])), rows=placed)
                print(sname, run, ep, s, 'ens steps %.1f surv %.3f ret %.1f | ref steps %.1f ret %.1f  (%.0fs)' % (
                    ens['steps'].mean(), (ens['steps'] >= 500).mean(), ens['ret'].mean(), rep[f'{run}/epoch{ep}']['ref_mean_steps'],
                    rep[f'{run}/epoch{ep}']['ref_mean_ret'], time.time() - t0), flush=True)
        # epoch-0 training episodes: stochastic policy, exploration std 0.509, un-normalised observations
        for run, (env_id, _) in cl.RUNS.items():
            ac = cl.load_actor_critic(run, 0, torch.device('cuda'))
            ac.pi.log_std.data.fill_(float(np.log(0.509)))
            ens = cl.fly(env_id, ac, n, seed=7, deterministic=False, identity_obs_norm=True, **kw)
            ref = np.array(cl.episodic_rows(run))
            rep[f'{run}/training_epoch0'] = dict(ens_steps_q=[float(q) for q in np.quantile(ens['steps'], [0.05, 0.25, 0.5, 0.75, 0.95])],
                                                 ref_steps_q=[float(q) for q in np.quantile(ref[:, 1], [0.05, 0.25, 0.5, 0.75, 0.95])],
                                                 ens_mean_steps=float(ens['steps'].mean()), ref_mean_steps=float(ref[:, 1].mean()),
                                                 ens_mean_ret=float(ens['ret'].mean()), ref_mean_ret=float(ref[:, 2].mean()))
            print(sname, run, 'training epoch 0', rep[f'{run}/training_epoch0'], flush=True)
    json.dump(report, open(out_path, 'w'), indent=1)


if __name__ == '__main__':
    main()
