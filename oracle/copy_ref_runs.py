"""Copy the reference's OWN logged training artefacts that the closed-loop acceptance test needs into
tests/golden/ref_runs/ (test infrastructure; outputs of the reference, not product code).

Runs in the build container (needs /root/reference).  Per run: the checkpoints ``torch_save/model{epoch}.pt``
(``state_dict`` of the reference's ActorCritic, ~50 KB each), ``validation_results.csv`` (8 deterministic-policy episodes
per validated epoch: level, steps, return - written by ``IWPGAlgorithm.validate``, algs/iwpg/iwpg.py:513-556), the first
rows of ``episodic_data.csv`` (the training episodes of epoch 0: random initial policy with exploration noise 0.509) and
the header + first rows of ``progress.csv``.  ``env_config.json`` is reduced to its scalar entries (the constants the env
was built with).  Usage: python oracle/copy_ref_runs.py
"""
import json
import os
import shutil

REF = '/root/reference/train_results_phoenix'
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, 'tests', 'golden', 'ref_runs')

RUNS = {
    'free_randomhj_61267': ('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0/ppo/2023_11_24_10_45/seed_61267', (0, 50, 100, 150)),
    'free_adv15_15431': ('DroneHoverBulletFreeEnvWithAdversary-v0/ppo/2023_11_21_15_24/seed_15431', (0, 50, 100, 150, 200)),
    'free_without_61305': ('DroneHoverBulletFreeEnvWithoutAdversary-v0/ppo/2023_11_24_10_46/seed_61305', (0, 50, 100, 150)),
}


def head(src, dst, n):
    with open(src) as f, open(dst, 'w') as g:
        for i, line in enumerate(f):
            if i >= n:
                break
            g.write(line)


def main():
    for name, (rel, epochs) in RUNS.items():
        src, dst = os.path.join(REF, rel), os.path.join(OUT, name)
        os.makedirs(dst, exist_ok=True)
        for e in epochs:
            shutil.copyfile(os.path.join(src, 'torch_save', f'model{e}.pt'), os.path.join(dst, f'model{e}.pt'))
        shutil.copyfile(os.path.join(src, 'validation_results.csv'), os.path.join(dst, 'validation_results.csv'))
        head(os.path.join(src, 'episodic_data.csv'), os.path.join(dst, 'episodic_data_head.csv'), 401)
        head(os.path.join(src, 'progress.csv'), os.path.join(dst, 'progress_head.csv'), 4)
        cfg = json.load(open(os.path.join(src, 'env_config.json')))
        small = {k: v for k, v in cfg.items() if isinstance(v, (int, float, str, bool)) or k == 'agent_params'}
        json.dump(dict(source=rel, env_config=small, config=json.load(open(os.path.join(src, 'config.json')))),
                  open(os.path.join(dst, 'run.json'), 'w'), indent=1, sort_keys=True)
        for p in os.listdir(dst):
            os.chmod(os.path.join(dst, p), 0o644)
        print(name, sorted(os.listdir(dst)))


if __name__ == '__main__':
    main()
