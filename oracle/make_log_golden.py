"""make_log_golden.py - golden run-directory files written by the reference's OWN EpochLogger.

TEST INFRASTRUCTURE ONLY.  Runs in the build container (needs /root/reference); imports
``phoenix_drone_simulation/utils/loggers.py`` with the stubs of ref_harness plus a single-process stand-in for
mpi4py (a MagicMock would make ``num_procs() == 1`` false and zero every statistic), feeds it a seeded
sequence of ``store`` / ``log_tabular`` / ``log_episode`` / ``log_validation`` / ``save_config`` calls shaped like
``IWPGAlgorithm.log`` (algs/iwpg/iwpg.py:309-348), and commits what it wrote under ``tests/golden/logs/`` together
with the call sequence (``calls.json``) that tests/test_run_logger.py replays through ``RunLogger``.

    python oracle/make_log_golden.py
"""
import json
import os
import shutil
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(ROOT, 'tests', 'golden', 'logs')
FILES = ['progress.csv', 'episodic_data.csv', 'validation_results.csv', 'penalty_data.csv', 'config.json']


def single_process_mpi():
    mpi4py, MPI = types.ModuleType('mpi4py'), types.ModuleType('mpi4py.MPI')

    class _Comm:
        def Get_rank(self):
            return 0

        def Get_size(self):
            return 1

    MPI.COMM_WORLD, MPI.SUM, MPI.MIN, MPI.MAX = _Comm(), 'sum', 'min', 'max'
    mpi4py.MPI = MPI
    sys.modules['mpi4py'], sys.modules['mpi4py.MPI'] = mpi4py, MPI


def call_sequence(seed: int = 7, epochs: int = 3):
    """JSON-able list of logger calls; floats are Python floats so both sides see identical inputs."""
    rng = np.random.default_rng(seed)
    calls = [['save_config', {'alg': 'ppo', 'seed': 61267, 'pi_lr': 3e-4, 'ac_kwargs': {'pi': {'hidden_sizes': [50, 50]}},
                              'env_id': 'DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', 'use_reward_scaling': True}]]
    episode = 0
    for ep in range(epochs):
        for _ in range(int(rng.integers(3, 9))):
            ret, length = float(-rng.uniform(5, 900)), int(rng.integers(1, 501))
            episode += 1
            calls.append(['store', {'EpRet': ret, 'EpLen': length}])
            calls.append(['log_episode', [episode, round(float(rng.integers(0, 21)) / 10.0, 1), length, ret]])
        calls.append(['store_array', {'Values/V': [float(v) for v in rng.standard_normal(64)]}])
        calls.append(['store_array', {'Values/V': [float(v) for v in rng.standard_normal(64)]}])
        calls.append(['store', {'Values/Adv': float(rng.standard_normal())}])
        calls.append(['store', {'Values/Adv': float(rng.standard_normal())}])
        for k in ('Loss/Pi', 'Loss/Value', 'Loss/DeltaPi', 'Loss/DeltaValue', 'Entropy', 'KL', 'PolicyRatio'):
            calls.append(['store', {k: float(rng.standard_normal())}])
        calls.append(['store', {'Misc/StopIter': int(rng.integers(1, 81))}])
        calls.append(['epoch', {'Epoch': ep + 1, 'Misc/Seed': 61267, 'LR': 3e-4, 'Misc/RewScaleMean': float(rng.uniform(-9, 0)),
                                'Misc/RewScaleStddev': float(rng.uniform(1, 40)), 'TotalEnvSteps': (ep + 1) * 64000,
                                'Time': 10 * (ep + 1), 'FPS': 6400 + ep}])
        if ep == 1:
            calls.append(['log_validation', [[ep + 1] * 3, [0.0, 1.0, 2.0], [500, 312, 77],
                                             [float(-rng.uniform(5, 900)) for _ in range(3)], [0.0, 3.0, 41.0]]])
    calls.append(['log_penalty', [1, 0.0, 0.01, 0.2, 0.0, 0.05, 0.26]])
    return calls


def replay(logger, calls):
    """Drive a reference EpochLogger or a RunLogger with the same calls (column order of IWPGAlgorithm.log)."""
    for name, arg in calls:
        if name == 'save_config':
            logger.save_config(arg)
        elif name == 'store':
            logger.store(**arg)
        elif name == 'store_array':
            logger.store(**{k: np.asarray(v, dtype=np.float32) for k, v in arg.items()})
        elif name == 'log_episode':
            logger.log_episode(*arg)
        elif name == 'log_validation':
            logger.log_validation(*arg)
        elif name == 'log_penalty':
            logger.log_penalty(*arg)
        elif name == 'epoch':
            logger.log_tabular('Epoch', arg['Epoch'])
            logger.log_tabular('EpRet', min_and_max=True, std=True)
            logger.log_tabular('EpLen', min_and_max=True)
            logger.log_tabular('Values/V', min_and_max=True)
            logger.log_tabular('Values/Adv', min_and_max=True)
            logger.log_tabular('Loss/Pi', std=False)
            for k in ('Loss/Value', 'Loss/DeltaPi', 'Loss/DeltaValue', 'Entropy', 'KL', 'Misc/StopIter'):
                logger.log_tabular(k)
            logger.log_tabular('Misc/Seed', arg['Misc/Seed'])
            logger.log_tabular('PolicyRatio')
            for k in ('LR', 'Misc/RewScaleMean', 'Misc/RewScaleStddev', 'TotalEnvSteps', 'Time', 'FPS'):
                logger.log_tabular(k, arg[k])
            logger.dump_tabular()


def main():
    single_process_mpi()
    sys.path.insert(0, HERE)
    import ref_harness
    ref_harness.install()
    from phoenix_drone_simulation.utils.loggers import EpochLogger
    calls = call_sequence()
    tmp = tempfile.mkdtemp()
    lg = EpochLogger(log_dir=tmp, exp_name='golden', use_tensor_board=False, verbose=False)
    replay(lg, calls)
    lg.close()
    os.makedirs(OUT, exist_ok=True)
    for f in FILES:
        shutil.copy(os.path.join(tmp, f), os.path.join(OUT, f))
    with open(os.path.join(OUT, 'calls.json'), 'w') as f:
        json.dump(calls, f)
    print('wrote', OUT, sorted(os.listdir(OUT)))


if __name__ == '__main__':
    main()
