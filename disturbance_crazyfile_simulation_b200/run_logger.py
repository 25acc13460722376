"""Run-directory log / wire formats of the reference trainer.

Writes the files the reference's ``play.py`` / plot scripts and checkpoints read, byte-compatible with
``utils/loggers.py`` of the reference:

* ``progress.csv``            one row per epoch, header = keys of the first row (loggers.py:318-330, 448-513)
* ``episodic_data.csv``       Episode, disturbance_level, Steps, Returns (loggers.py:257-265, 612-618)
* ``validation_results.csv``  Training_Epoch, disturbance_level, Steps, Returns, Costs (loggers.py:267-275, 620-627)
* ``penalty_data.csv``        header only unless ``log_penalty`` is used (loggers.py:277-284, 629-633)
* ``config.json`` / ``env_config.json``  ``json.dumps(..., separators=(',', ':\\t'), indent=4, sort_keys=True)``
  after ``ConverterJSON`` (loggers.py:76-120, 332-361)
* ``torch_save/model{itr}.pt``  the actor-critic ``state_dict`` (loggers.py:411-446)

Epoch statistics follow ``mpi_statistics_scalar`` (utils/mpi_tools.py:217-240): float32 samples, float32 sums,
population standard deviation; across ranks the sums are reduced with ``torch.distributed`` (NCCL or gloo) where
the reference uses MPI.  Only rank 0 touches the file system, like the reference's ``proc_id() == 0`` guards.
"""
from __future__ import annotations

import csv
import json
import os
import os.path as osp
from typing import Optional

import numpy as np

EPISODIC_HEADER = ['Episode', 'disturbance_level', 'Steps', 'Returns']
VALIDATION_HEADER = ['Training_Epoch', 'disturbance_level', 'Steps', 'Returns', 'Costs']
PENALTY_HEADER = ['Step', 'Penalty_crash', 'Penalty_rpy', 'Penalty_rpy_dot', 'Penalty_z', 'Penalty_velocity', 'Penalty']


def _is_json_serializable(v) -> bool:
    try:
        json.dumps(v)
        return True
    except Exception:
        return False


class _ConverterJSON:
    """loggers.py:76-112: anything JSON cannot hold becomes a name, a dict of its attributes, or ``str(obj)``."""

    def __init__(self):
        self.dumped = []

    def convert(self, obj, depth: int = 0):
        if depth > 5:
            return 'max_convert_depth_reached'
        if _is_json_serializable(obj):
            return obj
        if isinstance(obj, dict):
            return {self.convert(k): self.convert(v, depth + 1) for k, v in obj.items()}
        if isinstance(obj, tuple):
            return (self.convert(x) for x in obj)
        if isinstance(obj, list):
            return [self.convert(x) for x in obj]
        if hasattr(obj, '__name__') and 'lambda' not in obj.__name__:
            return self.convert(obj.__name__)
        if hasattr(obj, '__dict__') and obj.__dict__:
            if any(obj is d for d in self.dumped):
                return 'Already dumped elsewhere (stop recursion)'
            self.dumped.append(obj)
            return {str(obj): {self.convert(k): self.convert(v, depth + 1) for k, v in obj.__dict__.items()}}
        return str(obj)


def _reduce(values, op: str):
    """float32 all-reduce of a short list over the default process group (no-op without one)."""
    buf = np.asarray(values, dtype=np.float32).reshape(-1)
    try:
        import torch
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dev = 'cuda' if dist.get_backend() == 'nccl' else 'cpu'
            t = torch.from_numpy(buf.copy()).to(dev)
            dist.all_reduce(t, op={'sum': dist.ReduceOp.SUM, 'min': dist.ReduceOp.MIN, 'max': dist.ReduceOp.MAX}[op])
            buf = t.cpu().numpy()
    except ImportError:
        pass
    return buf


def statistics_scalar(x, with_min_and_max: bool = False):
    """mpi_statistics_scalar (utils/mpi_tools.py:217-240)."""
    x = np.array(x, dtype=np.float32)
    s, n = _reduce([np.sum(x), len(x)], 'sum')
    mean = s / n
    std = np.sqrt(_reduce([np.sum((x - mean) ** 2)], 'sum')[0] / n)
    if with_min_and_max:
        lo = _reduce([np.min(x) if len(x) > 0 else np.inf], 'min')[0]
        hi = _reduce([np.max(x) if len(x) > 0 else -np.inf], 'max')[0]
        return mean, std, lo, hi
    return mean, std


class RunLogger:
    """``EpochLogger`` of the reference (loggers.py:205-633) without console colouring and TensorBoard."""

    def __init__(self, log_dir: str, output_fname: str = 'progress.csv', exp_name: Optional[str] = None, rank: int = 0):
        self.log_dir, self.exp_name, self.rank = log_dir, exp_name, rank
        self.epoch, self.first_row = 0, True
        self.log_headers, self.log_current_row, self.epoch_dict = [], {}, {}
        self.output_file = None
        if rank == 0:
            os.makedirs(log_dir, exist_ok=True)
            self.output_file = open(osp.join(log_dir, output_fname), 'w')
            self.episodic_file = osp.join(log_dir, 'episodic_data.csv')
            self.validation_file = osp.join(log_dir, 'validation_results.csv')
            self.penalty_file = osp.join(log_dir, 'penalty_data.csv')
            for path, header in ((self.episodic_file, EPISODIC_HEADER), (self.validation_file, VALIDATION_HEADER),
                                 (self.penalty_file, PENALTY_HEADER)):
                with open(path, mode='w', newline='') as f:
                    csv.writer(f).writerow(header)
            self.validation_gifs = osp.join(log_dir, 'validation_gifs')
            os.makedirs(self.validation_gifs, exist_ok=True)

    def close(self) -> None:
        if self.output_file is not None:
            self.output_file.close()

    # ---- epoch table (progress.csv) -------------------------------------------------------------
    def store(self, **kwargs) -> None:
        for k, v in kwargs.items():
            self.epoch_dict.setdefault(k, []).append(v)

    def get_stats(self, key: str, with_min_and_max: bool = False):
        assert key in self.epoch_dict, f'key={key} not in dict'
        v = self.epoch_dict[key]
        vals = np.concatenate(v) if isinstance(v[0], np.ndarray) and len(v[0].shape) > 0 else v
        return statistics_scalar(vals, with_min_and_max)

    def _put(self, key: str, val) -> None:
        if self.first_row:
            self.log_headers.append(key)
        else:
            assert key in self.log_headers, f"Trying to introduce a new key {key} that you didn't include in the first iteration"
        assert key not in self.log_current_row, f'You already set {key} this iteration. Maybe you forgot to call dump_tabular()'
        self.log_current_row[key] = val

    def log_tabular(self, key: str, val=None, min_and_max: bool = False, std: bool = False) -> None:
        if val is not None:
            self._put(key, val)
        else:
            stats = self.get_stats(key, min_and_max)
            self._put(key + '/Mean' if (min_and_max or std) else key, stats[0])
            if std:
                self._put(key + '/Std', stats[1])
            if min_and_max:
                self._put(key + '/Min', stats[2])
                self._put(key + '/Max', stats[3])
        self.epoch_dict[key] = []

    def dump_tabular(self) -> None:
        if self.rank == 0:
            vals = list(self.log_current_row.values())
            self.epoch += 1
            if self.output_file is not None:
                if self.first_row:
                    self.output_file.write(','.join(self.log_headers) + '\n')
                self.output_file.write(','.join(map(str, vals)) + '\n')
                self.output_file.flush()
        self.log_current_row.clear()
        self.first_row = False

    # ---- per-episode / validation rows ---------------------------------------------------------
    def log_episode(self, num_episodes, distb_level, steps, returns) -> None:
        if self.rank == 0:
            with open(self.episodic_file, mode='a', newline='') as f:
                csv.writer(f).writerow([num_episodes, distb_level, steps, returns])

    def log_episodes(self, first_episode: int, distb_levels, steps, returns) -> None:
        """Batched form for the vectorised rollout: one row per finished episode, numbered from ``first_episode``."""
        if self.rank == 0:
            with open(self.episodic_file, mode='a', newline='') as f:
                csv.writer(f).writerows([first_episode + i, lv, st, rt]
                                        for i, (lv, st, rt) in enumerate(zip(distb_levels, steps, returns)))

    def log_validation(self, current_epoch, distb_level, steps, returns, costs) -> None:
        if self.rank == 0:
            with open(self.validation_file, mode='a', newline='') as f:
                csv.writer(f).writerows(list(zip(current_epoch, distb_level, steps, returns, costs)))

    def log_penalty(self, step, penalty_crash, penalty_rpy, penalty_rpy_dot, penalty_z, penalty_velocity, penalty) -> None:
        if self.rank == 0:
            with open(self.penalty_file, mode='a', newline='') as f:
                csv.writer(f).writerow([step, penalty_crash, penalty_rpy, penalty_rpy_dot, penalty_z, penalty_velocity, penalty])

    # ---- configuration and checkpoints ----------------------------------------------------------
    def save_dict_to_disk(self, dic: dict, file_name: str) -> None:
        assert isinstance(dic, dict), f'Got {type(dic)} but expected dict'
        if self.rank == 0:
            out = _ConverterJSON().convert(dic)
            if self.exp_name is not None:
                out['exp_name'] = self.exp_name
            with open(osp.join(self.log_dir, file_name), 'w') as f:
                f.write(json.dumps(out, separators=(',', ':\t'), indent=4, sort_keys=True))

    def save_config(self, config: dict, file_name: str = 'config.json') -> None:
        self.save_dict_to_disk(config, file_name)

    def save_env_config(self, env, file_name: str = 'env_config.json') -> None:
        """``env.unwrapped.__dict__`` of the reference; here the env's public configuration (``env.config_dict()`` when
        the env offers it, else its ``__dict__``)."""
        d = env.config_dict() if hasattr(env, 'config_dict') else dict(getattr(env, 'unwrapped', env).__dict__)
        self.save_dict_to_disk(d, file_name)

    def torch_save(self, state_dict: dict, itr: Optional[int] = None) -> Optional[str]:
        """``torch_save/model{itr}.pt`` holding a ``state_dict`` (the reference's final ``torch.save`` at loggers.py:444)."""
        if self.rank != 0:
            return None
        import torch
        fpath = osp.join(self.log_dir, 'torch_save')
        os.makedirs(fpath, exist_ok=True)
        fname = osp.join(fpath, 'model' + ('%d' % itr if itr is not None else '') + '.pt')
        torch.save(state_dict, fname)
        return fname
