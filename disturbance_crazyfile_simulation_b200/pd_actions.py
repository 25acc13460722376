"""Synthetic "PID actions" source for throughput runs and closed-loop tests (one CUDA kernel).

An attitude -> rate -> motor-mix PD law on the env's own noisy observation rows, structured like the
firmware cascade the reference cites (envs/control.py:13-26 gains, :33-50 ``rpy_control_factors_to_PWM``
with QUAD_FORMATION_X mixing, :94-100 ``PWM.act``), expressed in action units.  With the default gains the
RandomHJ hover env runs to its 500-step TimeLimit in most episodes (mean length ~486 on the CPU oracle).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import abi

HOVER_ACTION = 2.0 / 1.8 - 1.0
# hover action, kz, kvz, k_attitude [1/s], k_rate [action per rad/s], k_yaw_rate
DEFAULT_GAINS = (HOVER_ACTION, 0.5, 0.3, 8.0, 0.05, 0.05)


def pd_actions(obs: torch.Tensor, act: torch.Tensor | None = None, gains=DEFAULT_GAINS) -> torch.Tensor:
    """obs [n, 34|42] float32 CUDA (as returned by ``CrazyFlieVecEnv.step``) -> act [n, 4] in [-1, 1]."""
    assert obs.is_cuda and obs.dtype == torch.float32 and obs.dim() == 2 and obs.is_contiguous()
    n, od = obs.shape
    if act is None:
        act = torch.empty((n, 4), dtype=torch.float32, device=obs.device)
    assert act.is_cuda and act.dtype == torch.float32 and act.shape == (n, 4) and act.is_contiguous()
    g = (C.c_float * 6)(*gains)
    with torch.cuda.device(obs.device):
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        abi.check(abi.lib().cf_pd_actions_dev(C.c_void_p(obs.data_ptr()), od, n, C.c_void_p(act.data_ptr()), g, stream))
    return act
