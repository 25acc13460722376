"""Fused PPO GAE / return pass over a time-major rollout buffer (one CUDA kernel).

Replaces ``Buffer.finish_path`` + ``calculate_adv_and_value_targets`` + ``discount_cumsum``
(algs/core.py:498-535, 459-480, 105-119) for all envs and all paths of a rollout at once.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import abi


def gae(rew: torch.Tensor, val: torch.Tensor, done: torch.Tensor, end_val: Optional[torch.Tensor],
        last_val: torch.Tensor, gamma: float = 0.99, lam: float = 0.95, rew_scale: float = 0.0):
    """rew, val [T, N] float32; done [T, N] uint8 (1 = a path ends after this step); end_val [T, N]
    bootstrap value at path ends (0 for true terminals, V(s_T) for time-limit cuts, iwpg.py:392-402;
    None = all zeros); last_val [N] bootstrap of paths still open at T-1.  rew_scale > 0 applies the
    reference's reward scaling r * rew_scale clipped to +-10 inside the deltas
    (use_scaled_rewards, online_mean_std.py:32-48).  Returns (adv, target_v, disc_ret), each [T, N]."""
    assert rew.is_cuda and rew.dtype == torch.float32 and rew.dim() == 2
    T, N = rew.shape
    rew, val, last_val = rew.contiguous(), val.contiguous(), last_val.contiguous()
    done = done.to(torch.uint8).contiguous()
    assert val.shape == (T, N) and done.shape == (T, N) and last_val.shape == (N,)
    if end_val is not None:
        end_val = end_val.contiguous()
        assert end_val.shape == (T, N)
    adv, tgt, ret = torch.empty_like(rew), torch.empty_like(rew), torch.empty_like(rew)
    p = lambda t: None if t is None else C.c_void_p(t.data_ptr())
    with torch.cuda.device(rew.device):
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        abi.check(abi.lib().cf_gae_dev(p(rew), p(val), p(done), p(end_val), p(last_val), gamma, lam, rew_scale,
                                       T, N, p(adv), p(tgt), p(ret), stream))
    return adv, tgt, ret
