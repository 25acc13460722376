"""The rollout's networks as one CUDA kernel: ``ActorCritic.step`` of the reference
(algs/core.py:371-394) behind ``cf_policy_step_dev`` - observation standardisation, value MLP, policy MLP, Gaussian sample
and log-probability in ONE launch per control step, reading the observation rows the step kernel wrote and writing actions /
values / log-probabilities straight into the rollout buffers.  No CPU or eager fallback."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import abi


def _linears(seq):
    return [m for m in seq if isinstance(m, torch.nn.Linear)]


class FusedPolicy:
    """Binds the parameter tensors of an ``ActorCritic`` (trainer.ActorCritic or the reference's own class: same attribute
    names ``obs_oms.mean/std``, ``pi.net``, ``pi.log_std``, ``v.net``) to the kernel.  The kernel reads the tensors in place
    (torch's ``[out][in]`` layout), so optimiser steps are seen without any re-packing; call :meth:`rebind` if a parameter
    tensor is REPLACED rather than updated in place."""

    def __init__(self, env, ac):
        self.env, self.ac = env, ac
        self._L = abi.lib()
        self.rebind()

    def rebind(self) -> None:
        ac, dev = self.ac, self.env.device
        pi, v = _linears(ac.pi.net), _linears(ac.v.net)
        if len(pi) != 3 or len(v) != 3:
            raise ValueError('the fused kernel implements two hidden layers per network (the reference default)')
        if pi[0].out_features != pi[1].out_features or v[0].out_features != v[1].out_features:
            raise ValueError('hidden layers of one network must have equal width')
        od = self.env.obs_dim
        if pi[0].in_features != od or v[0].in_features != od or pi[2].out_features != 4 or v[2].out_features != 1:
            raise ValueError('network shapes do not match the env (obs_dim -> h -> h -> 4 / 1)')
        w = abi.CfPolicyWeights()
        keep = []

        def ptr(t: torch.Tensor):
            if t.device != dev or t.dtype != torch.float32 or not t.is_contiguous():
                raise ValueError('policy tensors must be contiguous float32 tensors on the env device')
            keep.append(t)
            return t.data_ptr()

        oms = getattr(ac, 'obs_oms', None)
        w.obs_mean = ptr(oms.mean.data) if oms is not None else None
        w.obs_std = ptr(oms.std.data) if oms is not None else None
        w.obs_eps = float(oms.eps) if oms is not None else 0.0
        for l in range(3):
            w.pi_w[l], w.pi_b[l] = ptr(pi[l].weight.data), ptr(pi[l].bias.data)
            w.v_w[l], w.v_b[l] = ptr(v[l].weight.data), ptr(v[l].bias.data)
        w.log_std = ptr(ac.pi.log_std.data)
        w.pi_hidden, w.v_hidden = pi[0].out_features, v[0].out_features
        self._w, self._keep = w, keep

    def step(self, obs: torch.Tensor, act: Optional[torch.Tensor] = None, val: Optional[torch.Tensor] = None,
             logp: Optional[torch.Tensor] = None, deterministic: bool = False):
        """(action [N,4], value [N], log_prob [N]) for raw observation rows ``obs`` [N, obs_dim]; outputs go into the given
        tensors (e.g. row ``t`` of the rollout buffers).  Call it before the ``env.step`` that consumes the actions: the
        exploration noise is keyed by the envs' event counters."""
        n, dev = self.env.num_envs, self.env.device
        if obs.shape != (n, self.env.obs_dim) or not obs.is_contiguous() or obs.dtype != torch.float32:
            raise ValueError('obs must be a contiguous float32 [num_envs, obs_dim] tensor')
        act = torch.empty((n, 4), dtype=torch.float32, device=dev) if act is None else act
        val = torch.empty((n,), dtype=torch.float32, device=dev) if val is None else val
        logp = torch.empty((n,), dtype=torch.float32, device=dev) if logp is None else logp
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        abi.check(self._L.cf_policy_step_dev(self.env._h, C.byref(self._w), C.c_void_p(obs.data_ptr()), 1 if deterministic else 0,
                                             C.c_void_p(act.data_ptr()), C.c_void_p(val.data_ptr()), C.c_void_p(logp.data_ptr()),
                                             stream))
        return act, val, logp

    # ------------------------------------------------------------------ fused policy-update step (cf_ppo_pi_grad_dev)
    def pi_grad(self, obs_norm: torch.Tensor, act: torch.Tensor, adv: torch.Tensor, logp_old: torch.Tensor, clip_ratio: float):
        """One iteration of ``update_policy_net`` up to the optimiser step (iwpg.py:448-459; ``compute_loss_pi`` of
        ppo.py:22-40 + ``backward``) in one kernel: fills ``.grad`` of the six ``pi.net`` tensors and returns
        (mean surrogate loss, mean ratio) as 0-d device tensors.  ``obs_norm`` are the STANDARDISED observations."""
        dev = self.env.device
        B = obs_norm.shape[0]
        for t, shape in ((obs_norm, (B, self.env.obs_dim)), (act, (B, 4)), (adv, (B,)), (logp_old, (B,))):
            if t.shape != shape or t.dtype != torch.float32 or not t.is_contiguous() or t.device != dev:
                raise ValueError(f'pi_grad: expected contiguous float32 {shape} on {dev}')
        lins = _linears(self.ac.pi.net)
        if getattr(self, '_ppo_ws', None) is None:
            self._ppo_ws = torch.empty(self._L.cf_ppo_workspace_floats(), dtype=torch.float32, device=dev)
            self._ppo_lr = torch.zeros(2, dtype=torch.float32, device=dev)
        for l in lins:
            for p in (l.weight, l.bias):
                if p.grad is None:
                    p.grad = torch.zeros_like(p)
        gw = (C.c_void_p * 3)(*[l.weight.grad.data_ptr() for l in lins])
        gb = (C.c_void_p * 3)(*[l.bias.grad.data_ptr() for l in lins])
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        abi.check(self._L.cf_ppo_pi_grad_dev(C.byref(self._w), self.env.obs_dim, C.c_void_p(obs_norm.data_ptr()),
                                             C.c_void_p(act.data_ptr()), C.c_void_p(adv.data_ptr()), C.c_void_p(logp_old.data_ptr()),
                                             B, float(clip_ratio), gw, gb, C.c_void_p(self._ppo_lr.data_ptr()),
                                             C.c_void_p(self._ppo_ws.data_ptr()), stream))
        return self._ppo_lr[0], self._ppo_lr[1]

    def v_grad(self, obs_norm: torch.Tensor, target_v: torch.Tensor, perm: Optional[torch.Tensor], start: int, count: int):
        """One mini-batch iteration of ``update_value_net`` up to the optimiser step (iwpg.py:478-510): fills ``.grad`` of the six
        ``v.net`` tensors for rows ``perm[start:start + count]`` of the epoch's batch (``perm`` None: rows start..start+count-1 need
        start == 0) and returns the mean squared error as a 0-d device tensor."""
        dev = self.env.device
        n = obs_norm.shape[0]
        if obs_norm.shape != (n, self.env.obs_dim) or obs_norm.dtype != torch.float32 or not obs_norm.is_contiguous() or obs_norm.device != dev:
            raise ValueError('v_grad: obs must be contiguous float32 [n, obs_dim] on the env device')
        if target_v.shape != (n,) or target_v.dtype != torch.float32 or not target_v.is_contiguous():
            raise ValueError('v_grad: target_v must be contiguous float32 [n]')
        if perm is None:
            if start != 0 or not 0 < count <= n:
                raise ValueError('v_grad without perm covers rows 0..count-1')
            idx = None
        else:
            if perm.dtype != torch.int64 or perm.dim() != 1 or not perm.is_contiguous() or start < 0 or count < 1 or start + count > perm.numel():
                raise ValueError('v_grad: perm must be a contiguous int64 vector covering start..start+count')
            idx = perm.data_ptr() + 8 * start
        lins = _linears(self.ac.v.net)
        if getattr(self, '_ppo_ws_v', None) is None:
            self._ppo_ws_v = torch.empty(self._L.cf_ppo_workspace_floats(), dtype=torch.float32, device=dev)
            self._ppo_lv = torch.zeros(2, dtype=torch.float32, device=dev)
        for l in lins:
            for p in (l.weight, l.bias):
                if p.grad is None:
                    p.grad = torch.zeros_like(p)
        gw = (C.c_void_p * 3)(*[l.weight.grad.data_ptr() for l in lins])
        gb = (C.c_void_p * 3)(*[l.bias.grad.data_ptr() for l in lins])
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        abi.check(self._L.cf_ppo_v_grad_dev(C.byref(self._w), self.env.obs_dim, C.c_void_p(obs_norm.data_ptr()),
                                            C.c_void_p(target_v.data_ptr()), C.c_void_p(idx), count, gw, gb,
                                            C.c_void_p(self._ppo_lv.data_ptr()), C.c_void_p(self._ppo_ws_v.data_ptr()), stream))
        return self._ppo_lv[0]
