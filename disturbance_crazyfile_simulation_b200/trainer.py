"""Batched on-device PPO around the B200 hover env (BASELINE.json configs[4]).

Shape of ``IWPGAlgorithm.learn_one_epoch`` (algs/iwpg/iwpg.py:282-307) with the PPO clip loss of
algs/ppo/ppo.py:22-40, re-hosted on the GPU: the rollout is ``T`` control steps of ``N`` envs per rank
(``cf_step_dev``), advantages come from the fused GAE kernel (``cf_gae_dev``), and ranks average
gradients / statistics over NCCL (``dist_utils``) where the reference uses MPI.  The two MLPs
(34->50->50->4 relu, 34->64->64->1 tanh: 10 969 parameters) are library GEMMs through torch, not the
product; ``state_dict`` keys equal the reference's (``pi.net.{0,2,4}``, ``pi.log_std``,
``v.net.{0,2,4}``, ``obs_oms.*``, ``ret_oms.*``) so its ``torch_save/model*.pt`` checkpoints load.
"""
from __future__ import annotations

import math
import time
from typing import Optional

import numpy as np
import torch
import torch.nn as nn

from . import dist_utils as du


class OnlineMeanStd(nn.Module):
    """utils/online_mean_std.py:7-95: running mean / std with the parallel (Chan et al.) update; the
    two MPI averages become two all-reduces over the ranks."""

    def __init__(self, epsilon: float = 1e-5, shape=()):
        super().__init__()
        self.mean = nn.Parameter(torch.zeros(*shape), requires_grad=False)
        self.std = nn.Parameter(torch.ones(*shape), requires_grad=False)
        self.count = nn.Parameter(torch.zeros(1), requires_grad=False)
        self.eps, self.bound, self.shape = epsilon, 10, shape

    def forward(self, x: torch.Tensor, subtract_mean: bool = True, clip: bool = False) -> torch.Tensor:
        y = (x - self.mean) / (self.std + self.eps) if subtract_mean else x / (self.std + self.eps)
        return torch.clamp(y, -self.bound, self.bound) if clip else y

    @torch.no_grad()
    def update(self, x: torch.Tensor) -> None:
        if self.shape[0] == 1:
            x = x.reshape(-1, 1)
        _, ws = du.world()
        n_B = float(x.shape[0] * ws)
        n_A = self.count.clone()
        n_AB = self.count + n_B
        batch_mean = du.mean_across_ranks(torch.mean(x, dim=0))
        delta = batch_mean - self.mean
        mean_new = self.mean + delta * n_B / n_AB
        batch_var = du.mean_across_ranks(torch.mean((x - mean_new) ** 2, dim=0))
        M2_AB = n_A * torch.square(self.std) + n_B * batch_var + delta ** 2 * (n_A * n_B / n_AB)
        # in place: the fused rollout kernel and captured graphs read these tensors by address
        self.mean.data.copy_(mean_new)
        self.count.data.copy_(n_AB)
        self.std.data.copy_(torch.sqrt(M2_AB / n_AB))


def _mlp(sizes, activation) -> nn.Sequential:
    """build_mlp_network (algs/core.py:84-98): [Linear, act] per layer, Identity after the last;
    kaiming_uniform weights (core.py:33-37)."""
    layers = []
    for j in range(len(sizes) - 1):
        lin = nn.Linear(sizes[j], sizes[j + 1])
        nn.init.kaiming_uniform_(lin.weight, a=math.sqrt(5))
        layers += [lin, activation() if j < len(sizes) - 2 else nn.Identity()]
    return nn.Sequential(*layers)


class GaussianActor(nn.Module):
    """MLPGaussianActor (algs/core.py:227-290)."""

    def __init__(self, obs_dim: int, act_dim: int, hidden=(50, 50)):
        super().__init__()
        self.log_std = nn.Parameter(torch.full((act_dim,), math.log(0.5)), requires_grad=False)
        self.net = _mlp([obs_dim] + list(hidden) + [act_dim], nn.ReLU)

    def set_log_std(self, frac: float) -> None:                 # core.py:269-276
        self.log_std.data.fill_(math.log(0.499 * frac + 0.01))

    def log_prob(self, mu: torch.Tensor, act: torch.Tensor) -> torch.Tensor:
        var = torch.exp(2.0 * self.log_std)
        return (-((act - mu) ** 2) / (2.0 * var) - self.log_std - 0.5 * math.log(2.0 * math.pi)).sum(-1)

    def entropy(self) -> torch.Tensor:
        """dist.entropy().mean() of the reference (iwpg.py:259): mean over the action dimensions, not their sum."""
        return (0.5 + 0.5 * math.log(2.0 * math.pi) + self.log_std).mean()


class Critic(nn.Module):
    """MLPCritic (algs/core.py:298-312)."""

    def __init__(self, obs_dim: int, hidden=(64, 64)):
        super().__init__()
        self.net = _mlp([obs_dim] + list(hidden) + [1], nn.Tanh)

    def forward(self, obs: torch.Tensor) -> torch.Tensor:
        return self.net(obs).squeeze(-1)


class ActorCritic(nn.Module):
    """ActorCritic (algs/core.py:315-412) with the PPO defaults of algs/ppo/defaults.py:6-19."""

    def __init__(self, obs_dim: int = 34, act_dim: int = 4, use_scaled_rewards: bool = True):
        super().__init__()
        self.obs_oms = OnlineMeanStd(shape=(obs_dim,))
        self.pi = GaussianActor(obs_dim, act_dim)
        self.v = Critic(obs_dim)
        self.ret_oms = OnlineMeanStd(shape=(1,)) if use_scaled_rewards else None

    @torch.no_grad()
    def step(self, obs: torch.Tensor, deterministic: bool = False):
        """(action, value, log_prob) for raw observations (core.py:371-394)."""
        o = self.obs_oms(obs)
        v = self.v(o)
        mu = self.pi.net(o)
        if deterministic:
            return mu, v, torch.ones_like(v)
        a = mu + torch.exp(self.pi.log_std) * torch.randn_like(mu)
        return a, v, self.pi.log_prob(mu, a)

    @torch.no_grad()
    def value(self, obs: torch.Tensor) -> torch.Tensor:
        return self.v(self.obs_oms(obs))


class PPOTrainer:
    """``ProximalPolicyOptimizationAlgorithm`` (algs/ppo/ppo.py) over a batched env.

    One epoch = ``T`` control steps of the rank's ``N`` envs (``steps_per_epoch = T * N * world``), then
    ``Buffer.finish_path`` for every path at once (fused GAE kernel), ``train_v_iterations`` x
    ``num_mini_batches`` value updates, ``train_pi_iterations`` full-batch policy updates, running
    statistics.  Hyper-parameter names and defaults follow iwpg.py:27-66 / ppo/defaults.py.
    """

    def __init__(self, env, rollout_len: int = 32, epochs: int = 300, gamma: float = 0.99, lam: float = 0.95,
                 clip_ratio: float = 0.2, pi_lr: float = 3e-4, vf_lr: float = 1e-3, train_pi_iterations: int = 80,
                 train_v_iterations: int = 5, num_mini_batches: int = 16, entropy_coef: float = 0.0,
                 use_reward_scaling: bool = True, use_linear_lr_decay: bool = True,
                 use_exploration_noise_anneal: bool = True, use_kl_early_stopping: bool = False,
                 target_kl: float = 0.01, check_freq: int = 25, seed: int = 0, actor_critic: Optional[ActorCritic] = None,
                 gae_fn=None, logger=None, save_freq: int = 10, val_freq: int = 0, val_kwargs: Optional[dict] = None,
                 fused_rollout: Optional[bool] = None, use_cuda_graphs: Optional[bool] = None, episode_rows_per_epoch: int = 256,
                 fused_update: Optional[bool] = None):
        # fused_rollout (default: on for a CUDA CrazyFlieVecEnv): the two networks of ac.step run as ONE kernel per control step
        # (policy.FusedPolicy / cf_policy_step_dev) that writes straight into the rollout buffers, the env writes its rows
        # straight into them too (no copies), and the T x 2 launches of a rollout are replayed from one CUDA graph;
        # use_cuda_graphs also captures one optimiser step of each network (forward, loss, backward, gradient all-reduce,
        # Adam) and replays it train_pi_iterations / train_v_iterations x num_mini_batches times
        # logger: a run_logger.RunLogger (progress.csv / config.json / torch_save/model{epoch}.pt in the reference's
        # formats, loggers.py:205-513); save_freq as iwpg.py:45,302-304
        # val_freq > 0: validate() every val_freq epochs (iwpg.py:66,297-301; the reference's default is 50); val_kwargs are
        # passed to validate() (num_evaluations, camera, frame_stride, ...)
        self.val_freq, self.val_kwargs, self._val_env = int(val_freq), dict(val_kwargs or {}), None
        self.episode_rows_per_epoch, self._episodes_logged = int(episode_rows_per_epoch), 0
        self.logger, self.save_freq, self.seed, self._t_start = logger, int(save_freq), seed, time.time()
        if gae_fn is None:                                   # the fused CUDA kernel; tests of the host logic inject a stand-in
            from .gae import gae as gae_fn
        self._gae = gae_fn
        self.env, self.T, self.N = env, int(rollout_len), env.num_envs
        self.dev = env.device
        self.rank, self.world = du.world()
        self.epochs, self.gamma, self.lam, self.clip_ratio = epochs, gamma, lam, clip_ratio
        self.train_pi_iterations, self.train_v_iterations = train_pi_iterations, train_v_iterations
        self.num_mini_batches, self.entropy_coef = num_mini_batches, entropy_coef
        self.use_reward_scaling, self.use_exploration_noise_anneal = use_reward_scaling, use_exploration_noise_anneal
        self.use_kl_early_stopping, self.target_kl, self.check_freq = use_kl_early_stopping, target_kl, check_freq
        self.pi_lr = pi_lr
        torch.manual_seed(seed + 10000 * self.rank)             # iwpg.py:135-137: seed += 10000 * proc_id()
        self.ac = (actor_critic or ActorCritic(env.obs_dim, 4, use_reward_scaling)).to(self.dev)
        du.broadcast_params(list(self.ac.parameters()))          # mpi_tools.sync_params, iwpg.py:222
        self.pi_optimizer = torch.optim.Adam(self.ac.pi.net.parameters(), lr=pi_lr)
        self.vf_optimizer = torch.optim.Adam(self.ac.v.parameters(), lr=vf_lr)
        self.scheduler = (torch.optim.lr_scheduler.LambdaLR(self.pi_optimizer, lambda e: 1 - e / self.epochs)
                          if use_linear_lr_decay else None)
        self._avg_pi = du.FlatGradAverager(list(self.ac.pi.net.parameters()))
        self._avg_v = du.FlatGradAverager(list(self.ac.v.parameters()))
        T, N, od = self.T, self.N, env.obs_dim
        f = dict(dtype=torch.float32, device=self.dev)
        on_gpu_env = self.dev.type == 'cuda' and hasattr(env, 'step_into') and hasattr(env, '_h')
        self.fused = on_gpu_env if fused_rollout is None else bool(fused_rollout)
        self.use_graphs = (self.dev.type == 'cuda') if use_cuda_graphs is None else bool(use_cuda_graphs)
        if self.fused and not on_gpu_env:
            raise ValueError('fused_rollout needs a CUDA CrazyFlieVecEnv')
        # fused_update (default: with the fused rollout and CUDA graphs): every optimiser step of either network runs forward + loss +
        # backward as one kernel (cf_ppo_pi_grad_dev / cf_ppo_v_grad_dev) instead of ~40 torch kernels; the gradient all-reduce
        # and Adam stay in torch
        self.fused_update = (self.fused and self.use_graphs) if fused_update is None else bool(fused_update)
        if self.fused_update and not (self.fused and self.use_graphs):
            raise ValueError('fused_update needs fused_rollout and use_cuda_graphs')
        self._obs_all = torch.empty((T + 1, N, od), **f)     # row t: observation the policy sees at step t; row T: after the last step
        self.obs_buf = self._obs_all[:T]
        self.act_buf = torch.empty((T, N, 4), **f)
        self.rew_buf, self.val_buf, self.logp_buf = torch.empty((T, N), **f), torch.empty((T, N), **f), torch.empty((T, N), **f)
        self.end_val_buf = torch.zeros((T, N), **f)
        self.done_buf = torch.empty((T, N), dtype=torch.uint8, device=self.dev)
        self.epoch = 0
        self.total_env_steps = 0
        self._obs = None
        self._fp = self._rollout_graph = self._pi_graph = self._v_graph = None
        if self.fused:
            from .policy import FusedPolicy
            self._fp = FusedPolicy(env, self.ac)
            self.term_buf = torch.zeros((T, N, od), **f)     # last observation of the envs that finished at step t
            self.trunc_buf = torch.zeros((T, N), dtype=torch.uint8, device=self.dev)
            self._scratch_act = torch.empty((N, 4), **f)
            self._scratch_logp = torch.empty((N,), **f)
            self.last_val = torch.empty((N,), **f)
        if self.use_graphs:
            # Adam with device-side step counters and a tensor learning rate, so that one captured step can be replayed
            self.pi_optimizer = torch.optim.Adam(self.ac.pi.net.parameters(), lr=torch.tensor(pi_lr, device=self.dev), capturable=True)
            self.vf_optimizer = torch.optim.Adam(self.ac.v.parameters(), lr=torch.tensor(vf_lr, device=self.dev), capturable=True)
            self.scheduler = (torch.optim.lr_scheduler.LambdaLR(self.pi_optimizer, lambda e: 1 - e / self.epochs)
                              if use_linear_lr_decay else None)
        if getattr(env, '_term', None) is None:
            raise ValueError('PPOTrainer needs make(..., return_terminal_obs=True): time-limit cuts bootstrap from V(s_T)')
        if self.logger is not None:                          # logger.save_config(self.params), iwpg.py:126
            self.logger.save_config(dict(alg='ppo', env_id=getattr(env, 'id', None), num_envs=self.N, world_size=self.world,
                                         rollout_len=self.T, epochs=epochs, gamma=gamma, lam=lam, clip_ratio=clip_ratio,
                                         pi_lr=pi_lr, vf_lr=vf_lr, train_pi_iterations=train_pi_iterations,
                                         train_v_iterations=train_v_iterations, num_mini_batches=num_mini_batches,
                                         entropy_coef=entropy_coef, use_reward_scaling=use_reward_scaling,
                                         use_linear_lr_decay=use_linear_lr_decay,
                                         use_exploration_noise_anneal=use_exploration_noise_anneal,
                                         use_kl_early_stopping=use_kl_early_stopping, target_kl=target_kl, seed=seed,
                                         save_freq=save_freq))

    # ------------------------------------------------------------------ rollout (iwpg.py:372-410)
    def _rollout_launches(self) -> None:
        """The 2 T launches of one rollout: policy kernel (obs row t -> action / value / log-prob row t), step kernel (action
        row t -> obs row t + 1, reward / done / truncated row t, terminal observations row t)."""
        env, fp = self.env, self._fp
        for t in range(self.T):
            fp.step(self._obs_all[t], self.act_buf[t], self.val_buf[t], self.logp_buf[t])
            env.step_into(self.act_buf[t], self._obs_all[t + 1], self.rew_buf[t], self.done_buf[t], None, self.term_buf[t],
                          self.trunc_buf[t])
        fp.step(self._obs_all[self.T], self._scratch_act, self.last_val, self._scratch_logp, deterministic=True)

    @torch.no_grad()
    def roll_out_fused(self) -> None:
        """roll_out with the networks fused into one kernel per step and no copies; from the second epoch on the whole rollout
        is one CUDA-graph replay.  Paths that end by the time limit bootstrap from V(terminal observation) (iwpg.py:392-397):
        those rows are rare (an env reaches 500 steps), so they are gathered after the rollout instead of evaluating the value
        net on every terminal observation of every step."""
        if self._obs is None:
            self._obs_all[0].copy_(self.env.reset())
            self._obs = self._obs_all[0]
        else:
            self._obs_all[0].copy_(self._obs_all[self.T])   # paths continue across epochs
        if self.use_graphs and self.epoch >= 1:
            if self._rollout_graph is None:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self._rollout_launches()
                self._rollout_graph = g
            self._rollout_graph.replay()
        else:
            self._rollout_launches()
        idx = self.trunc_buf.nonzero(as_tuple=True)
        self.end_val_buf.zero_()
        if idx[0].numel():
            self.end_val_buf[idx] = self.ac.value(self.term_buf[idx])
        self.total_env_steps += self.T * self.N * self.world

    @torch.no_grad()
    def roll_out(self) -> None:
        if self.fused:
            return self.roll_out_fused()
        env, ac = self.env, self.ac
        if self._obs is None:
            self._obs = env.reset()
        o = self._obs
        for t in range(self.T):
            a, v, logp = ac.step(o)
            self.obs_buf[t].copy_(o)
            self.act_buf[t].copy_(a)
            self.val_buf[t].copy_(v)
            self.logp_buf[t].copy_(logp)
            o, r, d, info = env.step(a)                     # done envs are already reset: o holds their first obs
            self.rew_buf[t].copy_(r)
            self.done_buf[t].copy_(d)
            # path end: v = 0 on a true terminal, V(s_T) on a time-limit cut (iwpg.py:392-397)
            v_term = ac.value(info['terminal_observation'])
            self.end_val_buf[t].copy_(torch.where(info['truncated'], v_term, torch.zeros_like(v_term)))
        self._obs = o
        self.last_val = ac.value(o)                         # epoch_ended: bootstrap the open paths (iwpg.py:391-394)
        self.total_env_steps += self.T * self.N * self.world

    # ------------------------------------------------------------------ update (iwpg.py:423-510)
    def compute_loss_pi(self, obs, act, adv, logp_old):
        mu = self.ac.pi.net(obs)
        logp = self.ac.pi.log_prob(mu, act)
        ratio = torch.exp(logp - logp_old)
        clip_adv = adv * torch.clamp(ratio, 1 - self.clip_ratio, 1 + self.clip_ratio)
        loss = -(torch.min(ratio * adv, clip_adv)).mean() - self.entropy_coef * self.ac.pi.entropy()
        return loss, mu, ratio

    def update(self) -> dict:
        T, N = self.T, self.N
        rew_scale = 0.0
        if self.use_reward_scaling:                          # ret_oms(rews, subtract_mean=False, clip=True), core.py:519-527
            rew_scale = 1.0 / (float(self.ac.ret_oms.std.item()) + self.ac.ret_oms.eps)
        adv, target_v, disc_ret = self._gae(self.rew_buf, self.val_buf, self.done_buf, self.end_val_buf, self.last_val,
                                            self.gamma, self.lam, rew_scale)
        raw_obs = self.obs_buf.reshape(T * N, -1)
        obs = self.ac.obs_oms(raw_obs).detach()              # pre_process_data, iwpg.py:350-370
        act, adv = self.act_buf.reshape(T * N, 4), adv.reshape(T * N)
        target_v, logp_old = target_v.reshape(T * N), self.logp_buf.reshape(T * N)
        info = {}
        # value net: train_v_iterations epochs of num_mini_batches shuffled mini-batches (iwpg.py:478-510)
        B = T * N
        mbs = B // self.num_mini_batches
        with torch.no_grad():
            info['loss_v_before'] = float(((self.ac.v(obs) - target_v) ** 2).mean())
        if self.use_graphs:
            return self._update_graphed(info, raw_obs, obs, act, adv, target_v, logp_old, disc_ret, B, mbs)
        for _ in range(self.train_v_iterations):
            perm = torch.randperm(B, device=self.dev)
            for start in range(0, mbs * self.num_mini_batches, mbs):
                idx = perm[start:start + mbs]
                self.vf_optimizer.zero_grad(set_to_none=False)
                loss_v = ((self.ac.v(obs[idx]) - target_v[idx]) ** 2).mean()
                loss_v.backward()
                self._avg_v()                                # mpi_avg_grads(self.ac.v), iwpg.py:504
                self.vf_optimizer.step()
        # policy net: full-batch steps (iwpg.py:436-476)
        with torch.no_grad():
            mu_old = self.ac.pi.net(obs)
        stop_iter = self.train_pi_iterations
        for i in range(self.train_pi_iterations):
            self.pi_optimizer.zero_grad(set_to_none=False)
            loss_pi, mu, ratio = self.compute_loss_pi(obs, act, adv, logp_old)
            if i == 0:
                info['loss_pi_before'] = float(loss_pi.detach())
            loss_pi.backward()
            self._avg_pi()                                   # mpi_avg_grads(self.ac.pi.net), iwpg.py:458
            self.pi_optimizer.step()
            if self.use_kl_early_stopping:
                if self._kl(obs, mu_old) > self.target_kl:
                    stop_iter = i + 1
                    break
        info.update(loss_pi=float(loss_pi.detach()), stop_iter=stop_iter, policy_ratio=float(ratio.detach().mean()))
        # running statistics on the RAW data (iwpg.py:412-421)
        self.ac.obs_oms.update(raw_obs)
        if self.use_reward_scaling:
            self.ac.ret_oms.update(disc_ret.reshape(-1))
        return info

    def _kl(self, obs, mu_old) -> float:
        """kl_divergence(p_old, p_new).mean() over all [B, act_dim] elements (iwpg.py:461-466); equal std on both sides."""
        with torch.no_grad():
            kl = (((self.ac.pi.net(obs) - mu_old) ** 2) / (2.0 * torch.exp(2.0 * self.ac.pi.log_std))).mean()
        return float(du.mean_across_ranks(kl.reshape(1)))

    def _update_graphed(self, info, raw_obs, obs, act, adv, target_v, logp_old, disc_ret, B, mbs) -> dict:
        """The same update with ONE optimiser step of each network captured in a CUDA graph (forward, loss, backward, flattened
        gradient all-reduce, Adam with device-side step count) and replayed: train_v_iterations x (one graph of num_mini_batches
        mini-batch steps over a fresh permutation), then train_pi_iterations x the full-batch policy step."""
        if self._pi_graph is None:
            # static inputs of the captured steps (the buffers the epoch's data is copied into)
            self._g_obs = torch.empty_like(obs); self._g_act = torch.empty_like(act); self._g_adv = torch.empty_like(adv)
            self._g_tgt = torch.empty_like(target_v); self._g_logp = torch.empty_like(logp_old)
            self._g_perm = torch.arange(B, device=self.dev)      # refilled from torch.randperm before every replay
        for dst, src in ((self._g_obs, obs), (self._g_act, act), (self._g_adv, adv), (self._g_tgt, target_v), (self._g_logp, logp_old)):
            dst.copy_(src)
        if self._pi_graph is None:
            side = torch.cuda.Stream(self.dev)
            side.wait_stream(torch.cuda.current_stream(self.dev))
            with torch.cuda.stream(side):                    # warm-up outside capture (allocator, cuBLAS handles, Adam state)
                params = [p.detach().clone() for p in self.ac.parameters()]
                for _ in range(2):
                    self._v_minibatches(mbs)
                    self._pi_step()
                for p, q in zip(self.ac.parameters(), params):
                    p.data.copy_(q)
                for opt in (self.pi_optimizer, self.vf_optimizer):      # forget the warm-up steps
                    for st in opt.state.values():
                        for v in st.values():
                            if torch.is_tensor(v):
                                v.zero_()
            torch.cuda.current_stream(self.dev).wait_stream(side)
            self._v_graph, self._pi_graph = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._v_graph):
                self._v_minibatches(mbs)
            with torch.cuda.graph(self._pi_graph):
                self._g_loss_pi, self._g_ratio = self._pi_step()
        for _ in range(self.train_v_iterations):
            self._g_perm.copy_(torch.randperm(B, device=self.dev))      # same generator use as the eager update
            self._v_graph.replay()
        with torch.no_grad():
            mu_old = self.ac.pi.net(obs) if self.use_kl_early_stopping else None
        stop_iter = self.train_pi_iterations
        for i in range(self.train_pi_iterations):
            self._pi_graph.replay()
            if i == 0:
                info['loss_pi_before'] = float(self._g_loss_pi)
            if self.use_kl_early_stopping and self._kl(obs, mu_old) > self.target_kl:
                stop_iter = i + 1
                break
        info.update(loss_pi=float(self._g_loss_pi), stop_iter=stop_iter, policy_ratio=float(self._g_ratio))
        self.ac.obs_oms.update(raw_obs)
        if self.use_reward_scaling:
            self.ac.ret_oms.update(disc_ret.reshape(-1))
        return info

    def _v_minibatches(self, mbs: int) -> None:
        for start in range(0, mbs * self.num_mini_batches, mbs):
            if self.fused_update:                            # forward + loss + backward of v.net in one kernel (cf_ppo_v_grad_dev)
                self._fp.v_grad(self._g_obs, self._g_tgt, self._g_perm, start, mbs)
                self._avg_v()
                self.vf_optimizer.step()
                continue
            idx = self._g_perm[start:start + mbs]
            self.vf_optimizer.zero_grad(set_to_none=False)
            loss_v = ((self.ac.v(self._g_obs[idx]) - self._g_tgt[idx]) ** 2).mean()
            loss_v.backward()
            self._avg_v()
            self.vf_optimizer.step()

    def _pi_step(self):
        if self.fused_update:
            # forward + clipped surrogate + backward of pi.net in one kernel (cf_ppo_pi_grad_dev): .grad is written directly
            loss, ratio = self._fp.pi_grad(self._g_obs, self._g_act, self._g_adv, self._g_logp, self.clip_ratio)
            self._avg_pi()
            self.pi_optimizer.step()
            return loss - self.entropy_coef * self.ac.pi.entropy().detach(), ratio
        self.pi_optimizer.zero_grad(set_to_none=False)
        loss_pi, _, ratio = self.compute_loss_pi(self._g_obs, self._g_act, self._g_adv, self._g_logp)
        loss_pi.backward()
        self._avg_pi()
        self.pi_optimizer.step()
        return loss_pi.detach(), ratio.detach().mean()

    def _sync(self) -> None:
        if self.dev.type == 'cuda':
            torch.cuda.synchronize(self.dev)

    # ------------------------------------------------------------------ epoch (iwpg.py:282-307)
    def learn_one_epoch(self) -> dict:
        t0 = time.time()
        if self.use_exploration_noise_anneal:
            self.ac.pi.set_log_std(1.0 - self.epoch / self.epochs)     # ac.update(frac), core.py:396-412
        self.env.episode_stats(reset=True)
        self.roll_out()
        self._sync()
        t1 = time.time()
        info = self.update()
        self._sync()
        t2 = time.time()
        st = self.env.episode_stats()
        red = du.reduce_episode_stats(st['episodes'], st['return_sum'], st['length_sum'], float('inf'), float('-inf'), self.dev)
        current_lr = self.scheduler.get_last_lr()[0] if self.scheduler is not None else self.pi_lr
        if self.scheduler is not None:
            self.scheduler.step()
        if self.epoch % self.check_freq == 0 and not du.check_distributed_parameters(self.ac):
            raise RuntimeError('actor-critic parameters differ across ranks (iwpg.py:242-251)')
        self.epoch += 1
        info.update(epoch=self.epoch, EpRet=red['mean_return'], EpLen=red['mean_length'], episodes=red['episodes'],
                    TotalEnvSteps=self.total_env_steps, rollout_s=t1 - t0, update_s=t2 - t1,
                    FPS=self.T * self.N * self.world / (t2 - t0))
        if self.logger is not None:
            self._log_epoch(info, current_lr)
        if self.val_freq > 0 and (self.epoch - 1) % self.val_freq == 0:      # epochs 0, val_freq, ... (iwpg.py:297-301)
            info['validation'] = self.validate(**self.val_kwargs)
        return info

    # ------------------------------------------------------------------ validation (iwpg.py:512-568)
    @torch.no_grad()
    def validate(self, num_evaluations: int = 8, capture: bool = True, camera=None, frame_stride: int = 1,
                 gif_duration_ms: int = 500, max_steps: Optional[int] = None, log_penalties: bool = True,
                 epoch_label: Optional[int] = None):
        """``IWPGAlgorithm.validate`` / ``val_once``: ``num_evaluations`` episodes of the deterministic policy, one row each
        in ``validation_results.csv`` and, with ``capture``, one ``{epoch}-{episode}-{level}.gif`` each under
        ``validation_gifs/`` holding a ``capture_image()`` frame per control step.  The reference runs the episodes one after
        the other on its single env; here they run side by side on a ``num_evaluations``-env twin of the training env
        (same id, same knobs, same HJ tables, no auto-reset) and every step's frames come from one rasteriser launch.
        ``camera`` (raster.make_camera) sets the frame size; default is the reference's 1920x1080.  Rank 0 only, no
        collective.  ``log_penalties``: the per-step penalty terms the reference writes to ``penalty_data.csv``
        (iwpg.py:531-532) are recomputed on the host from the twin env's state after every step (vec_env.penalty_terms)."""
        if self.rank != 0:
            return None
        # Training_Epoch column / GIF names: the reference's 0-based self.num_epoch of the epoch just finished (iwpg.py:545,562)
        label = max(self.epoch - 1, 0) if epoch_label is None else int(epoch_label)
        env, k = self.env, int(num_evaluations)
        if self._val_env is None or self._val_env.num_envs != k:
            from .vec_env import CrazyFlieVecEnv
            kw = dict(getattr(env, '_kwargs', {}))
            kw['auto_reset'] = 0
            self._val_env = CrazyFlieVecEnv(env.id, num_envs=k, device=env.device_index, seed=self.seed + 7919,
                                            env_id_base=1 << 40, share_tables_with=env, **kw)
        venv = self._val_env
        was_training = self.ac.training
        self.ac.eval()
        o = venv.reset()
        levels = venv.disturbance_level
        alive = torch.ones(k, dtype=torch.bool, device=self.dev)
        ret = torch.zeros(k, device=self.dev)
        cost = torch.zeros(k, device=self.dev)
        length = torch.zeros(k, dtype=torch.int32, device=self.dev)
        frames = [[] for _ in range(k)]
        pen_rows = [[] for _ in range(k)]
        limit = int(max_steps or venv._max_episode_steps)
        for t in range(limit):
            if capture and t % frame_stride == 0:
                img = venv.capture_images(None, camera=camera).cpu().numpy()        # [k, H, W, 3], one launch
                live = alive.cpu().numpy()
                for i in range(k):
                    if live[i]:
                        frames[i].append(img[i].copy())      # a view would pin the whole [k, H, W, 3] batch
            a, _, _ = self.ac.step(o, deterministic=True)
            o, r, d, info = venv.step(a)
            if log_penalties and self.logger is not None:    # logger.log_penalty(...) per step, iwpg.py:531-532
                pen = venv.penalty_terms(a.cpu().numpy(), (d & ~info['truncated']).cpu().numpy())
                live = alive.cpu().numpy()
                for i in range(k):
                    if live[i]:
                        pen_rows[i].append((t + 1, pen['penalty_crash'][i], pen['penalty_rpy'][i], pen['penalty_rpy_dot'][i],
                                            pen['penalty_z'][i], pen['penalty_velocity'][i], pen['penalty'][i]))
            ret += torch.where(alive, r, torch.zeros_like(r))
            cost += torch.where(alive, info['cost'], torch.zeros_like(r))
            length += alive.to(torch.int32)
            alive &= ~d.bool()
            if not bool(alive.any()):
                break
        self.ac.train(was_training)
        out = dict(epoch=label, disturbance_level=[float(x) for x in levels], steps=[float(x) for x in length.cpu()],
                   returns=[float(x) for x in ret.cpu()], costs=[float(x) for x in cost.cpu()], gifs=[])
        if self.logger is not None:
            self.logger.log_validation([label] * k, out['disturbance_level'], out['steps'], out['returns'], out['costs'])
            for rows in pen_rows:                            # episode after episode, like the reference's sequential val_once
                for row in rows:
                    self.logger.log_penalty(*row)
            if capture:
                from PIL import Image
                for i in range(k):
                    if not frames[i]:
                        continue
                    ims = [Image.fromarray(f) for f in frames[i]]
                    path = f"{self.logger.validation_gifs}/{label}-{i}-{out['disturbance_level'][i]}.gif"
                    ims[0].save(path, save_all=True, append_images=ims[1:], duration=gif_duration_ms, loop=0)
                    out['gifs'].append(path)
        return out

    def _log_epoch(self, info: dict, current_lr: float) -> None:
        """One progress.csv row with the column names of IWPGAlgorithm.log (iwpg.py:309-348) for the quantities the
        batched trainer tracks, then the periodic checkpoint (iwpg.py:302-304).  Episode statistics are the means the
        step kernel accumulates on the device; episodic_data.csv gets a bounded sample of per-episode rows per epoch."""
        lg = self.logger
        if self.episode_rows_per_epoch > 0 and hasattr(self.env, 'last_episodes') and self.rank == 0:
            # episodic_data.csv (loggers.py:612-618; iwpg.py:407-408 writes one row per finished training episode): a bounded
            # sample per epoch - the last finished episode of the envs that ended one during this rollout
            ended = self.done_buf.any(dim=0).nonzero().flatten()[: self.episode_rows_per_epoch].cpu().numpy()
            if ended.size:
                ret, ln, lv = self.env.last_episodes()
                lg.log_episodes(self._episodes_logged + 1, [float(round(x, 1)) for x in lv[ended]], [int(x) for x in ln[ended]],
                                [float(x) for x in ret[ended]])
                self._episodes_logged += int(ended.size)
        lg.log_tabular('Epoch', info['epoch'])
        lg.log_tabular('EpRet/Mean', float(info['EpRet']))
        lg.log_tabular('EpLen/Mean', float(info['EpLen']))
        lg.log_tabular('Loss/Pi', info['loss_pi_before'])
        lg.log_tabular('Loss/Value', info['loss_v_before'])
        lg.log_tabular('Loss/DeltaPi', info['loss_pi'] - info['loss_pi_before'])
        lg.log_tabular('Misc/StopIter', info['stop_iter'])
        lg.log_tabular('Misc/Seed', self.seed)
        lg.log_tabular('PolicyRatio', info['policy_ratio'])
        lg.log_tabular('LR', current_lr)
        if self.use_reward_scaling:
            lg.log_tabular('Misc/RewScaleMean', float(self.ac.ret_oms.mean.item()))
            lg.log_tabular('Misc/RewScaleStddev', float(self.ac.ret_oms.std.item()))
        if self.use_exploration_noise_anneal:
            lg.log_tabular('Misc/ExplorationNoiseStd', float(torch.exp(self.ac.pi.log_std[0]).item()))
        lg.log_tabular('TotalEnvSteps', info['TotalEnvSteps'])
        lg.log_tabular('Time', int(time.time() - self._t_start))
        lg.log_tabular('FPS', int(info['FPS']))
        lg.dump_tabular()
        ep0 = self.epoch - 1                                  # 0-based index of the epoch just finished, the reference's self.epoch
        if ep0 % self.save_freq == 0 or ep0 == self.epochs - 1:   # model0, model{save_freq}, ..., model{epochs-1} (iwpg.py:302-304)
            lg.torch_save(self.ac.state_dict(), itr=ep0)

    def learn(self, epochs: Optional[int] = None, log=None):
        for _ in range(epochs or self.epochs):
            info = self.learn_one_epoch()
            if log and self.rank == 0:
                log(info)
        return self.ac, self.env

    # ------------------------------------------------------------------ checkpoints (loggers.py:371-446)
    def save(self, path: str) -> None:
        if self.rank == 0:
            torch.save(self.ac.state_dict(), path)

    def load(self, path: str) -> None:
        sd = torch.load(path, map_location=self.dev)
        self.ac.load_state_dict(sd, strict=False)
