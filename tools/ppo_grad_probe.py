"""Times cf_ppo_pi_grad_dev alone on a random batch (default 65536 x 32 samples, the PPO bench's batch); the ncu target."""
import sys
import torch
from disturbance_crazyfile_simulation_b200 import make
from disturbance_crazyfile_simulation_b200.policy import FusedPolicy
from disturbance_crazyfile_simulation_b200.trainer import ActorCritic

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536 * 32
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
env = make('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', num_envs=128, seed=1, hj_tables='synthetic')
torch.manual_seed(0)
ac = ActorCritic(env.obs_dim, 4).to(env.device)
fp = FusedPolicy(env, ac)
obs = torch.randn(B, env.obs_dim, device=env.device)
with torch.no_grad():
    mu = ac.pi.net(obs)
    act = mu + torch.exp(ac.pi.log_std) * torch.randn(B, 4, device=env.device)
    logp = ac.pi.log_prob(mu, act) + 0.1 * torch.randn(B, device=env.device)
adv = torch.randn(B, device=env.device)
for _ in range(2):
    fp.pi_grad(obs, act, adv, logp, 0.2)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    fp.pi_grad(obs, act, adv, logp, 0.2)
e1.record(); torch.cuda.synchronize()
print(f'B={B} pi_grad {e0.elapsed_time(e1) / reps:.4f} ms per launch pair')
env.close()
