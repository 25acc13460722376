"""B200-native batched CrazyFlie hover environment (drop-in for the gym envs of
Hu-Hanyang/disturbance-CrazyFile-simulation) + fused PPO GAE.  All compute is hand-written
sm_100a CUDA behind the C ABI of ``include/cf_abi.h``; see DESIGN.md."""
from .vec_env import ENV_IDS, CrazyFlieVecEnv, GymAdapter, make, register_with_gym  # noqa: F401

__all__ = ['ENV_IDS', 'CrazyFlieVecEnv', 'GymAdapter', 'make', 'register_with_gym']
