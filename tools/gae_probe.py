"""Times cf_gae_dev alone: T = 32, N = 524 288 (352 MB of rollout buffers, larger than L2), CUDA events over 20 launches."""
import sys
import torch
from disturbance_crazyfile_simulation_b200.gae import gae

T, N = 32, int(sys.argv[1]) if len(sys.argv) > 1 else 524288
dev = torch.device('cuda', 0)
g = torch.Generator(device=dev).manual_seed(0)
rew = torch.randn(T, N, device=dev, generator=g)
val = torch.randn(T, N, device=dev, generator=g)
done = (torch.rand(T, N, device=dev, generator=g) < 0.01).to(torch.uint8)
last = torch.randn(N, device=dev, generator=g)
for _ in range(5):
    gae(rew, val, done, None, last)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best = 1e9
for rep in range(3):
    e0.record()
    for _ in range(20):
        gae(rew, val, done, None, last)
    e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / 20)
print(f'T={T} N={N} gae {1e3 * best:.2f} us  {21 * T * N / (best * 1e-3) / 1e9:.0f} GB/s')
