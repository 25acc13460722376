"""Collective helpers of the trainer: what ``utils/mpi_tools.py`` does over MPI in the reference,
done over ``torch.distributed`` (NCCL on the GPUs, gloo in the CPU tests).

There is NO collective on the env step path: envs are sharded by global env id and never interact.
The collectives below carry gradients, running statistics and episode statistics only
(SURVEY.md section 2, table of MPI call sites).  Every helper is a no-op for world size 1.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence, Tuple

import torch
import torch.distributed as dist


def world() -> Tuple[int, int]:
    """(rank, world_size); (0, 1) when torch.distributed is not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n_total: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous env-index range [base, base + count) of ``rank``: the first ``n_total % world_size`` ranks
    own one env more.  RNG is keyed by the global env id, so results do not depend on ``world_size``."""
    if not (0 <= rank < world_size):
        raise ValueError('rank out of range')
    q, r = divmod(n_total, world_size)
    count = q + (1 if rank < r else 0)
    base = rank * q + min(rank, r)
    return base, count


def broadcast_params(tensors: Iterable[torch.Tensor], src: int = 0) -> None:
    """``mpi_tools.sync_params`` (mpi_tools.py:153-157): one Bcast per tensor there, ONE here."""
    _, ws = world()
    if ws == 1:
        return
    ts = [t for t in tensors]
    flat = torch.cat([t.detach().reshape(-1).float() for t in ts])
    dist.broadcast(flat, src=src)
    off = 0
    for t in ts:
        n = t.numel()
        t.data.copy_(flat[off:off + n].view_as(t))
        off += n


class FlatGradAverager:
    """``mpi_tools.mpi_avg_grads`` (mpi_tools.py:175-187) does one Allreduce per parameter tensor;
    here the gradients of a module are packed into one contiguous buffer and averaged with ONE
    all-reduce per optimiser step (messages are <= 43 KB: latency-bound, so fewer launches win)."""

    def __init__(self, params: Sequence[torch.nn.Parameter]):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        n = sum(p.numel() for p in self.params)
        p0 = self.params[0]
        self.flat = torch.zeros(n, dtype=torch.float32, device=p0.device)

    def __call__(self) -> None:
        _, ws = world()
        if ws == 1:
            return
        off = 0
        for p in self.params:
            n = p.numel()
            if p.grad is None:
                self.flat[off:off + n].zero_()
            else:
                self.flat[off:off + n].copy_(p.grad.reshape(-1))
            off += n
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
        self.flat.div_(ws)
        off = 0
        for p in self.params:
            n = p.numel()
            if p.grad is not None:
                p.grad.copy_(self.flat[off:off + n].view_as(p))
            off += n


def mean_across_ranks(x: torch.Tensor) -> torch.Tensor:
    """``mpi_avg_torch_tensor`` (mpi_tools.py:190-200): in-place average over ranks."""
    _, ws = world()
    if ws > 1:
        dist.all_reduce(x, op=dist.ReduceOp.SUM)
        x.div_(ws)
    return x


def reduce_episode_stats(episodes: float, return_sum: float, length_sum: float, ret_min: float, ret_max: float,
                         device) -> dict:
    """``mpi_statistics_scalar`` (mpi_tools.py:217-240) for the per-epoch episode statistics, packed into
    two small all-reduces (SUM and MAX) instead of four per logged key."""
    _, ws = world()
    s = torch.tensor([episodes, return_sum, length_sum], dtype=torch.float64, device=device)
    m = torch.tensor([-ret_min, ret_max], dtype=torch.float64, device=device)
    if ws > 1:
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
        dist.all_reduce(m, op=dist.ReduceOp.MAX)
    n = float(s[0].item())
    return {'episodes': int(n), 'mean_return': float(s[1].item()) / n if n else float('nan'),
            'mean_length': float(s[2].item()) / n if n else float('nan'),
            'min_return': -float(m[0].item()), 'max_return': float(m[1].item())}


def check_distributed_parameters(module: torch.nn.Module) -> bool:
    """``IWPGAlgorithm.check_distributed_parameters`` (iwpg.py:242-251): the sum of all parameters must be
    identical on every rank (min == max).  One packed all-reduce pair."""
    _, ws = world()
    if ws == 1:
        return True
    total = torch.stack([p.detach().double().sum() for p in module.parameters()]).sum().reshape(1)
    lo, hi = total.clone(), total.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    return bool(torch.allclose(lo, hi, rtol=0.0, atol=1e-6))


def bind_to_gpu_cpus(device_index: int) -> dict:
    """Pin the calling process to the CPUs NVML reports as local to CUDA device ``device_index`` (its NUMA node), so
    that pinned host buffers allocated afterwards (``cf_host_alloc``) and the threads that fill them sit next to the
    GPU's PCIe root.  One rank per GPU on a multi-socket box otherwise lands wherever the scheduler puts it, and the
    host-pointer step path (10 MB per step per rank) crosses the socket interconnect.  Best effort: returns what was
    done, never raises; a no-op when NVML, the affinity call or the device mapping is unavailable."""
    import os
    info = {'bound': False}
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            h = pynvml.nvmlDeviceGetHandleByUUID(('GPU-' + uuid).encode())
        except Exception:                                    # no uuid on this torch, or a MIG-style id
            if os.environ.get('CUDA_VISIBLE_DEVICES'):
                return {'bound': False, 'why': 'device mapping unknown under CUDA_VISIBLE_DEVICES'}
            h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        ncpu = os.cpu_count() or 1
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        use = cpus & os.sched_getaffinity(0)
        if use and len(use) < len(os.sched_getaffinity(0)):
            os.sched_setaffinity(0, use)
            info = {'bound': True, 'cpus': len(use), 'first_cpu': min(use), 'last_cpu': max(use)}
        else:
            info = {'bound': False, 'why': 'GPU-local CPU set is empty or already covers the process mask', 'cpus': len(use)}
    except Exception as e:                                   # noqa: BLE001 - placement is an optimisation, never an error
        info = {'bound': False, 'why': f'{type(e).__name__}: {e}'[:120]}
    return info
