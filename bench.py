"""bench.py - env-steps/s of the HJ-adversary hover env (BASELINE.json metric) on N B200s.

    python bench.py --gpus N --steps K --warmup W            # this framework (CUDA, C ABI)
    python bench.py --impl reference --gpus N --steps K ...  # the reference algorithm on the host cores

A "step" is one control step (env.step) of one shard of `--envs` envs (default 65 536 per GPU,
BASELINE.json configs[2]: DroneHoverBulletFreeEnvWithRandomHJAdversary-v0).  To keep every timed
step HBM-cold, each rank round-robins over `--shards` independent shards whose combined state +
I/O exceeds the 126 MB L2, so a shard's data has been evicted when its turn comes again.
One JSON line is printed by rank 0 (see DESIGN.md "measurement" for every key).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ENV_ID = 'DroneHoverBulletFreeEnvWithRandomHJAdversary-v0'
HOVER = 2.0 / 1.8 - 1.0
# ALGORITHMIC bytes per env-step (DESIGN.md "roofline"): state 216 B in +
# 216 B out, DR params 56 B in, action 16 B in, obs 136 B + reward 4 B + done 1 B + cost 4 B out,
# HJ packed sign table 1 B gathered.
BYTES_PER_ENV_STEP = 216 + 216 + 56 + 16 + 136 + 4 + 1 + 4 + 1
# dram__bytes_read.sum + dram__bytes_write.sum per k_step launch (bytes), mean of the 3 launches of the committed
# ncu --set full capture profiles/r1_v5_ncu_summary.md (caches flushed per replay): 42.36 MB read (this shard's state +
# actions + parameters, and the next shard's 21 MB prefetched into L2) + 0.05 MB written during the kernel; the
# ~25 MB of new state and I/O rows stay dirty in the 126 MB L2 and are written back while later launches run.
TRAFFIC_NCU = 42.41e6


try:
    _ORIG_AFFINITY = os.sched_getaffinity(0)                  # the CPU baseline runs on all of these, bound or not
except AttributeError:
    _ORIG_AFFINITY = set(range(os.cpu_count() or 1))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=2000)
    ap.add_argument('--warmup', type=int, default=200)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--envs', type=int, default=65536, help='envs per GPU per step')
    ap.add_argument('--shards', type=int, default=8, help='independent env shards per GPU (L2 eviction)')
    ap.add_argument('--kfuse', type=int, default=32, help='control steps per launch in the fused side measurement')
    ap.add_argument('--e2e-steps', type=int, default=100)
    ap.add_argument('--cpu-seconds', type=float, default=12.0, help='budget of the CPU baseline sample')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--graph-passes', type=int, default=4, help='round-robin passes over the shards captured in the CUDA graph')
    ap.add_argument('--streams', type=int, default=2,
                    help='independent shards are stepped on this many streams (parallel branches of the CUDA graph), so one '
                         "shard's load/store phases overlap another's compute")
    ap.add_argument('--no-prefetch', action='store_true',
                    help='do not chain the shards with cf_set_prefetch_next (each step kernel asks the L2 to prefetch the next shard)')
    ap.add_argument('--workload', default='hover_hj', choices=['hover_hj', 'initial', 'ppo'],
                    help='hover_hj = BASELINE.json configs[2] (the headline); initial = configs[3] (DroneHoverBulletEnvWithAdversaryInitial-v0, '
                         '131072 envs/GPU); ppo = configs[4] (rollout + fused GAE + PPO update, gradient all-reduce)')
    ap.add_argument('--ground-effect', type=int, default=0, help='workload initial: use_ground_effect switch (reference: 0)')
    ap.add_argument('--actions', default='hover_noise', choices=['hover_noise', 'uniform', 'pd'],
                    help='action source (SURVEY 8d config 3): hover + 0.05 N(0,1) resident in HBM (headline); U(-1,1) resident in HBM '
                         '(short episodes, reset-heavy); pd = on-device attitude-rate PD on the noisy obs, one extra kernel per step '
                         '(episodes run to the 500-step limit)')
    ap.add_argument('--no-bind', action='store_true', help='do not pin the process to the CPUs local to its GPU (NVML affinity)')
    ap.add_argument('--no-graph', action='store_true', help='launch every step directly instead of replaying a CUDA graph')
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--id={index}', f'--query-gpu={self.Q}', '--format=csv,noheader,nounits',
                                          '-lms', '100'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0: float, t1: float) -> dict:
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 - 0.05 <= t <= t1 + 0.15] or [r for _, r in self.rows]
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in rows:
            f = [x.strip() for x in r.split(',')]
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': statistics.median(sm) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': sorted(reasons), 'samples': len(sm)}


def measured_peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)['hbm_gbs']), 'MEASURED_PEAKS.json hbm_gbs (of measured)'
    return 6650.0, 'fallback 6.65 TB/s of B200_PROFILING.md (of fallback)'


def cpu_reference_run(envs: int, steps: int, warmup: int, seconds: float, env_id: str = ENV_ID):
    """The reference algorithm on the host cores: oracle-B (float64 C restatement of the reference's
    Python + restated Bullet step, OpenMP over envs) on a bounded sample of the same workload."""
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import cf_oracle as co
    from disturbance_crazyfile_simulation_b200 import abi, hj_tables
    cfg = abi.config_default(env_id)
    V = hj_tables.synthetic_table(1.5, 15)
    for li in range(1, 21):          # one table shared by all levels: the lookup cost does not depend on its values
        co.set_table(li, V)
    try:                                                     # all host cores again if the process was bound to the GPU's node
        os.sched_setaffinity(0, _ORIG_AFFINITY)
    except (AttributeError, OSError):
        pass
    co.lib().cfo_set_num_threads(os.cpu_count() or 1)        # all host cores, whatever OMP_NUM_THREADS the launcher exported
    orc = co.OracleEnvs(cfg, envs, seed=1)
    orc.reset()
    rng = np.random.default_rng(0)
    act = (HOVER + 0.05 * rng.standard_normal((envs, 4)))
    for _ in range(warmup):
        orc.step(act)
    t0 = time.perf_counter()
    done_steps = 0
    for _ in range(steps):
        orc.step(act)
        done_steps += 1
        if time.perf_counter() - t0 > seconds:
            break
    dt = time.perf_counter() - t0
    return envs * done_steps / dt, co.lib().cfo_num_threads(), done_steps, dt


def main_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    envs = 16384
    steps = max(1, min(args.steps, 2000))
    value, cores, done_steps, dt = cpu_reference_run(envs, steps, min(args.warmup, 5), 60.0)
    sample = f'{envs} envs x {done_steps} control steps of {ENV_ID} (noise, DR, auto-reset on), synthetic HJ table'
    line = {
        'impl': 'reference', 'metric': 'env-steps/sec (HJ-adversary hover)', 'value': value, 'unit': 'env-steps/s',
        'n_gpus': args.gpus, 'steps': done_steps, 'warmup': min(args.warmup, 5), 'ms_per_step': 1e3 * dt / done_steps,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': f'{ENV_ID}, bounded CPU sample of {envs} envs per step', 'envs_per_step': envs},
        'cpu_baseline': {'value': value, 'unit': 'env-steps/s', 'cores': cores, 'kind': 'port', 'sample': sample,
                         'note': 'the reference itself (pure Python + PyBullet 3.21) cannot be installed here; this is its '
                                 'algorithm restated in C (oracle-B), which is far faster than the reference\'s own '
                                 '~1.2e3 (no adversary) / ~1.5e2 (HJ) steps/s per process (BASELINE.md)'},
        'e2e': {'value': value, 'unit': 'env-steps/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    print(json.dumps(line))


WORKLOADS = {
    'hover_hj': dict(env_id='DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', envs=65536, tag='BASELINE.json configs[2]'),
    'initial': dict(env_id='DroneHoverBulletEnvWithAdversaryInitial-v0', envs=131072, tag='BASELINE.json configs[3]'),
}


def main_ours(args):
    import torch
    import torch.distributed as dist
    from disturbance_crazyfile_simulation_b200 import abi, hj_tables, make

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device - the product path has no CPU fallback')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    # one rank per GPU: sit on the GPU's own NUMA node before any pinned host buffer is allocated (e2e path)
    from disturbance_crazyfile_simulation_b200 import dist_utils
    host_binding = dist_utils.bind_to_gpu_cpus(local) if not args.no_bind else {'bound': False, 'why': '--no-bind'}
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    L = abi.lib()
    wl = WORKLOADS[args.workload]
    ENV_ID = wl['env_id']
    if args.workload == 'initial':
        args.envs = wl['envs'] if args.envs == 65536 else args.envs
        args.shards = 4 if args.shards == 8 else args.shards
    env_kw = {'use_ground_effect': args.ground_effect} if args.workload == 'initial' else {}
    n, S, K, W = args.envs, args.shards, args.steps, max(args.warmup, 3)
    shards = []
    for s in range(S):
        # global env ids: rank-major, then shard, so results do not depend on the number of GPUs;
        # one HJ table set per GPU, shared by its shards
        env = make(ENV_ID, num_envs=n, seed=2024, env_id_base=(rank * S + s) * n,
                   share_tables_with=shards[0] if shards else None, **env_kw)
        shards.append(env)
    od = shards[0].obs_dim
    g = torch.Generator(device=dev).manual_seed(1 + rank)
    if args.actions == 'uniform':
        acts = [(2.0 * torch.rand((n, 4), device=dev, generator=g) - 1.0).contiguous() for _ in range(S)]
    else:
        acts = [(HOVER + 0.05 * torch.randn((n, 4), device=dev, generator=g)).contiguous() for _ in range(S)]
    PD = args.actions == 'pd'
    pd_gains = (C.c_float * 6)(HOVER, 0.5, 0.3, 8.0, 0.05, 0.05)
    obs = [torch.empty((n, od), device=dev) for _ in range(S)]
    rew = [torch.empty(n, device=dev) for _ in range(S)]
    done = [torch.empty(n, dtype=torch.uint8, device=dev) for _ in range(S)]
    cost = [torch.empty(n, device=dev) for _ in range(S)]
    P = lambda t: C.c_void_p(t.data_ptr())

    NS = max(1, min(args.streams, S))
    GP = max(1, args.graph_passes)
    side_streams = [torch.cuda.Stream(device=dev) for _ in range(NS - 1)]

    def step_shard(i):
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        rc = L.cf_step_dev(shards[i]._h, P(acts[i]), P(obs[i]), P(rew[i]), P(done[i]), P(cost[i]), None, None, stream)
        if rc:
            abi.check(rc)
        if PD:                                           # the next action of this shard from the observation just written
            abi.check(L.cf_pd_actions_dev(P(obs[i]), od, n, P(acts[i]), pd_gains, stream))

    for e in shards:
        e.reset()
    if not args.no_prefetch and S > 1:
        for i in range(S):                               # shard i's step kernel prefetches shard i+1's state and actions into L2
            nxt = i + NS if i + NS < S else i % NS               # the shard this shard's stream steps next
            shards[i].set_prefetch_next(shards[nxt], acts[nxt])

    def run(k_steps, offset=0):
        """k_steps control steps, shard (offset+t) % S on stream t % NS (fork / join around the batch)"""
        main = torch.cuda.current_stream()
        for st in side_streams:
            st.wait_stream(main)
        for t in range(k_steps):
            i = (offset + t) % S
            if i % NS == 0:                                  # a shard is always stepped on the same stream: its steps stay ordered
                step_shard(i)
            else:
                with torch.cuda.stream(side_streams[i % NS - 1]):
                    step_shard(i)
        for st in side_streams:
            main.wait_stream(st)

    # ---- warm-up (direct launches), then one CUDA graph of S launches (one control step of every shard) --
    run(W)
    torch.cuda.synchronize()
    graph = None
    if not args.no_graph:
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            run(S * GP, offset=W)
        graph.replay()                                   # untimed warm-up replay
        torch.cuda.synchronize()

    def run_timed(k_steps):
        """exactly k_steps control steps: whole-graph replays (S steps each) + direct launches for the rest"""
        if graph is None:
            run(k_steps, offset=W)
            return
        for _ in range(k_steps // (S * GP)):
            graph.replay()
        run(k_steps % (S * GP), offset=W)

    # per-launch device time of the step kernel with CUDA events on the launching stream (direct launches)
    for e in shards:
        e.kernel_timing(True, reset=True)
    run(min(K, 200), offset=W)
    torch.cuda.synchronize()
    kt = [e.kernel_timing(False) for e in shards]
    kern_ms_events = sum(m * c for m, c in kt) / max(1, sum(c for _, c in kt))

    # ---- the timed region: exactly K steps between barrier + synchronize -----------------------------
    sampler = ClockSampler(local) if rank == 0 else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0 = time.time()
    ev0.record()
    run_timed(K)
    ev1.record()
    barrier()
    t1 = time.time()
    ms = ev0.elapsed_time(ev1)
    launches = K * (2 if PD else 1)                      # one k_step launch per control step (graph nodes included) [+ k_pd_actions]
    clocks = sampler.stop(t0, t1) if sampler else None
    kern_ms = ms / K                                     # launch-to-launch time of the dominant kernel inside the timed region
    t_ms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms_max = float(t_ms.item())
    value = n * K * world / (ms_max * 1e-3)

    # ---- side measurements (rank 0 only reports them) ---------------------------------------------
    # (a) L2-resident back-to-back stepping of ONE shard; (b) K control steps fused in one launch
    side = {}
    e0 = shards[0]
    ev0.record(); [step_shard(0) for _ in range(200)]; ev1.record(); torch.cuda.synchronize()
    side['one_shard_l2_resident_env_steps_per_s'] = n * 200 / (ev0.elapsed_time(ev1) * 1e-3)
    kf = args.kfuse
    actk = (HOVER + 0.05 * torch.randn((kf, n, 4), device=dev, generator=g)).contiguous()
    obsk = torch.empty((kf, n, od), device=dev); rewk = torch.empty((kf, n), device=dev)
    donek = torch.empty((kf, n), dtype=torch.uint8, device=dev); costk = torch.empty((kf, n), device=dev)
    e0.step_k(actk, obsk, rewk, donek, costk)
    ev0.record()
    for _ in range(5):
        e0.step_k(actk, obsk, rewk, donek, costk)
    ev1.record(); torch.cuda.synchronize()
    side[f'fused_k{kf}_env_steps_per_s'] = n * kf * 5 / (ev0.elapsed_time(ev1) * 1e-3)
    # (c) fused GAE pass (cf_gae_dev) over a [T][N] rollout buffer larger than L2: 21 algorithmic bytes per sample
    from disturbance_crazyfile_simulation_b200.gae import gae
    Tg, Ng = 32, 8 * n
    g_rew = torch.randn((Tg, Ng), device=dev); g_val = torch.randn((Tg, Ng), device=dev)
    g_done = (torch.rand((Tg, Ng), device=dev) < 0.01).to(torch.uint8); g_last = torch.randn(Ng, device=dev)
    gae(g_rew, g_val, g_done, None, g_last)
    ev0.record()
    for _ in range(5):
        gae(g_rew, g_val, g_done, None, g_last)
    ev1.record(); torch.cuda.synchronize()
    gae_ms = ev0.elapsed_time(ev1) / 5
    side['gae'] = {'T': Tg, 'N': Ng, 'ms': gae_ms, 'samples_per_s': Tg * Ng / (gae_ms * 1e-3),
                   'achieved_gb_s': 21 * Tg * Ng / (gae_ms * 1e-3) / 1e9, 'bytes_per_sample': 21}
    del g_rew, g_val, g_done, g_last
    stats = e0.episode_stats()
    side['mean_episode_length_shard0'] = stats['mean_length']
    # (d) HJ value-table generator: full-size (15^6) updates of the FasTrack solve, rank 0 only.
    # Algorithmic bytes per update: V read once + V' written once (l0 is recomputed, the stencil's neighbour reads are
    # cache hits by design) = 2 * 4 * 15^6 = 91.1 MB.
    if rank == 0:
        from disturbance_crazyfile_simulation_b200 import hj_solver
        hs = hj_solver.HjSolver(hj_solver.fastrack_problem(1.5, 15), device=local)
        hs.run(2.0, 1e-3, max_updates=20)
        ms0, n0 = hs.timing()
        hs.run(2.0, 1e-3, max_updates=200)
        ms1, n1 = hs.timing()
        upd_ms = (ms1 * n1 - ms0 * n0) / max(1, n1 - n0)
        side['hj_solver'] = {'grid': '15^6', 'ms_per_update': upd_ms, 'cells_per_s': 15 ** 6 / (upd_ms * 1e-3),
                             'bytes_per_update': 2 * 4 * 15 ** 6, 'achieved_gb_s': 2 * 4 * 15 ** 6 / (upd_ms * 1e-3) / 1e9,
                             'updates_timed': n1 - n0,
                             'note': 'float32, bit-identical to the NumPy restatement of optimized_dp graph_6D (accuracy low); '
                                     'latency / instruction bound (13-point 6-D stencil, IEEE divisions kept for exactness), '
                                     'not HBM bound'}
        hs.close()
        # (e) frames of capture_image(): 64 frames of 1920x1080x3 uint8 per launch, poses resident on the
        # device; algorithmic bytes = the frame buffer written once (3 B per pixel)
        from disturbance_crazyfile_simulation_b200 import raster
        rs = raster.Rasteriser(local)
        n_fr = 64
        rp = torch.zeros((n_fr, 7), device=dev)
        rp[:, :3] = torch.rand((n_fr, 3), device=dev) * 2 - 1
        rp[:, 2] = rp[:, 2].abs() + 0.3
        rp[:, 3:] = torch.nn.functional.normalize(torch.randn((n_fr, 4), device=dev), dim=1)
        r_out = torch.empty((n_fr, 1080, 1920, 3), dtype=torch.uint8, device=dev)
        for _ in range(3):
            rs.render(rp, out=r_out)
        ev0.record()
        for _ in range(10):
            rs.render(rp, out=r_out)
        ev1.record(); torch.cuda.synchronize()
        r_ms = ev0.elapsed_time(ev1) / 10
        side['raster'] = {'frames_per_launch': n_fr, 'triangles': rs.n_tris, 'ms_per_launch': r_ms,
                          'frames_per_s': n_fr / (r_ms * 1e-3), 'bytes_per_frame': 1080 * 1920 * 3,
                          'achieved_gb_s': n_fr * 1080 * 1920 * 3 / (r_ms * 1e-3) / 1e9,
                          'note': 'byte-identical to oracle/raster_oracle.py; parity with PyBullet OpenGL frames unpinned '
                                  '(graphics unavailable on this box)'}
        rs.close()
        del r_out

    # ---- e2e: the public host-pointer API (cf_step_host) with pinned HOST buffers ---------------
    def pinned(shape, dtype):
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        ptr = L.cf_host_alloc(nbytes)
        assert ptr, L.cf_last_error()
        return np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(ptr)).view(dtype).reshape(shape), ptr
    # D buffer sets in flight: shard i's download overlaps shard i+1's upload and kernel (cf_step_host_async + cf_host_sync);
    # every step still uploads its actions from pinned host memory and downloads obs / rew / done / cost into it
    D = min(3, S)
    sets, ptrs = [], []
    for _ in range(D):
        bufs = [pinned((n, 4), np.float32), pinned((n, od), np.float32), pinned((n,), np.float32), pinned((n,), np.uint8),
                pinned((n,), np.float32)]
        sets.append([b for b, _ in bufs]); ptrs += [q for _, q in bufs]
        sets[-1][0][:] = acts[0].cpu().numpy()
    ke = max(3, min(args.e2e_steps, K))

    def e2e_loop(count):
        for i in range(count):
            if i >= D:
                shards[(i - D) % S].host_sync()              # the results of step i-D are in host memory
            shards[i % S].step_host_async(*sets[i % D])
        for i in range(max(0, count - D), count):
            shards[i % S].host_sync()
    e2e_loop(2 * D)
    barrier()
    te0 = time.perf_counter()
    e2e_loop(ke)
    barrier()
    te = time.perf_counter() - te0
    t_e = torch.tensor([te], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
    e2e_value = n * ke * world / float(t_e.item())
    for q in ptrs:
        L.cf_host_free(q)

    if rank == 0:
        peak, peak_src = measured_peaks()
        achieved = BYTES_PER_ENV_STEP * n / (kern_ms * 1e-3) / 1e9
        line = {
            'metric': 'env-steps/sec (HJ-adversary hover)', 'value': value, 'unit': 'env-steps/s', 'n_gpus': world,
            'steps': K, 'warmup': W, 'ms_per_step': ms_max / K, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': f'{ENV_ID}, {n} envs/GPU per step ({wl["tag"]}); sensor noise, 10% domain '
                                   'randomisation, ' + ('Boltzmann HJ level' if args.workload == 'hover_hj' else f'HJ level 1.5, ground effect {args.ground_effect}') + ', in-kernel auto-reset, TimeLimit 500 all on',
                       'envs_per_gpu_per_step': n, 'shards_per_gpu': S,
                       'l2': f'inputs larger than L2: round-robin over {S} shards, {S * n * (304 + 16 + 4 * od + 9) / 1e6:.0f} MB '
                             'of state + I/O per GPU vs 126 MB L2', 'actions': {'hover_noise': 'hover + 0.05*N(0,1), resident in HBM', 'uniform': 'U(-1,1), resident in HBM',
                                   'pd': 'on-device attitude-rate PD on the noisy observation (cf_pd_actions_dev, one extra kernel per step)'}[args.actions],
                       'hj_tables': '20 synthetic float32 (15,)*6 tables (real blobs missing from the reference checkout), '
                                    'packed 2-bit sign tables in HBM/L2',
                       'step_mode': 'one kernel launch per control step' + ('' if args.no_graph else f', launches replayed from a CUDA graph of {S * GP} ({GP} passes over the {S} shards)')
                                    + (f'; the {S} independent shards alternate over {NS} streams' if NS > 1 else ''),
                       'shard_pipelining': ('off' if args.no_prefetch or S == 1 else
                                            'each step kernel issues cp.async.bulk.prefetch.L2 for the state + actions of the shard stepped next '
                                            '(cf_set_prefetch_next); every byte still comes from HBM once per step')},
            'clocks': clocks, 'gpu_launches': int(launches),
            'e2e': {'value': e2e_value, 'unit': 'env-steps/s', 'h2d_bytes_per_step': n * 16,
                    'd2h_bytes_per_step': n * (4 * od + 4 + 1 + 4), 'steps': ke,
                    'api': f'cf_step_host_async + cf_host_sync (C ABI, pinned host buffers, {D} steps in flight over the shards; '
                           'uploads and downloads inside the timed region)',
                    'host_binding_rank0': host_binding},
            'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                         'traffic': TRAFFIC_NCU, 'kernel': 'cf::k_step<true, CF_HJ_SIGN_TABLE>', 'avg_kernel_ms': kern_ms,
                         'avg_kernel_ms_events_per_launch': kern_ms_events,
                         'bytes_per_env_step': BYTES_PER_ENV_STEP, 'peak_source': peak_src,
                         'note': 'avg_kernel_ms = CUDA-event time of the timed region / launches in it (launch gaps included); '
                                 'avg_kernel_ms_events_per_launch = CUDA events around each direct launch on the launching stream; '
                                 'traffic = dram bytes per launch from the committed ncu --set full capture (profiles/)'},
            'side': side,
        }
        if not args.no_cpu_baseline:
            v, cores, steps_done, dt = cpu_reference_run(8192, 10000, 2, args.cpu_seconds, ENV_ID)
            line['cpu_baseline'] = {'value': v, 'unit': 'env-steps/s', 'cores': cores, 'kind': 'port',
                                    'sample': f'8192 envs x {steps_done} control steps of the same workload ({dt:.1f} s), '
                                              'oracle-B float64 C restatement, OpenMP'}
        print(json.dumps(line))
    for e in shards:
        e.close()
    if world > 1:
        dist.destroy_process_group()


def main_ppo(args):
    """BASELINE.json configs[4]: the PPO loop on DroneHoverBulletFreeEnvWithRandomHJAdversary-v0 - GPU rollout (policy MLP
    + cf_step_dev), fused GAE (cf_gae_dev), value / policy updates with the gradient all-reduce over the ranks.
    A "step" here is one epoch (T control steps of N envs per GPU + the update); value = env-steps/s end to end."""
    import torch
    import torch.distributed as dist
    from disturbance_crazyfile_simulation_b200 import make
    from disturbance_crazyfile_simulation_b200.trainer import PPOTrainer

    world = int(os.environ.get('WORLD_SIZE', '1')); rank = int(os.environ.get('RANK', '0')); local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)
    n, T = args.envs, 32
    env = make(ENV_ID, num_envs=n, seed=2024, env_id_base=rank * n, return_terminal_obs=True)
    tr = PPOTrainer(env, rollout_len=T, epochs=300, seed=1)
    K, W = max(1, min(args.steps, 20)), max(1, min(args.warmup, 3))
    for _ in range(W):
        tr.learn_one_epoch()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    infos = [tr.learn_one_epoch() for _ in range(K)]
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t.item())
    if rank == 0:
        roll = sum(i['rollout_s'] for i in infos) / K
        upd = sum(i['update_s'] for i in infos) / K
        print(json.dumps({
            'metric': 'env-steps/sec (PPO end to end, HJ-adversary hover)', 'value': n * T * K * world / dt, 'unit': 'env-steps/s',
            'n_gpus': world, 'steps': K, 'warmup': W, 'ms_per_step': 1e3 * dt / K, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': f'PPO on {ENV_ID} (BASELINE.json configs[4]): {n} envs/GPU x T={T} control steps per epoch, fused GAE, '
                                   '5x16 value mini-batch steps + 80 full-batch policy steps (reference defaults), one flattened '
                                   'gradient all-reduce per optimiser step', 'envs_per_gpu': n, 'rollout_len': T},
            'side': {'rollout_s_per_epoch': roll, 'update_s_per_epoch': upd, 'rollout_env_steps_per_s': n * T * world / roll,
                     'EpRet_last': infos[-1]['EpRet'], 'EpLen_last': infos[-1]['EpLen'], 'reference_logged_fps': '129-172 (BASELINE.md)'},
            'gpu_launches': int(env.launch_count)}))
    env.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    a = parse()
    if a.impl == 'reference':
        main_reference(a)
    elif a.workload == 'ppo':
        main_ppo(a)
    else:
        main_ours(a)
