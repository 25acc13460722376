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


def ncu_kstep_metrics():
    """dram bytes and warp-instructions per k_step launch from the newest committed ncu --set full capture
    (profiles/*kstep_ncu_metrics.json, written by tools/ncu_summary.py --json from the .ncu-rep): a measurement of a
    profiled run of this same command, quoted with its file name - not a measurement of THIS run."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, 'profiles', '*kstep_ncu_metrics.json')))
    if not files:
        return None
    with open(files[-1]) as f:
        m = json.load(f)
    m['file'] = os.path.relpath(files[-1], ROOT)
    return m


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
    ap.add_argument('--shards', type=int, default=16, help='independent env shards per GPU (L2 eviction; every shard is its own handle)')
    ap.add_argument('--kfuse', type=int, default=32, help='control steps per launch in the fused side measurement')
    ap.add_argument('--e2e-steps', type=int, default=200)
    ap.add_argument('--e2e-separate', action='store_true', help='e2e: four separate pinned output arrays per buffer set (four downloads per step) instead of one adjacent block')
    ap.add_argument('--e2e-depth', type=int, default=8, help='host-pointer steps in flight (pinned buffer sets; at most one per shard) of the e2e measurement')
    ap.add_argument('--cpu-seconds', type=float, default=12.0, help='budget of the CPU baseline sample')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--graph-passes', type=int, default=4, help='round-robin passes over the shards captured in the CUDA graph')
    ap.add_argument('--streams', type=int, default=16,
                    help='independent shards are stepped on this many streams (parallel branches of the CUDA graph), so one '
                         "shard's load/store phases overlap another's compute")
    ap.add_argument('--prefetch', action='store_true',
                    help='chain the shards with cf_set_prefetch_next (each step kernel asks the L2 to prefetch the next shard): helped the '
                         'round-1 kernel at 4 CTAs/SM on two streams, measured -2 %% with 5 CTAs/SM on four streams (profiles/r2_kstep_ab.json)')
    ap.add_argument('--no-prefetch', action='store_true', help='(default since round 2; kept for old command lines)')
    ap.add_argument('--fused-update', type=int, default=-1, help='ppo workload: 1 / 0 = policy gradient by the fused kernel / by torch autograd (default: fused)')
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
    ap.add_argument('--repeats', type=int, default=0,
                    help='timed regions of exactly --steps launches each (median reported); 0 = as many as make the regions last ~60 ms')
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


def cpu_reference_run(envs: int, steps: int, warmup: int, seconds: float, env_id: str = ENV_ID, env_kw=None):
    """The reference algorithm on the host cores: oracle-B (float64 C restatement of the reference's
    Python + restated Bullet step, OpenMP over envs) on a bounded sample of the same workload.
    Nothing of the product package is imported or loaded here: config defaults and the synthetic HJ table come from
    oracle/ (cfo_config_default, cf_oracle.synthetic_table)."""
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import cf_oracle as co
    cfg = co.config_default(env_id)
    for k, v in (env_kw or {}).items():
        setattr(cfg, k, v)
    V = co.synthetic_table(1.5, 15)
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
    wl = WORKLOADS[args.workload if args.workload in WORKLOADS else 'hover_hj']
    env_id = wl['env_id']
    envs = wl['envs'] if args.envs == 65536 else args.envs           # the product arm's envs per GPU per step
    env_kw = {'use_ground_effect': args.ground_effect} if args.workload == 'initial' else {}
    steps = max(1, min(args.steps, 2000))
    warm = min(args.warmup, 5)
    value, cores, done_steps, dt = cpu_reference_run(envs, steps, warm, 120.0, env_id, env_kw)
    sample = (f'{envs} envs x {done_steps} control steps of {env_id} (noise, DR, auto-reset on), synthetic HJ table; every step is a '
              f'full {envs}-env step of the product arm\'s workload')
    loaded = sorted({ln.split()[-1] for ln in open('/proc/self/maps') if ln.rstrip().endswith('.so') and ROOT in ln}) \
        if os.path.exists('/proc/self/maps') else []
    line = {
        'impl': 'reference', 'metric': 'env-steps/sec (HJ-adversary hover)', 'value': value, 'unit': 'env-steps/s',
        'n_gpus': args.gpus, 'steps': done_steps, 'warmup': warm, 'ms_per_step': 1e3 * dt / done_steps,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': f'{env_id}, {envs} envs/GPU per step ({wl["tag"]}); sensor noise, 10% domain randomisation, '
                               + ('Boltzmann HJ level' if env_id == ENV_ID else f'HJ level 1.5, ground effect {args.ground_effect}')
                               + ', auto-reset, TimeLimit 500 all on - the product arm\'s config, stepped on the host cores',
                   'envs_per_gpu_per_step': envs},
        'cpu_baseline': {'value': value, 'unit': 'env-steps/s', 'cores': cores, 'kind': 'port', 'sample': sample,
                         'note': 'the reference itself (pure Python + PyBullet 3.21) cannot be installed here; this is its '
                                 'algorithm restated in C (oracle-B, float64, OpenMP over envs), far faster than the reference\'s own '
                                 '~1.2e3 (no adversary) / ~1.5e2 (HJ) steps/s per process (BASELINE.md)'},
        'e2e': {'value': value, 'unit': 'env-steps/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'native_so_in_repo_loaded': [os.path.relpath(x, ROOT) for x in loaded],
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
                   share_tables_with=shards[0] if shards else None, hj_tables='synthetic', **env_kw)
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
    if args.prefetch and not args.no_prefetch and S > 1:
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

    # ---- warm-up (direct launches); then every timed launch is replayed from CUDA graphs ------------------
    run(W)
    torch.cuda.synchronize()
    CH = S * GP                                          # launches per graph chunk
    graphs = {}                                          # (launches, first shard) -> captured graph

    def graph_for(k_steps, offset):
        key = (k_steps, offset % S)
        if key not in graphs:
            g_ = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g_):
                run(k_steps, offset=offset % S)
            g_.replay()                                  # untimed warm-up replay (one more control step of those shards)
            graphs[key] = g_
        return graphs[key]

    def plan(k_steps, offset):
        """[(launches, first shard)]: exactly k_steps control steps continuing the round-robin at `offset`, in chunks of
        at most CH launches; a shard always steps on the same stream, so any run of the round-robin is a legal graph."""
        out, o, left = [], offset, k_steps
        while left > 0:
            c = min(CH, left)
            out.append((c, o % S))
            o += c; left -= c
        return out

    def run_timed(k_steps, offset):
        """exactly k_steps control steps, every launch from a CUDA graph (or direct with --no-graph)"""
        if args.no_graph:
            run(k_steps, offset=offset)
            return
        for c, o in plan(k_steps, offset):
            graphs[(c, o)].replay()

    # per-launch device time of the step kernel with CUDA events on the launching stream (direct launches)
    for e in shards:
        e.kernel_timing(True, reset=True)
    run(min(max(K, 64), 200), offset=W)
    torch.cuda.synchronize()
    kt = [e.kernel_timing(False) for e in shards]
    kern_ms_events = sum(m * c for m, c in kt) / max(1, sum(c for _, c in kt))

    # ---- the timed region: exactly K steps between barrier + synchronize, repeated R times --------------------
    # A region of K = 20 launches lasts ~0.3 ms, shorter than the jitter of its own bracketing; the region is therefore
    # repeated R times back to back (R chosen so that the R regions last >= ~60 ms), each region timed with its own
    # CUDA-event pair on the launching stream, and the MEDIAN region is reported (`repeats`: R).  The round-robin over
    # the shards continues across regions, so every launch of every region finds its shard evicted from the L2.
    est_ms = max(kern_ms_events, 1e-3) * K
    R = 1 if args.repeats == 1 else (args.repeats if args.repeats > 0 else int(min(400, max(3, round(60.0 / est_ms)))))
    if not args.no_graph:
        for r in range(R):
            for c, o in plan(K, W + r * K):
                graph_for(c, o)
        torch.cuda.synchronize()
    sampler = ClockSampler(local) if rank == 0 else None
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(R + 1)]
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0 = time.time()
    evs[0].record()
    for r in range(R):
        run_timed(K, W + r * K)
        evs[r + 1].record()
    barrier()
    t1 = time.time()
    regions = [evs[r].elapsed_time(evs[r + 1]) for r in range(R)]
    ms = statistics.median(regions)
    deferred = os.environ.get('CF_DEFER_RESET', '1') != '0'   # single-step launches: k_step + k_reset_list (the in-place resets of the step)
    launches = K * ((2 if PD else 1) + (1 if deferred else 0))   # graph nodes per control step: k_step [+ k_reset_list] [+ k_pd_actions]
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
    for _ in range(5):                                       # warm-up: the first launches of a fresh 352 MB working set run 3-4 % slow
        gae(g_rew, g_val, g_done, None, g_last)
    gae_sets = []
    for _ in range(3):                                       # 3 sets of 20 launches on torch's stream, the median set is reported
        ev0.record()
        for _ in range(20):
            gae(g_rew, g_val, g_done, None, g_last)
        ev1.record(); torch.cuda.synchronize()
        gae_sets.append(ev0.elapsed_time(ev1) / 20)
    gae_ms = statistics.median(gae_sets)
    peak_gbs = measured_peaks()[0]
    side['gae'] = {'T': Tg, 'N': Ng, 'ms': gae_ms, 'ms_sets': gae_sets, 'launches_per_set': 20, 'samples_per_s': Tg * Ng / (gae_ms * 1e-3),
                   'achieved_gb_s': 21 * Tg * Ng / (gae_ms * 1e-3) / 1e9, 'bytes_per_sample': 21,
                   'frac_of_hbm_peak': 21 * Tg * Ng / (gae_ms * 1e-3) / 1e9 / peak_gbs}
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
    D = max(1, min(args.e2e_depth, S))
    sets, ptrs = [], []
    for _ in range(D):
        a_buf, a_ptr = pinned((n, 4), np.float32)
        ptrs.append(a_ptr)
        if args.e2e_separate or n % 4:
            bufs = [pinned((n, od), np.float32), pinned((n,), np.float32), pinned((n,), np.uint8), pinned((n,), np.float32)]
            o_buf, r_buf, d_buf, c_buf = (b for b, _ in bufs); ptrs += [q for _, q in bufs]
        else:
            # the four output arrays adjacent in ONE pinned block, in the order the library documents (obs | rew | cost | done):
            # cf_step_host_async then downloads them with a single copy
            blk, b_ptr = pinned((n * (4 * od + 9),), np.uint8)
            ptrs.append(b_ptr)
            o_buf = blk[:4 * n * od].view(np.float32).reshape(n, od)
            r_buf = blk[4 * n * od:4 * n * (od + 1)].view(np.float32)
            c_buf = blk[4 * n * (od + 1):4 * n * (od + 2)].view(np.float32)
            d_buf = blk[4 * n * (od + 2):]
        sets.append([a_buf, o_buf, r_buf, d_buf, c_buf])
        sets[-1][0][:] = acts[0].cpu().numpy()
    ke = max(3 * D, args.e2e_steps)                    # its own step count: a region of K = 20 steps would be 7 pipeline fills

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
    e2e_h2d, e2e_d2h = n * 16, n * (4 * od + 4 + 1 + 4)
    e2e_api = (f'cf_step_host_async + cf_host_sync (C ABI, pinned host buffers, {D} steps in flight over the shards; '
               + ('four output arrays, four downloads per step; ' if (args.e2e_separate or n % 4) else
                  'obs | rew | cost | done adjacent in one pinned block: one download per step; ')
               + 'uploads and downloads inside the timed region)')

    if rank == 0:
        peak, peak_src = measured_peaks()
        achieved = BYTES_PER_ENV_STEP * n / (kern_ms * 1e-3) / 1e9
        achieved_launch = BYTES_PER_ENV_STEP * n / (kern_ms_events * 1e-3) / 1e9
        nm = ncu_kstep_metrics()
        sm_hz = 1e6 * float((clocks or {}).get('sm_mhz') or 1965.0)
        issue_floor_ms = (nm['inst_executed_per_launch'] * (n / nm['envs_per_launch']) / (148 * 4 * sm_hz) * 1e3) if nm else None
        if args.no_graph:
            step_mode = 'one kernel launch per control step, launched directly'
        else:
            chunks = sorted({c for c, _ in plan(K, W)}, reverse=True)
            step_mode = (f'one kernel launch per control step; all {K} launches of a timed region are replayed from CUDA graphs '
                         f'(chunks of {" + ".join(str(c) for c in chunks)} launches, round-robin over the {S} shards continued across regions)')
        line = {
            'metric': 'env-steps/sec (HJ-adversary hover)', 'value': value, 'unit': 'env-steps/s', 'n_gpus': world,
            'steps': K, 'warmup': W, 'ms_per_step': ms_max / K, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'repeats': R, 'region_ms': {'median': ms, 'min': min(regions), 'max': max(regions), 'first': regions[0]},
            'config': {'workload': f'{ENV_ID}, {n} envs/GPU per step ({wl["tag"]}); sensor noise, 10% domain '
                                   'randomisation, ' + ('Boltzmann HJ level' if args.workload == 'hover_hj' else f'HJ level 1.5, ground effect {args.ground_effect}') + ', in-kernel auto-reset, TimeLimit 500 all on',
                       'envs_per_gpu_per_step': n, 'shards_per_gpu': S,
                       'l2': f'inputs larger than L2: round-robin over {S} shards, {S * n * (304 + 16 + 4 * od + 9) / 1e6:.0f} MB '
                             'of state + I/O per GPU vs 126 MB L2', 'actions': {'hover_noise': 'hover + 0.05*N(0,1), resident in HBM', 'uniform': 'U(-1,1), resident in HBM',
                                   'pd': 'on-device attitude-rate PD on the noisy observation (cf_pd_actions_dev, one extra kernel per step)'}[args.actions],
                       'hj_tables': '20 synthetic float32 (15,)*6 tables (real blobs missing from the reference checkout), '
                                    'packed 2-bit sign tables in HBM/L2',
                       'step_mode': step_mode + (f'; the {S} independent shards alternate over {NS} streams' if NS > 1 else ''),
                       'timing': f'{R} regions of exactly {K} launches each, back to back between one barrier + synchronize pair, each region '
                                 'timed by its own CUDA-event pair; value = median region, max over ranks',
                       'shard_pipelining': ('off' if not args.prefetch or args.no_prefetch or S == 1 else
                                            'each step kernel issues cp.async.bulk.prefetch.L2 for the state + actions of the shard stepped next '
                                            '(cf_set_prefetch_next); every byte still comes from HBM once per step')},
            'clocks': clocks, 'gpu_launches': int(launches),
            'e2e': {'value': e2e_value, 'unit': 'env-steps/s', 'h2d_bytes_per_step': e2e_h2d,
                    'd2h_bytes_per_step': e2e_d2h, 'steps': ke,
                    'api': e2e_api, 'host_binding_rank0': host_binding},
            'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                         'frac_per_launch': achieved_launch / peak,
                         'issue_frac': (issue_floor_ms / kern_ms) if issue_floor_ms else None,
                         'traffic': (nm['dram_bytes_per_launch'] if nm else None),
                         'traffic_source': (nm['file'] if nm else None),
                         'kernel': 'cf::k_step<true, CF_HJ_SIGN_TABLE>', 'avg_kernel_ms': kern_ms,
                         'avg_kernel_ms_events_per_launch': kern_ms_events,
                         'issue_floor_ms': issue_floor_ms,
                         'warp_inst_per_launch': (nm['inst_executed_per_launch'] * (n / nm['envs_per_launch']) if nm else None),
                         'bytes_per_env_step': BYTES_PER_ENV_STEP, 'peak_source': peak_src,
                         'note': 'frac = sustained: algorithmic bytes per launch / (median timed region / launches in it), the launches of the '
                                 'two streams overlap; frac_per_launch = the same bytes / the un-overlapped duration of one launch (CUDA events '
                                 'around direct launches on the launching stream); issue_frac = the second ceiling: warp-instructions per '
                                 'launch (ncu smsp__inst_executed.sum of the committed capture) / (148 SMs x 4 schedulers x the SM clock '
                                 'sampled during this run) / avg_kernel_ms; traffic = dram bytes per launch of that same committed ncu '
                                 '--set full capture (traffic_source), not of this run'},
            'side': side,
        }
        if not args.no_cpu_baseline:
            v, cores, steps_done, dt = cpu_reference_run(8192, 10000, 2, args.cpu_seconds, ENV_ID, env_kw)
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
    env = make(ENV_ID, num_envs=n, seed=2024, env_id_base=rank * n, return_terminal_obs=True, hj_tables='synthetic')
    tr = PPOTrainer(env, rollout_len=T, epochs=300, seed=1, fused_update=None if args.fused_update < 0 else bool(args.fused_update))
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
    # side: the fused network kernel alone (CUDA events, 200 launches on the rollout's last observation rows) against the fp32 FFMA peak
    pol = {}
    if getattr(tr, '_fp', None) is not None:
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(20):
            tr._fp.step(tr._obs_all[T], tr._scratch_act, tr.last_val, tr._scratch_logp)
        ev0.record()
        for _ in range(200):
            tr._fp.step(tr._obs_all[T], tr._scratch_act, tr.last_val, tr._scratch_logp)
        ev1.record(); torch.cuda.synchronize()
        us = 1e3 * ev0.elapsed_time(ev1) / 200
        od = env.obs_dim
        macs = od * 50 + 50 * 50 + 50 * 4 + od * 64 + 64 * 64 + 64            # per env, reference layer sizes
        peak = 148 * 128 * 2 * 1.965e9
        pol = {'policy_kernel_us': us, 'policy_kernel_env_steps_per_s': n / (us * 1e-6), 'useful_flop_per_env': 2 * macs,
               'ffma_frac_of_fp32_peak': 2 * macs * n / (us * 1e-6) / peak, 'fp32_peak_tflops_nominal': peak / 1e12}
        upd_floor = (80 * 3 * 2 * (od * 50 + 50 * 50 + 50 * 4) + 5 * 3 * 2 * (od * 64 + 64 * 64 + 64)) * n * T / peak
        pol['update_fp32_ffma_floor_s'] = upd_floor
        if tr.fused_update:                                  # the fused policy-gradient kernel alone on the last epoch's batch
            for _ in range(3):
                tr._fp.pi_grad(tr._g_obs, tr._g_act, tr._g_adv, tr._g_logp, tr.clip_ratio)
            ev0.record()
            for _ in range(20):
                tr._fp.pi_grad(tr._g_obs, tr._g_act, tr._g_adv, tr._g_logp, tr.clip_ratio)
            ev1.record(); torch.cuda.synchronize()
            ms = ev0.elapsed_time(ev1) / 20
            useful = 3 * 2 * (od * 50 + 50 * 50 + 50 * 4) * n * T
            pol.update(pi_grad_kernel_ms=ms, pi_grad_ffma_frac_of_fp32_peak=useful / (ms * 1e-3) / peak)
    # the ranks apply the same all-reduced gradients from the same initial weights: their parameters must stay identical
    rank_diff = 0.0
    if world > 1:
        flat = torch.cat([p.detach().reshape(-1) for p in tr.ac.parameters()])
        ref = flat.clone()
        dist.broadcast(ref, src=0)
        dmax = (flat - ref).abs().max().reshape(1)
        dist.all_reduce(dmax, op=dist.ReduceOp.MAX)
        rank_diff = float(dmax.item())
        pol['param_max_diff_across_ranks'] = rank_diff
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
                     'EpRet_last': infos[-1]['EpRet'], 'EpLen_last': infos[-1]['EpLen'], 'reference_logged_fps': '129-172 (BASELINE.md)',
                     'fused_rollout': bool(tr.fused), 'cuda_graphs': bool(tr.use_graphs), 'fused_update': bool(tr.fused_update), **pol},
            'gpu_launches': int(env.launch_count)}), flush=True)
    # captured graphs hold the NCCL all-reduce of the optimiser steps: release them before the communicator goes (destroying
    # the process group under live graphs hung the 2-GPU run of round 2 until its timeout)
    tr._pi_graph = tr._v_graph = tr._rollout_graph = None
    del tr
    import gc
    gc.collect()
    torch.cuda.synchronize()
    env.close()
    if world > 1:
        dist.barrier()
        sys.stdout.flush()
        os._exit(0)                                          # every rank has printed / finished: skip communicator teardown


if __name__ == '__main__':
    a = parse()
    if a.impl == 'reference':
        main_reference(a)
    elif a.workload == 'ppo':
        main_ppo(a)
    else:
        main_ours(a)
