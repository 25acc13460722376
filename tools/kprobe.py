"""kprobe.py - per-CTA timeline of the step kernel (cf_set_trace_buffer) and quick launch timing.

    python tools/kprobe.py [--env-id ID] [--envs N] [--shards S] [--steps K] [--kw name=value ...]

Prints the average step time of a CUDA-graph replay loop, then traces a few launches and reports
the distribution of CTA lifetimes, the phases of the slowest CTAs and the per-SM busy span.
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from disturbance_crazyfile_simulation_b200 import abi, make  # noqa: E402

HOVER = 2.0 / 1.8 - 1.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--env-id', default='DroneHoverBulletFreeEnvWithRandomHJAdversary-v0')
    ap.add_argument('--envs', type=int, default=65536)
    ap.add_argument('--shards', type=int, default=8)
    ap.add_argument('--steps', type=int, default=400)
    ap.add_argument('--traces', type=int, default=3)
    ap.add_argument('--streams', type=int, nargs='*', default=[1])
    ap.add_argument('--graph-reps', type=int, nargs='*', default=[1], help='round-robin passes over the shards captured per graph')
    ap.add_argument('--prefetch', action='store_true', help='each shard prefetches the next one (cf_set_prefetch_next)')
    ap.add_argument('--kw', nargs='*', default=[])
    a = ap.parse_args()
    kw = {}
    for item in a.kw:
        k, v = item.split('=')
        kw[k] = float(v) if '.' in v else int(v)
    L = abi.lib()
    dev = torch.device('cuda', 0)
    n, S = a.envs, a.shards
    shards = []
    for s in range(S):
        shards.append(make(a.env_id, num_envs=n, seed=2024, env_id_base=s * n, hj_tables='synthetic',
                           share_tables_with=shards[0] if shards else None, **kw))
    od = shards[0].obs_dim
    g = torch.Generator(device=dev).manual_seed(1)
    acts = [(HOVER + 0.05 * torch.randn((n, 4), device=dev, generator=g)).contiguous() for _ in range(S)]
    obs = [torch.empty((n, od), device=dev) for _ in range(S)]
    rew = [torch.empty(n, device=dev) for _ in range(S)]
    done = [torch.empty(n, dtype=torch.uint8, device=dev) for _ in range(S)]
    cost = [torch.empty(n, device=dev) for _ in range(S)]
    P = lambda t: C.c_void_p(t.data_ptr())

    def step(i):
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        abi.check(L.cf_step_dev(shards[i]._h, P(acts[i]), P(obs[i]), P(rew[i]), P(done[i]), P(cost[i]), None, None, st))

    if kw.get('rng_mode') == abi.CF_RNG_INJECTED:
        # PROBE ONLY: the step kernel with its draws read from buffers instead of generated (what a separate
        # generator kernel would feed it); the same random buffers every step, so this is not a valid simulation
        r = np.random.default_rng(0)
        for e in shards:
            e.set_noise_override(step_z=r.standard_normal((abi.CF_STEP_NORMALS, n)).astype(np.float32),
                                 step_u=r.random((abi.CF_STEP_UNIFORMS, n), dtype=np.float32),
                                 reset_z=r.standard_normal((abi.CF_RESET_NORMALS, n)).astype(np.float32),
                                 reset_u=r.random((abi.CF_RESET_UNIFORMS, n), dtype=np.float32))
    for e in shards:
        e.reset()
    if a.prefetch:
        ns0 = a.streams[0]
        for i in range(S):
            nxt = i + ns0 if i + ns0 < S else i % ns0       # the shard this shard's stream steps next (first --streams value)
            shards[i].set_prefetch_next(shards[nxt], acts[nxt])
    for t in range(200):
        step(t % S)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for ns in a.streams:
      for gr in a.graph_reps:
        side = [torch.cuda.Stream(device=dev) for _ in range(ns - 1)]
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            main = torch.cuda.current_stream()
            for st in side:
                st.wait_stream(main)
            for j in range(S * gr):
                i = j % S
                if i % ns == 0:                              # a shard always steps on the same stream
                    step(i)
                else:
                    with torch.cuda.stream(side[i % ns - 1]):
                        step(i)
            for st in side:
                main.wait_stream(st)
        graph.replay(); torch.cuda.synchronize()
        reps = max(1, a.steps // (S * gr))
        e0.record()
        for _ in range(reps):
            graph.replay()
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (reps * S * gr)
        print(f'[{a.env_id} {kw}] graph of {S * gr} launches, {ns} stream(s): {us:.2f} us/step  {n / us * 1e6:.3e} env-steps/s', flush=True)
    # direct launches, one at a time with a sync in between (isolated kernel time)
    ts = []
    for t in range(16):
        e0.record(); step(t % S); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    print(f'isolated launches: median {np.median(ts):.2f} us, min {min(ts):.2f} us')

    nt = L.cf_num_tiles(shards[0]._h)
    for rep in range(a.traces):
        i = rep % S
        tr = torch.zeros((nt, 16), dtype=torch.int64, device=dev)
        abi.check(L.cf_set_trace_buffer(shards[i]._h, P(tr)))
        step(i)
        torch.cuda.synchronize()
        abi.check(L.cf_set_trace_buffer(shards[i]._h, None))
        T = tr.cpu().numpy()
        t0 = T[:, 1].min()
        us_ = lambda c: (T[:, c] - t0) / 1e3
        start, landed, b1, b2, end, drawn = us_(1), us_(2), us_(7), us_(8), us_(9), us_(11)
        wfin = np.stack([us_(3 + w) for w in range(4)], axis=1)
        nd = T[:, 10]
        print(f'--- trace {rep}: kernel span {end.max():.2f} us | CTA start spread {start.max():.2f} | ctr landed median {np.median(us_(12)):.2f} max {us_(12).max():.2f} | draws done median {np.median(drawn):.2f} max {drawn.max():.2f} | landed median {np.median(landed):.2f} '
              f'max {landed.max():.2f} | warp step phase (from draws done) median {np.median(wfin - us_(13)[:, None]):.2f} '
              f'p99 {np.percentile(wfin - us_(13)[:, None], 99):.2f} max {(wfin - us_(13)[:, None]).max():.2f} | '
              f'barrier1 median {np.median(b1):.2f} max {b1.max():.2f} | reset phase median {np.median((b2 - b1)[nd > 0]) if (nd > 0).any() else 0:.2f} '
              f'max {(b2 - b1).max():.2f} | tail (end - resets done) median {np.median(end - b2):.2f} max {(end - b2).max():.2f} | '
              f'end median {np.median(end):.2f}')
        pc = lambda c: ' '.join(f'{np.percentile(us_(c), q):.2f}' for q in (0, 10, 50, 90, 100))
        print(f'    percentiles 0/10/50/90/100 us: ctr {pc(12)} | draws part 0 {pc(11)} | tile landed {pc(2)} | hj issued + draws part 1 {pc(13)}')
        if (nd > 0).any():
            r1, r2 = us_(14)[nd > 0], us_(15)[nd > 0]
            print(f'    reset sub-phases (tiles with resets, first group): R1 draws {np.median(r1 - b1[nd > 0]):.2f} | R2 tasks {np.median(r2 - r1):.2f} | '
                  f'R3 gyro + rest {np.median(b2[nd > 0] - r2):.2f} us')
        for kk in range(0, 7):
            sel = nd == kk
            if sel.any():
                print(f'    tiles with {kk} dones: {sel.sum():4d}  reset phase median {np.median((b2 - b1)[sel]):.2f} us  end median {np.median(end[sel]):.2f}')
        order = np.argsort(-end)[:4]
        for c in order:
            print(f'    tile {c:4d} sm {T[c, 0]:3d} start {start[c]:6.2f} landed {landed[c]:6.2f} warps {np.round(wfin[c], 2)} '
                  f'b1 {b1[c]:6.2f} b2 {b2[c]:6.2f} end {end[c]:6.2f} dones {nd[c]}')
        sm_end = {}
        for c in range(nt):
            sm_end[T[c, 0]] = max(sm_end.get(T[c, 0], 0.0), end[c])
        v = np.array(sorted(sm_end.values()))
        print(f'    per-SM finish: n_sm {len(v)} median {np.median(v):.2f} p90 {np.percentile(v, 90):.2f} max {v.max():.2f}; '
              f'CTAs per SM max {np.bincount(T[:, 0]).max()}')
    for e in shards:
        e.close()


if __name__ == '__main__':
    main()
