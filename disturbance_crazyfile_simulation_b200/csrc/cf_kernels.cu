// cf_kernels.cu - kernels of the batched CrazyFlie hover environment and their launchers (sm_100a).
//
// Layout in HBM: per handle one float array S[n_tiles][76][CF_TILE] (word-major tiles of CF_TILE
// envs, see cf_device.cuh), the HJ tables (packed 2-bit sign tables in one slab and/or fp32 value
// tables), and caller-owned env-major I/O rows act[n][4], obs[n][obs_dim], rew[n], done[n], cost[n].
//
// Step kernel: one CTA per tile, one thread per env.  The 60 dynamic state words of the tile
// (30 KB, contiguous) arrive in shared memory with ONE bulk async copy (cp.async.bulk + mbarrier,
// SASS UBLKCP) while the threads fetch their action and domain-randomised parameters; the step
// runs against shared memory (conflict-free word-major accesses), the observation rows are
// assembled in shared memory, and the tile's new state (30 KB) and observation rows (17 KB) leave
// with one bulk copy each.  No inter-thread communication on the step path.
#include "cf_device.cuh"
#include "cf_kernels.h"

#include <cstdlib>

#ifndef CF_WS_DEFAULT
#define CF_WS_DEFAULT 0                  /* 1: the sensor-noise variants launch the warp-specialised kernel (CF_STEP_WS=0/1 overrides at run time) */
#endif

namespace cf {

// ---------------------------------------------------------------------------------- bulk-copy PTX
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// global -> shared, completion counted in bytes on the mbarrier (16-byte aligned, 16-byte multiple)
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
#ifdef CF_STREAM_EVICT_FIRST
    uint64_t pol;                                           // the tile is read once per step: first in line for eviction
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
#else
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
#endif
}
// shared -> global, tracked by the thread's bulk async-group
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 ::"l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
// hint: bring [p, p + bytes) into the L2 (16-byte aligned, 16-byte multiple); no completion to wait for
__device__ __forceinline__ void bulk_prefetch_l2(const void* p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// make generic-proxy shared-memory writes visible to the async proxy (before a bulk store reads them)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) :: "memory"); return t; }
__device__ __forceinline__ unsigned smid() { unsigned r; asm volatile("mov.u32 %0, %%smid;" : "=r"(r)); return r; }

// ---------------------------------------------------------------------------------- kernels
// all envs of the padded tile array get a valid constructor state
__global__ void k_init(float* __restrict__ S, int n_pad, DevCfg cfg) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pad) return;
    init_env(env_ref_global(S, i), cfg);
}

template <bool NOISE>
__global__ void __launch_bounds__(128) k_reset(float* __restrict__ S, int n, const __grid_constant__ DevCfg cfg,
                                              const uint8_t* __restrict__ mask, float* __restrict__ obs,
                                              const float* inj_z, const float* inj_u) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (mask && !mask[i]) return;
    constexpr int OD = NOISE ? 34 : 42;
    const EnvRef g = env_ref_global(S, i);
    long long gid = cfg.gid_base + i;
    consume_fresh(g, cfg, gid, NOISE);
    uint32_t ctr = g.u(CF_S_RNG_CTR);
    if (!cfg.injected) g.u(CF_S_RNG_CTR, ctr + 1u);
    DrawSrc dw = make_draws(cfg, gid, ctr, 2u, inj_z, inj_u, n, i, CF_RESET_UNIFORMS);
    float z[CF_RESET_NORMALS], u[(CF_RESET_UNIFORMS + 3) / 4 * 4];
    reset_draws(dw, cfg, z, u);
    reset_env<NOISE>(g, g, cfg, z, u, obs + (size_t)i * OD);
}

template <int OD>
__device__ __forceinline__ void copy_row(float* __restrict__ dst, const float* src) {
    float2* d = reinterpret_cast<float2*>(dst);
    const float2* s = reinterpret_cast<const float2*>(src);
#pragma unroll
    for (int j = 0; j < OD / 2; ++j) d[j] = s[j];
}

#ifndef CF_DRAW_UNROLL
#define CF_DRAW_UNROLL 1                 /* Philox blocks per trip of the step's draw loop */
#endif
#ifndef CF_DRAW_HOOK
#define CF_DRAW_HOOK 6                   /* blocks generated before the loop waits for the state tile */
#endif
#ifndef CF_WS_CTAS
#define CF_WS_CTAS 3                     /* warp-specialised kernel: CTAs of 3 * CF_TILE threads per SM (56 registers per thread at launch) */
#endif
#ifndef CF_WS_PRODUCER_REGS
#define CF_WS_PRODUCER_REGS 40
#endif
#ifndef CF_WS_CONSUMER_REGS
#define CF_WS_CONSUMER_REGS 88           /* 128 x 88 + 256 x 40 = 384 x 56 */
#endif
// barrier over the threads that own an env: the whole CTA, or the consumer warps (named barrier 1) of the warp-specialised kernel
template <bool WS>
__device__ __forceinline__ void tile_sync() {
    if (WS) asm volatile("bar.sync 1, %0;" ::"n"(CF_TILE) : "memory");
    else __syncthreads();
}
#ifndef CF_MIN_CTAS
#define CF_MIN_CTAS (640 / CF_TILE)      /* 5 CTAs = 640 threads / SM: <= 96 registers and <= 44.4 KB of shared memory per CTA */
#endif

// env.step for all envs, `k` control steps per launch with the tile's state held in shared memory.
// Timeline of a CTA: (1) thread 0 starts the bulk copy of the tile's 39 staged state words (19.5 KB) and an L2 prefetch of its
// 21 history words; every thread fetches its event counter, action and domain-randomised parameters with ordinary loads and
// generates the step's standard draws - ONE rolled loop over the 13 Philox blocks (Box-Muller for the first 11) into
// shared-memory scratch; after block 5 the loop waits for the tile, issues the HJ sign-table gather and carries on, so the first
// half of the draws hides the tile's flight and the second half the gather's; (2) the step runs against shared memory, reading
// and writing the history words straight in HBM; (3) envs that finished an episode are reset in place by the whole CTA: up to
// four at a time, 23 threads per env generate the reset's Philox blocks, then the reset's independent tasks (pose / parameters,
// the two observations, action history) run on different warps with one env per lane; (4) the tile's staged words and the
// observation rows leave with 128-bit coalesced stores.
// WS (warp-specialised, sensor-noise variants only): the CTA has 3 * CF_TILE threads.  Threads CF_TILE.. are PRODUCERS, two per
// env, that do nothing but generate the step's eleven Philox + Box-Muller blocks (even / odd blocks) into the draw slots and
// announce them in four stages on mbarriers; threads 0..CF_TILE-1 are the CONSUMERS that run the step as before, minus the draw
// loop, waiting for a stage where its normals are first read.  Registers follow the roles (setmaxnreg: producers shrink to
// CF_WS_PRODUCER_REGS, consumers grow to CF_WS_CONSUMER_REGS), so three such CTAs = 36 warps fit an SM.  A launch of 65 536 envs
// is 2 048 warps of one role in the single-role kernel - 3.5 per scheduler, each a chain of dependent instructions; here the
// draws (45 % of a CTA's life there) run beside the physics on 6 144 warps.
// INRESET: finished envs are reset in place by the CTA (needed when k > 1: the episode continues inside the launch).  Without
// it (single-step launches) the step kernel only appends the env to a list and the reset is done by k_reset_list behind it on
// the same stream: the reset is a 2.3 us latency chain that 63 % of the tiles of a 65 536-env step ran with 126 of 128 threads
// idle (21 % of the sustained step time, measured on one launch of 524 288 envs with and without finished episodes), and its
// code is a quarter of the kernel; in the list kernel a warp owns a finished env and other streams' step kernels fill the SMs.
template <bool NOISE, int HJ_MODE, bool WS, bool INRESET>
__global__ void __launch_bounds__(WS ? 3 * CF_TILE : CF_TILE, WS ? CF_WS_CTAS : CF_MIN_CTAS)
k_step(float* __restrict__ S, int n, int k, const __grid_constant__ DevCfg cfg, const __grid_constant__ DevTables tab,
       const __grid_constant__ StepIO io) {
    static_assert(!WS || NOISE, "the warp-specialised kernel is built for the sensor-noise variants");
    constexpr int OD = NOISE ? 34 : 42;
    constexpr int NW = CF_TILE / 32;
    constexpr int RC = CF_TILE >= 128 ? 4 : 2;              // finished envs reset concurrently: 23 threads each for the draws + 1
    constexpr bool kParallelReset = NOISE && NW >= 2;
    static_assert(!kParallelReset || 24 * RC <= CF_TILE, "reset groups");
    static_assert(RC * 128 <= CF_EXTRA_SLOTS * CF_TILE, "reset scratch aliases the extra draw slots");
    constexpr uint32_t kStateBytes = CF_BULK_WORDS * CF_TILE * 4, kHistBytes = CF_HIST_WORDS * CF_TILE * 4;
    extern __shared__ __align__(128) float smem[];
    float* s_state = smem;                                  // [39][CF_TILE]
    float* s_obs = s_state + CF_BULK_WORDS * CF_TILE;       // [CF_TILE][OD]; before the rows are written: draw slots 0..31
    float* s_extra = s_obs + OD * CF_TILE;                  // [12][CF_TILE] draw slots 32..43; after the step phase: reset scratch
    float* s_draw = s_extra;                                // [RC][96] reset draws (z[48], u[44]) of the envs being reset
    float* s_ctx = s_draw + RC * 96;                        // [RC][32] ResetCtx of the envs being reset
    float* s_thr = s_extra + CF_EXTRA_SLOTS * CF_TILE;      // [6][16]
    uint64_t* mbar = reinterpret_cast<uint64_t*>(s_thr + 96);   // [0] state tile landed, [1..4] draw stages (WS)
    int* s_ndone = reinterpret_cast<int*>(mbar + 5);        // [2]
    uint8_t* s_list = reinterpret_cast<uint8_t*>(s_ndone + 2);   // [CF_TILE] thread ids of finished envs
    uint32_t* s_hj = reinterpret_cast<uint32_t*>(s_list + CF_TILE);   // [CF_TILE] landing slots of the sign-table gather (cp.async)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, tile = blockIdx.x;
    if (WS && tid >= CF_TILE) {
        // ---------------------------------------------------------------- producer warps
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(CF_WS_PRODUCER_REGS));
        const int pe = (tid - CF_TILE) & (CF_TILE - 1), half = (tid - CF_TILE) / CF_TILE, pi = tile * CF_TILE + pe;
        const float* ptile = S + (size_t)tile * (CF_STATE_WORDS * CF_TILE);
        const uint32_t pctr = __float_as_uint(__ldcg(ptile + cf_dev_word(CF_S_RNG_CTR) * CF_TILE + pe));
        StepDraws ps;
        ps.s = s_obs + (pe >> 5) * (32 * OD) + (pe & 31);
        ps.x = s_extra + pe;
        ps.stage_bar = nullptr; ps.stage_parity = 0u;
        __syncthreads();                                    // the stage barriers are initialised (thread 0)
        for (int t = 0; t < k; ++t) {
            const DrawSrc dw = make_draws(cfg, cfg.gid_base + pi, pctr + (uint32_t)t, 0u, io.inj_step_z, io.inj_step_u, n, pi, CF_STEP_UNIFORMS);
#pragma unroll 1
            for (int b = half; b < CF_DRAW_BLOCKS; b += 2) {
                if (pi < n) ps.put(b, dw.z(cfg, b));
                // stages end with blocks 0 | 1, 4 | 3, 6 | 5, 10 | 9 of the even | odd producer: one arrival per thread and stage
                const int st = b <= 1 ? 0 : (b == 3 || b == 4 ? 1 : (b == 5 || b == 6 ? 2 : (b >= 9 ? 3 : -1)));
                if (st >= 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(mbar + 1 + st)) : "memory");
            }
            if (t + 1 < k) __syncthreads();                 // the consumers have stored the rows: the slots may be overwritten
        }
        return;
    }
    const int i = tile * CF_TILE + tid;
#ifdef CF_GATHER_TO_REGISTER                                  /* tools/ only: A/B against the register-destination gather */
    uint32_t* const my_hj = nullptr;
#else
    uint32_t* const my_hj = s_hj + tid;
#endif
    float* gtile = S + (size_t)tile * (CF_STATE_WORDS * CF_TILE);
    const bool valid = i < n;
    const unsigned wmask = __ballot_sync(0xffffffffu, valid);
    const bool bulk_obs = io.bulk_obs && (tile + 1) * CF_TILE <= n;
    unsigned long long* tr = io.trace ? io.trace + CF_TRACE_WORDS * (size_t)tile : nullptr;    // per-CTA timeline (debug)
    if (tr && tid == 0) { tr[0] = smid(); tr[1] = gtimer(); }
    // the event counter first (the draws wait for it), then the tile: it should not queue behind 10 MB of bulk copies
#ifdef CF_PROBE_NOCTR
    uint32_t ctr0 = 7u + (uint32_t)k;                       // tools/ only: how much of a CTA's life is the wait for its counter word
#else
    uint32_t ctr0 = WS ? 0u : __float_as_uint(__ldcg(gtile + cf_dev_word(CF_S_RNG_CTR) * CF_TILE + tid));   // WS: read from the tile
#endif
    if (tid == 0) {
        s_ndone[0] = 0;
        mbar_init(mbar, 1);
        if (WS) {
#pragma unroll
            for (int st = 0; st < 4; ++st) mbar_init(mbar + 1 + st, 2 * CF_TILE);
        }
        mbar_arrive_expect_tx(mbar, kStateBytes);
        bulk_g2s(s_state, gtile, kStateBytes, mbar);
        bulk_prefetch_l2(gtile + CF_BULK_WORDS * CF_TILE, kHistBytes);
    }
    const EnvRef e{s_state + tid, gtile + tid};
    const EnvRef g{gtile + tid, gtile + tid};
    for (int q = tid; q < 96; q += CF_TILE) s_thr[q] = cfg.hj_thr[q];               // from the constant bank: no global load in front of the HJ lookup
    DrParams p;
    p.load(g);                                              // coalesced loads, in flight with the bulk copy
    const long long gid = cfg.gid_base + i;
    float* row = s_obs + tid * OD;
    StepDraws sd;
    sd.s = s_obs + warp * (32 * OD) + lane;
    sd.x = s_extra + tid;
    sd.stage_bar = WS ? mbar + 1 : nullptr; sd.stage_parity = 0u;
    if (WS) {
        __syncthreads();                                    // all roles: barriers initialised, thresholds in place
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(CF_WS_CONSUMER_REGS));
    }
    if (tr && tid == 0) {
        asm volatile("" :: "r"(ctr0)); tr[12] = gtimer();
    }
    int total_done = 0;
    for (int t = 0; t < k; ++t) {
        const size_t r = (size_t)t * n + i;
        bool my_reset = false;
        float4 av = make_float4(0.f, 0.f, 0.f, 0.f);
        if (valid) av = __ldg(reinterpret_cast<const float4*>(io.act) + r);
        HjFetch hjf;
        hjf.raw = 0u; hjf.shift = 0; hjf.slot = nullptr;
        // step event of step t: counter ctr0 + t, tag 0.  One rolled loop over the 11 normal blocks keeps the kernel inside the
        // instruction cache (each unrolled Philox + Box-Muller body is 2.2 KB of code); CF_DRAW_UNROLL blocks per trip.
        if (WS) {
            if (t == 0) {
                mbar_wait(mbar, 0);
                if (tr && tid == 0) tr[2] = gtimer();
                ctr0 = e.u(CF_S_RNG_CTR);
                if (valid) consume_fresh(e, cfg, gid, NOISE);
            }
            if (valid) hjf = step_issue_hj<HJ_MODE>(e, cfg, tab, s_thr, my_hj);
            sd.stage_parity = (uint32_t)(t & 1);
        }
        const DrawSrc dw = make_draws(cfg, gid, ctr0 + (uint32_t)t, 0u, io.inj_step_z, io.inj_step_u, n, i, CF_STEP_UNIFORMS);
        if (!WS) {
            constexpr int U = CF_DRAW_UNROLL, kHook = CF_DRAW_HOOK;
            static_assert(kHook % U == 0 && kHook < CF_DRAW_BLOCKS, "the tile hook sits at the top of a trip");
#pragma unroll 1
            for (int b0 = 0; b0 < CF_DRAW_BLOCKS; b0 += U) {
                if (b0 == kHook) {
                    if (t == 0) {
                        if (tr && tid == 0) tr[11] = gtimer();
                        __syncthreads();                    // mbarrier initialised, thresholds in place
                        mbar_wait(mbar, 0);
                        if (tr && tid == 0) tr[2] = gtimer();
                        if (valid) consume_fresh(e, cfg, gid, NOISE);
                    }
                    // the table gather is in flight behind the second half of the draws
                    if (valid) hjf = step_issue_hj<HJ_MODE>(e, cfg, tab, s_thr, my_hj);
                }
                if (valid) {
                    float4 v[U];
#pragma unroll
                    for (int u = 0; u < U; ++u)
                        if (b0 + u < CF_DRAW_BLOCKS && step_block_live<NOISE>(b0 + u)) v[u] = dw.z(cfg, b0 + u);
#pragma unroll
                    for (int u = 0; u < U; ++u)
                        if (b0 + u < CF_DRAW_BLOCKS && step_block_live<NOISE>(b0 + u)) sd.put(b0 + u, v[u]);
                }
            }
        }
        if (tr && tid == 0 && t == 0) tr[13] = gtimer();
        if (valid) {
            const float a[4] = {av.x, av.y, av.z, av.w};
            e.u(CF_S_RNG_CTR, ctr0 + (uint32_t)t + 1u);     // an in-place reset in this launch: counter after the step, tag 1
            const StepOut so = step_env<NOISE, HJ_MODE>(e, p, cfg, tab, hjf, a, sd, dw, row, wmask);
            // this warp's table gather has been consumed: HBM may now stream in the inputs of the handle stepped next
            // (cf_set_prefetch_next); each warp asks for a quarter of the CTA's share
            if (io.pf_S && lane == 0 && t == 0) {
                constexpr uint32_t kQ = CF_STATE_WORDS * CF_TILE * 4 / NW;
                for (int pt = tile; pt < io.pf_tiles; pt += gridDim.x) {
                    bulk_prefetch_l2(reinterpret_cast<const char*>(io.pf_S + (size_t)pt * (CF_STATE_WORDS * CF_TILE)) + warp * kQ, kQ);
                    if (io.pf_act && warp == 0 && (pt + 1) * CF_TILE <= io.pf_n) bulk_prefetch_l2(io.pf_act + (size_t)pt * CF_TILE * 4, CF_TILE * 16);
                }
            }
            io.rew[r] = so.rew;
            io.done[r] = (uint8_t)so.done;
            if (io.cost) io.cost[r] = so.cost;
            if (io.truncated) io.truncated[i] = (uint8_t)so.truncated;
            if (so.done) {
                io.fin_ret[i] = so.ep_ret;
                io.fin_len[i] = min(so.ep_len, 0xffff) | (misc_level(e.u(CF_S_MISC)) << 16);   // length | level index of the finished episode
                double* agg = io.agg + 4 * (tile % CF_AGG_SLOTS);
                atomicAdd(agg + 0, 1.0);
                atomicAdd(agg + 1, (double)so.ep_ret);
                atomicAdd(agg + 2, (double)so.ep_len);
                if (cfg.auto_reset) {
                    if (io.term_obs) copy_row<OD>(io.term_obs + (size_t)i * OD, row);
                    if (INRESET) {
                        s_list[atomicAdd(s_ndone, 1)] = (uint8_t)tid;
                        my_reset = true;
                    } else {
                        io.reset_list[atomicAdd(io.reset_count, 1)] = i;
                    }
                }
            }
        }
        if (tr && t == 0) { __syncwarp(); if (lane == 0) tr[3 + warp] = gtimer(); }
        fence_async_smem();
        tile_sync<WS>();                                    // step phase of the tile complete, list of finished envs final
        if (tr && tid == 0 && t == 0) tr[7] = gtimer();
        const int nd = INRESET ? s_ndone[0] : 0;
        if (INRESET && nd > 0) {                            // uniform over the CTA
            if (kParallelReset) {
                for (int c0 = 0; c0 < nd; c0 += RC) {
                    const int cnt = min(RC, nd - c0);
                    // R1: the reset events' draws (23 Philox blocks per env, one thread each) + what is read from the old state
                    if (tid < 23 * cnt) {
                        const int j = tid / 23, b = tid - 23 * j, et = s_list[c0 + j];
                        const EnvRef ee{s_state + et, gtile + et};
                        const DrawSrc dr = make_draws(cfg, cfg.gid_base + tile * CF_TILE + et, ee.u(CF_S_RNG_CTR), 1u, io.inj_reset_z,
                                                      io.inj_reset_u, n, tile * CF_TILE + et, CF_RESET_UNIFORMS);
                        float4 v;
                        float* dst;
                        if (b < 12) { v = dr.z(cfg, b); dst = s_draw + j * 96 + 4 * b; }
                        else { v = dr.u(cfg, b - 12); dst = s_draw + j * 96 + CF_RESET_NORMALS + 4 * (b - 12); }
                        dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; dst[3] = v.w;
                    } else if (tid >= 23 * RC && tid < 23 * RC + cnt) {
                        const int j = tid - 23 * RC, et = s_list[c0 + j];
                        ResetCtx* c = reinterpret_cast<ResetCtx*>(s_ctx + j * 32);
                        reset_task_stale(EnvRef{s_state + et, gtile + et}, *c);
                    }
                    tile_sync<WS>();
                    if (tr && tid == 0 && t == 0 && c0 == 0) tr[14] = gtimer();
                    // R2, one env per lane: pose + motor state (warp 0) | domain randomisation, counters, level (warp 1) |
                    // gyro-free part of the first (warp 2) and second (warp 3) observation + action history
                    if (lane < cnt) {
                        const int et = s_list[c0 + lane];
                        const float* zs = s_draw + lane * 96;
                        const float* us = zs + CF_RESET_NORMALS;
                        const EnvRef ee{s_state + et, gtile + et};
                        float* rw = s_obs + et * OD;
#pragma unroll
                        for (int task = 0; task < 4; ++task) {
                            if (warp >= (NW < 4 ? NW : 4) || task % (NW < 4 ? NW : 4) != warp) continue;
                            if (task == 0) reset_task_pose(ee, cfg, zs, us, reinterpret_cast<ResetCtx*>(s_ctx + lane * 32)->wb);
                            else if (task == 1) reset_task_params(ee, EnvRef{gtile + et, gtile + et}, cfg, us);
                            else if (task == 2) reset_task_obs_nongyro<0>(ee, cfg, zs, us, rw);
                            else reset_task_obs_nongyro<1>(ee, cfg, zs, us, rw);
                        }
                    }
                    tile_sync<WS>();
                    if (tr && tid == 0 && t == 0 && c0 == 0) tr[15] = gtimer();
                    // R3: the gyro samples of both observations (need the new body rate and the old gyro state)
                    if (warp == 0 && lane < cnt) {
                        const int et = s_list[c0 + lane];
                        reset_task_gyro(EnvRef{s_state + et, gtile + et}, *reinterpret_cast<const ResetCtx*>(s_ctx + lane * 32), cfg,
                                        s_draw + lane * 96, s_obs + et * OD);
                    }
                    if (c0 + RC < nd) tile_sync<WS>();      // the scratch is reused by the next group
                }
            } else {
                for (int j = warp; j < nd; j += NW) {       // one env per warp: 23 lanes draw, lane 0 resets
                    const int et = s_list[j];
                    const EnvRef ee{s_state + et, gtile + et};
                    const EnvRef gg{gtile + et, gtile + et};
                    float* zs = s_draw + (warp % RC) * 96;
                    float* us = zs + CF_RESET_NORMALS;
                    const DrawSrc dr = make_draws(cfg, cfg.gid_base + tile * CF_TILE + et, ee.u(CF_S_RNG_CTR), 1u, io.inj_reset_z,
                                                  io.inj_reset_u, n, tile * CF_TILE + et, CF_RESET_UNIFORMS);
                    if (lane < 12) {
                        float4 v = dr.z(cfg, lane);
                        zs[4 * lane] = v.x; zs[4 * lane + 1] = v.y; zs[4 * lane + 2] = v.z; zs[4 * lane + 3] = v.w;
                    } else if (lane < 23) {
                        float4 v = dr.u(cfg, lane - 12);
                        us[4 * (lane - 12)] = v.x; us[4 * (lane - 12) + 1] = v.y; us[4 * (lane - 12) + 2] = v.z; us[4 * (lane - 12) + 3] = v.w;
                    }
                    __syncwarp();
                    if (lane == 0) reset_env<NOISE>(ee, gg, cfg, zs, us, s_obs + et * OD);
                    __syncwarp();
                }
            }
            total_done += nd;
            fence_async_smem();
            tile_sync<WS>();                                // resets complete
            if (tid == 0) s_ndone[0] = 0;
            if (my_reset && t + 1 < k) p.load(g);
        }
        if (tr && tid == 0 && t == 0) tr[8] = gtimer();
        // The tile's rows and staged state leave with 128-bit stores straight from shared memory (coalesced, conflict-free):
        // unlike a bulk copy they do not have to be waited for, so the CTA's slot is free as soon as they are issued.
        if (bulk_obs) {
            float4* dst = reinterpret_cast<float4*>(io.obs + ((size_t)t * n + (size_t)tile * CF_TILE) * OD);
            const float4* src = reinterpret_cast<const float4*>(s_obs);
#pragma unroll
            for (int j = tid; j < CF_TILE * OD / 4; j += CF_TILE) dst[j] = src[j];
        } else if (valid) {
            copy_row<OD>(io.obs + r * OD, row);
        }
        if (t + 1 == k) {
            float4* dst = reinterpret_cast<float4*>(gtile);
            const float4* src = reinterpret_cast<const float4*>(s_state);
#pragma unroll
            for (int j = tid; j < CF_BULK_WORDS * CF_TILE / 4; j += CF_TILE) dst[j] = src[j];
        }
        if (t + 1 < k) __syncthreads();                     // rows / extra scratch are overwritten by the next step's draws
    }
    if (tr && tid == 0) { tr[9] = gtimer(); tr[10] = (unsigned long long)total_done; }
}

// Deferred in-place resets of a single-step launch: one WARP per finished env (grid-stride over the list the step kernel wrote).
// Lanes 0..22 generate the reset event's 23 Philox blocks (tag 1, the env's counter as the step left it) into the warp's scratch,
// lane 0 then runs the reset against the tile in HBM and writes the observation row - the same tasks in the same arithmetic as the
// in-kernel reset (reset_env composes them), so the two paths are bit-identical.  The last CTA to finish zeroes the list counters.
template <bool NOISE>
__global__ void __launch_bounds__(128) k_reset_list(float* __restrict__ S, int n, const __grid_constant__ DevCfg cfg,
                                                    const int* __restrict__ list, int* __restrict__ count, float* __restrict__ obs,
                                                    const float* inj_z, const float* inj_u) {
    constexpr int OD = NOISE ? 34 : 42;
    __shared__ float s_draw[4][96];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nd = *reinterpret_cast<volatile const int*>(count);
    float* zs = s_draw[warp];
    float* us = zs + CF_RESET_NORMALS;
    for (int j = blockIdx.x * 4 + warp; j < nd; j += gridDim.x * 4) {
        const int i = list[j];
        const EnvRef g = env_ref_global(S, i);
        const DrawSrc dr = make_draws(cfg, cfg.gid_base + i, g.u(CF_S_RNG_CTR), 1u, inj_z, inj_u, n, i, CF_RESET_UNIFORMS);
        if (lane < 12) {
            const float4 v = dr.z(cfg, lane);
            zs[4 * lane] = v.x; zs[4 * lane + 1] = v.y; zs[4 * lane + 2] = v.z; zs[4 * lane + 3] = v.w;
        } else if (lane < 23) {
            const float4 v = dr.u(cfg, lane - 12);
            us[4 * (lane - 12)] = v.x; us[4 * (lane - 12) + 1] = v.y; us[4 * (lane - 12) + 2] = v.z; us[4 * (lane - 12) + 3] = v.w;
        }
        __syncwarp();
        if (lane == 0) reset_env<NOISE>(g, g, cfg, zs, us, obs + (size_t)i * OD);
        __syncwarp();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(count + 1, 1) == (int)gridDim.x - 1) { count[0] = 0; count[1] = 0; __threadfence(); }
    }
}

// Packed 2-bit sign table of one level from its fp32 value table, with the reference's own
// float32/float64 rule (distur_gener.py:40-125): one thread per byte = 4 consecutive cells.
__global__ void k_hj_signs(const float* __restrict__ V, uint8_t* __restrict__ P, int n_bytes) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_bytes) return;
    unsigned out = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        long c = 4L * b + j;
        if (c >= CF_HJ_CELLS) break;
        int i4 = (int)((c / 15) % 15), i3 = (int)((c / 225) % 15);
        int bits = (hj_deriv_positive(V, c, i3, 225, c_dx[3]) ? 1 : 0) | (hj_deriv_positive(V, c, i4, 15, c_dx[4]) ? 2 : 0);
        out |= (unsigned)bits << (2 * j);
    }
    P[b] = (uint8_t)out;
}

// Grid.get_index + sign lookup for given float64 (roll,pitch,yaw,p,q,r) states (the float64 rule)
__global__ void k_hj_lookup(const double* __restrict__ states, int n, const float* V, const uint8_t* P,
                            int* __restrict__ idx_out, float* __restrict__ sign_out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int idx[6];
#pragma unroll
    for (int d = 0; d < 6; ++d) { idx[d] = hj_index_dim(states[(size_t)i * 6 + d], d); idx_out[(size_t)i * 6 + d] = idx[d]; }
    int bits = P ? hj_signbits_packed(P, idx) : hj_signbits_value(V, idx);
    sign_out[2 * (size_t)i] = (bits & 1) ? -1.0f : 1.0f;
    sign_out[2 * (size_t)i + 1] = (bits & 2) ? -1.0f : 1.0f;
}
// the same for float32 states through the step kernel's own float32 threshold rule
__global__ void k_hj_lookup_f32(const float* __restrict__ states, int n, const __grid_constant__ DevCfg cfg, const float* V,
                                const uint8_t* P, const float* __restrict__ thr, int* __restrict__ idx_out,
                                float* __restrict__ sign_out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int idx[6];
#pragma unroll
    for (int d = 0; d < 6; ++d) { idx[d] = hj_index_f32(states[(size_t)i * 6 + d], d, thr, cfg); idx_out[(size_t)i * 6 + d] = idx[d]; }
    int bits = P ? hj_signbits_packed(P, idx) : hj_signbits_value(V, idx);
    sign_out[2 * (size_t)i] = (bits & 1) ? -1.0f : 1.0f;
    sign_out[2 * (size_t)i + 1] = (bits & 2) ? -1.0f : 1.0f;
}

// env-major host record [n][76] <-> tiled device state
__global__ void k_unpack_state(const float* __restrict__ S, int n, float* __restrict__ rec) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* g = env_global(const_cast<float*>(S), i);
    for (int w = 0; w < CF_STATE_WORDS; ++w) rec[(size_t)i * CF_STATE_WORDS + w] = g[cf_dev_word(w) * CF_TILE];
}
__global__ void k_pack_state(float* __restrict__ S, int n, const float* __restrict__ rec) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float* g = env_global(S, i);
    for (int w = 0; w < CF_STATE_WORDS; ++w) g[cf_dev_word(w) * CF_TILE] = rec[(size_t)i * CF_STATE_WORDS + w];
}
__global__ void k_levels(const float* __restrict__ S, int n, float* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = (float)misc_level(__float_as_uint(env_global(const_cast<float*>(S), i)[cf_dev_word(CF_S_MISC) * CF_TILE])) / 10.0f;
}

// PPO GAE (Buffer.finish_path, algs/core.py:498-535; discount_cumsum :105-119), time-major
// [T][N]: one thread per env walks its column backwards; every access is coalesced across envs.
__global__ void k_gae(const float* __restrict__ rew, const float* __restrict__ val, const uint8_t* __restrict__ done,
                      const float* __restrict__ end_val, const float* __restrict__ last_val, float gamma, float lam,
                      float rew_scale, int T, int N, float* __restrict__ adv, float* __restrict__ tgt,
                      float* __restrict__ disc_ret) {
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    // float64 accumulators like scipy.signal.lfilter in the reference; the kernel is HBM-bound
    double next_val = (double)last_val[n], run_adv = 0.0, run_ret = next_val;
    const double g = (double)gamma, gl = (double)gamma * (double)lam;
#pragma unroll 4
    for (int t = T - 1; t >= 0; --t) {
        size_t k = (size_t)t * N + n;
        float r = rew[k], v = val[k];
        if (done[k]) { next_val = end_val ? (double)end_val[k] : 0.0; run_adv = 0.0; run_ret = next_val; }
        run_ret = (double)r + g * run_ret;
        double rs = (double)r;
        if (rew_scale > 0.0f) rs = (double)fminf(10.0f, fmaxf(-10.0f, __fmul_rn(r, rew_scale)));
        double delta = rs + g * next_val - (double)v;
        run_adv = delta + gl * run_adv;
        adv[k] = (float)run_adv;
        tgt[k] = (float)(run_adv + (double)v);
        disc_ret[k] = (float)run_ret;
        next_val = (double)v;
    }
}

// The same scan with four env columns per thread: 128-bit loads and stores (one uchar4 for the flags), four independent
// recurrences per thread and the loads of the next time steps issued ahead of the arithmetic.  Needs N % 4 == 0 and 16-byte
// aligned arrays (launch_gae falls back to k_gae otherwise).  Same float64 recurrences, bit-identical results.
#ifndef CF_GAE_U
#define CF_GAE_U 2
#endif
// (U, min CTAs/SM, threads) sweep on one box, T = 32 x N = 524 288, 20 launches (tools/gpu_r2_gae_sweep.sh): (2,7,128) 62.9 us =
// 5.60 TB/s = 0.856 of the measured HBM peak; (1,8,128) 62.9; (2,14,64) 62.6; (2,8,128) 65.7; (2,4,256) 67.1; U = 3 / 4 spill: 80-83 us.
// At N = 2 097 152 the same kernel reaches 5.94 TB/s = 0.91: what is missing at the bench size is the ramp and tail of a 63 us kernel.
#ifndef CF_GAE_THREADS
#define CF_GAE_THREADS 128
#endif
#ifndef CF_GAE_MIN_CTAS
#define CF_GAE_MIN_CTAS 7                /* <= 72 registers: N = 524 288 columns (1 024 CTAs) are one wave on 148 SMs */
#endif
struct GaeLane {
    double next_val, run_adv, run_ret;
};
__device__ __forceinline__ void gae_lane(GaeLane& s, float r, float v, bool d, float ev, bool has_ev, double g, double gl,
                                         float rew_scale, float& adv, float& tgt, float& ret) {
    if (d) { s.next_val = has_ev ? (double)ev : 0.0; s.run_adv = 0.0; s.run_ret = s.next_val; }
    s.run_ret = (double)r + g * s.run_ret;
    double rs = (double)r;
    if (rew_scale > 0.0f) rs = (double)fminf(10.0f, fmaxf(-10.0f, __fmul_rn(r, rew_scale)));
    const double delta = rs + g * s.next_val - (double)v;
    s.run_adv = delta + gl * s.run_adv;
    adv = (float)s.run_adv;
    tgt = (float)(s.run_adv + (double)v);
    ret = (float)s.run_ret;
    s.next_val = (double)v;
}
__global__ void __launch_bounds__(CF_GAE_THREADS, CF_GAE_MIN_CTAS) k_gae4(const float4* __restrict__ rew, const float4* __restrict__ val,
                                              const uchar4* __restrict__ done, const float4* __restrict__ end_val,
                                              const float4* __restrict__ last_val, float gamma, float lam, float rew_scale, int T,
                                              int N4, float4* __restrict__ adv, float4* __restrict__ tgt, float4* __restrict__ disc_ret) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N4) return;
    const double g = (double)gamma, gl = (double)gamma * (double)lam;
    const float4 lv = last_val[n];
    GaeLane s[4] = {{(double)lv.x, 0.0, (double)lv.x}, {(double)lv.y, 0.0, (double)lv.y}, {(double)lv.z, 0.0, (double)lv.z},
                    {(double)lv.w, 0.0, (double)lv.w}};
    const bool has_ev = end_val != nullptr;
    constexpr int U = CF_GAE_U;                             // time steps whose loads are in flight together
    for (int t0 = T - 1; t0 >= 0; t0 -= U) {
        float4 r[U], v[U], ev[U];
        uchar4 d[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int t = t0 - u;
            if (t < 0) break;
            const size_t k = (size_t)t * N4 + n;
            r[u] = __ldcs(rew + k); v[u] = __ldcs(val + k); d[u] = __ldcs(done + k);
            ev[u] = has_ev ? __ldcs(end_val + k) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int t = t0 - u;
            if (t < 0) break;
            const size_t k = (size_t)t * N4 + n;
            float4 a, tg, rt;
            gae_lane(s[0], r[u].x, v[u].x, d[u].x != 0, ev[u].x, has_ev, g, gl, rew_scale, a.x, tg.x, rt.x);
            gae_lane(s[1], r[u].y, v[u].y, d[u].y != 0, ev[u].y, has_ev, g, gl, rew_scale, a.y, tg.y, rt.y);
            gae_lane(s[2], r[u].z, v[u].z, d[u].z != 0, ev[u].z, has_ev, g, gl, rew_scale, a.z, tg.z, rt.z);
            gae_lane(s[3], r[u].w, v[u].w, d[u].w != 0, ev[u].w, has_ev, g, gl, rew_scale, a.w, tg.w, rt.w);
            __stcs(adv + k, a); __stcs(tgt + k, tg); __stcs(disc_ret + k, rt);
        }
    }
}

// Synthetic "PID actions" source of the throughput runs (BASELINE.json north_star: "random/PID actions"): an
// attitude -> rate -> motor-mix PD law on the env's own NOISY observation, one thread per env.  Structure after
// the firmware cascade the reference cites (envs/control.py:13-26 gains table, :33-50 rpy_control_factors_to_PWM,
// QUAD_FORMATION_X mixing; :94-100 PWM = 30000 + 30000 a), expressed in action units; gains are arguments.
// obs rows [n][obs_dim], o(k) = (pos3, quat xyzw, vel3, gyro3) at float offset `off`.
__global__ void k_pd_actions(const float* __restrict__ obs, int obs_dim, int off, int n, float* __restrict__ act, PdGains g) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* o = obs + (size_t)i * obs_dim + off;
    const float z = o[2], qx = o[3], qy = o[4], qz = o[5], qw = o[6], vz = o[9], wx = o[10], wy = o[11], wz = o[12];
    const float roll = atan2f(2.0f * (qw * qx + qy * qz), 1.0f - 2.0f * (qx * qx + qy * qy));
    const float pitch = asinf(clampf(2.0f * (qw * qy - qz * qx), -1.0f, 1.0f));
    const float T = g.hover + g.kz * (1.0f - z) - g.kvz * vz;
    const float r = g.krate * (-g.katt * roll - wx), p = g.krate * (-g.katt * pitch - wy), y = g.kyaw * (-wz);
    float4 a;
    a.x = clampf(T - r - p - y, -1.0f, 1.0f);
    a.y = clampf(T - r + p + y, -1.0f, 1.0f);
    a.z = clampf(T + r + p - y, -1.0f, 1.0f);
    a.w = clampf(T + r - p + y, -1.0f, 1.0f);
    reinterpret_cast<float4*>(act)[i] = a;
}

// Parity hooks for the production RNG path (tests/test_rng_bits.py): the raw Philox words and the standard draws of one
// event, produced by the very device functions the step / reset kernels call (Rng::raw, u01, box_muller).
// One thread per (env, block): blocks 0..nzb-1 are normal blocks, nzb..nzb+nub-1 the uniform blocks (counter 16 + b).
__global__ void k_debug_draws(const __grid_constant__ DevCfg cfg, int first_env, int count, uint32_t ctr, uint32_t tag, int nzb,
                              int nub, uint32_t* __restrict__ words, float* __restrict__ draws) {
    const int nb = nzb + nub, t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count * nb) return;
    const int j = t / nb, b = t - j * nb;
    const Rng rng = make_rng(cfg.gid_base + first_env + j, ctr, tag);
    const bool normal = b < nzb;
    const uint4 w = rng.raw(cfg, normal ? (uint32_t)b : 16u + (uint32_t)(b - nzb));
    const float4 v = normal ? rng.normals(cfg, (uint32_t)b) : rng.uniforms(cfg, (uint32_t)(b - nzb));
    reinterpret_cast<uint4*>(words)[t] = w;
    reinterpret_cast<float4*>(draws)[t] = v;
}
// quat [n][4] (x, y, z, w) -> the two Euler conversions of a step (quat_to_rpy_bullet, quat2euler_hj), [n][3] each
__global__ void k_debug_euler(const float4* __restrict__ quat, long long n, float* __restrict__ rpy_bullet, float* __restrict__ rpy_hj) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const float4 v = quat[t];
    const float q[4] = {v.x, v.y, v.z, v.w};
    float a[3], b[3];
    quat_to_rpy_bullet(q, a);
    quat2euler_hj(q, b);
#pragma unroll
    for (int i = 0; i < 3; ++i) { rpy_bullet[3 * t + i] = a[i]; rpy_hj[3 * t + i] = b[i]; }
}
__global__ void k_debug_box_muller(const uint4* __restrict__ words, long long n, float4* __restrict__ z, float4* __restrict__ u) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const uint4 w = words[t];
    z[t] = box_muller(w);
    u[t] = make_float4(u01(w.x), u01(w.y), u01(w.z), u01(w.w));
}

// ---------------------------------------------------------------------------------- launchers
static inline int blocks_for(int n, int threads) { return (n + threads - 1) / threads; }

cudaError_t upload_constants(const double grid[6][CF_HJ_PTS], const double dx[6], const double cdf[21]) {
    cudaError_t err = cudaMemcpyToSymbol(c_grid, grid, sizeof(double) * 6 * CF_HJ_PTS);
    if (err != cudaSuccess) return err;
    err = cudaMemcpyToSymbol(c_dx, dx, sizeof(double) * 6);
    if (err != cudaSuccess) return err;
    return cudaMemcpyToSymbol(c_boltz_cdf, cdf, sizeof(double) * 21);
}

void launch_init(float* S, int n_pad, const DevCfg& cfg, cudaStream_t st) {
    k_init<<<blocks_for(n_pad, 128), 128, 0, st>>>(S, n_pad, cfg);
}
void launch_reset(float* S, int n, const DevCfg& cfg, const uint8_t* mask, float* obs, const float* inj_z,
                  const float* inj_u, cudaStream_t st) {
    if (cfg.obs_noise) k_reset<true><<<blocks_for(n, 128), 128, 0, st>>>(S, n, cfg, mask, obs, inj_z, inj_u);
    else k_reset<false><<<blocks_for(n, 128), 128, 0, st>>>(S, n, cfg, mask, obs, inj_z, inj_u);
}

size_t step_smem_bytes(int obs_dim) {
    // staged state words + observation rows (= draw slots 0..31) + extra draw slots (= reset scratch) + thresholds + mbarrier +
    // finished-env counter + list + gather landing slots
    return (size_t)(CF_BULK_WORDS + obs_dim + CF_EXTRA_SLOTS) * CF_TILE * 4 + 96 * 4 + 5 * 8 + 8 + CF_TILE + CF_TILE * 4;
}

template <bool NOISE, int HJ_MODE, bool WS, bool INRESET>
static cudaError_t launch_step_variant(float* S, int n, int k, const DevCfg& cfg, const DevTables& tab, const StepIO& io,
                                       cudaStream_t st) {
    static bool configured[64] = {};                     // per device and variant: opt in to > 48 KB of dynamic shared memory
    const size_t smem = step_smem_bytes(NOISE ? 34 : 42);
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        cudaError_t err = cudaFuncSetAttribute(k_step<NOISE, HJ_MODE, WS, INRESET>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return err;
        cudaFuncSetAttribute(k_step<NOISE, HJ_MODE, WS, INRESET>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        configured[dev] = true;
    }
    k_step<NOISE, HJ_MODE, WS, INRESET><<<blocks_for(n, CF_TILE), WS ? 3 * CF_TILE : CF_TILE, smem, st>>>(S, n, k, cfg, tab, io);
    if (!INRESET && cfg.auto_reset) {
        // a warp per finished env (grid-stride): ~1 % of the envs finish per step at the headline workload, up to 2 % with random
        // actions - one warp per 64 envs keeps it at about one list entry per warp
        const int grid = (n + 255) / 256;
        k_reset_list<NOISE><<<grid, 128, 0, st>>>(S, n, cfg, io.reset_list, io.reset_count, io.obs, io.inj_reset_z, io.inj_reset_u);
    }
    return cudaSuccess;
}
cudaError_t launch_step(float* S, int n, int k, const DevCfg& cfg, const DevTables& tab, const StepIO& io, cudaStream_t st) {
    const bool sign = cfg.hj_mode == CF_HJ_SIGN_TABLE;
    static const bool ws = [] { const char* v = getenv("CF_STEP_WS"); return v ? atoi(v) != 0 : (CF_WS_DEFAULT != 0); }();
    static const bool defer_ok = [] { const char* v = getenv("CF_DEFER_RESET"); return v ? atoi(v) != 0 : true; }();
    // single-step launches hand finished envs to k_reset_list; K-step launches reset in place (the episode goes on in the launch)
    const bool inreset = k > 1 || !defer_ok || !io.reset_list;
#define CF_LAUNCH(NOISE_, WS_)                                                                                                   \
    do {                                                                                                                         \
        if (sign) return inreset ? launch_step_variant<NOISE_, CF_HJ_SIGN_TABLE, WS_, true>(S, n, k, cfg, tab, io, st)           \
                                 : launch_step_variant<NOISE_, CF_HJ_SIGN_TABLE, WS_, false>(S, n, k, cfg, tab, io, st);         \
        return inreset ? launch_step_variant<NOISE_, CF_HJ_VALUE_TABLE, WS_, true>(S, n, k, cfg, tab, io, st)                    \
                       : launch_step_variant<NOISE_, CF_HJ_VALUE_TABLE, WS_, false>(S, n, k, cfg, tab, io, st);                  \
    } while (0)
    if (cfg.obs_noise) {
        if (ws) CF_LAUNCH(true, true);
        CF_LAUNCH(true, false);
    }
    CF_LAUNCH(false, false);
#undef CF_LAUNCH
}
void launch_hj_signs(const float* V, uint8_t* P, cudaStream_t st) {
    int nb = (CF_HJ_CELLS + 3) / 4;
    k_hj_signs<<<blocks_for(nb, 256), 256, 0, st>>>(V, P, nb);
}
void launch_hj_lookup(const double* states, int n, const float* V, const uint8_t* P, int* idx, float* sign,
                      cudaStream_t st) {
    k_hj_lookup<<<blocks_for(n, 128), 128, 0, st>>>(states, n, V, P, idx, sign);
}
void launch_hj_lookup_f32(const float* states, int n, const DevCfg& cfg, const float* V, const uint8_t* P, const float* thr,
                          int* idx, float* sign, cudaStream_t st) {
    k_hj_lookup_f32<<<blocks_for(n, 128), 128, 0, st>>>(states, n, cfg, V, P, thr, idx, sign);
}
void launch_unpack_state(const float* S, int n, float* rec, cudaStream_t st) {
    k_unpack_state<<<blocks_for(n, 128), 128, 0, st>>>(S, n, rec);
}
void launch_pack_state(float* S, int n, const float* rec, cudaStream_t st) {
    k_pack_state<<<blocks_for(n, 128), 128, 0, st>>>(S, n, rec);
}
// pose rows (pos xyz, quat xyzw) of `count` envs for the rasteriser; idx == nullptr: envs 0..count-1
__global__ void k_gather_poses(const float* __restrict__ S, int n, const int* __restrict__ idx, int count, float* __restrict__ poses) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= count) return;
    int i = idx ? idx[j] : j;
    i = min(max(i, 0), n - 1);
    const EnvRef g = env_ref_global(const_cast<float*>(S), i);
#pragma unroll
    for (int k = 0; k < 3; ++k) poses[7 * j + k] = g.f(CF_S_POS + k);
#pragma unroll
    for (int k = 0; k < 4; ++k) poses[7 * j + 3 + k] = g.f(CF_S_QUAT + k);
}
void launch_gather_poses(const float* S, int n, const int* idx, int count, float* poses, cudaStream_t st) {
    k_gather_poses<<<blocks_for(count, 128), 128, 0, st>>>(S, n, idx, count, poses);
}
void launch_levels(const float* S, int n, float* out, cudaStream_t st) {
    k_levels<<<blocks_for(n, 128), 128, 0, st>>>(S, n, out);
}
void launch_gae(const float* rew, const float* val, const uint8_t* done, const float* end_val, const float* last_val,
                float gamma, float lam, float rew_scale, int T, int N, float* adv, float* tgt, float* disc_ret,
                cudaStream_t st) {
    const uintptr_t al = (uintptr_t)rew | (uintptr_t)val | (uintptr_t)end_val | (uintptr_t)last_val | (uintptr_t)adv | (uintptr_t)tgt |
                         (uintptr_t)disc_ret;
    if (N % 4 == 0 && (al & 15) == 0 && ((uintptr_t)done & 3) == 0) {
        k_gae4<<<blocks_for(N / 4, CF_GAE_THREADS), CF_GAE_THREADS, 0, st>>>(
            reinterpret_cast<const float4*>(rew), reinterpret_cast<const float4*>(val), reinterpret_cast<const uchar4*>(done),
            reinterpret_cast<const float4*>(end_val), reinterpret_cast<const float4*>(last_val), gamma, lam, rew_scale, T, N / 4,
            reinterpret_cast<float4*>(adv), reinterpret_cast<float4*>(tgt), reinterpret_cast<float4*>(disc_ret));
        return;
    }
    k_gae<<<blocks_for(N, 128), 128, 0, st>>>(rew, val, done, end_val, last_val, gamma, lam, rew_scale, T, N, adv, tgt,
                                              disc_ret);
}
void launch_debug_draws(const DevCfg& cfg, int first_env, int count, uint32_t ctr, uint32_t tag, int nzb, int nub, uint32_t* words,
                        float* draws, cudaStream_t st) {
    k_debug_draws<<<blocks_for(count * (nzb + nub), 128), 128, 0, st>>>(cfg, first_env, count, ctr, tag, nzb, nub, words, draws);
}
void launch_debug_euler(const float* quat, long long n, float* rpy_bullet, float* rpy_hj, cudaStream_t st) {
    k_debug_euler<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(reinterpret_cast<const float4*>(quat), n, rpy_bullet, rpy_hj);
}
void launch_debug_box_muller(const uint32_t* words, long long n, float* z, float* u, cudaStream_t st) {
    k_debug_box_muller<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(reinterpret_cast<const uint4*>(words), n,
                                                                   reinterpret_cast<float4*>(z), reinterpret_cast<float4*>(u));
}
void launch_pd_actions(const float* obs, int obs_dim, int off, int n, float* act, const PdGains& g, cudaStream_t st) {
    k_pd_actions<<<blocks_for(n, 128), 128, 0, st>>>(obs, obs_dim, off, n, act, g);
}

}  // namespace cf
