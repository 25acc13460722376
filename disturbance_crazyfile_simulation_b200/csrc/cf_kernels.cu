// cf_kernels.cu - kernels of the batched CrazyFlie hover environment and their launchers (sm_100a).
//
// Layout in HBM: per handle one float4 array S[19][n_envs] (structure-of-arrays in 128-bit
// groups, see cf_device.cuh), HJ tables (fp32 value tables and/or packed 2-bit sign tables),
// and caller-owned env-major I/O rows act[n][4], obs[n][obs_dim], rew[n], done[n], cost[n].
// One thread per env; no inter-thread communication on the step path.
#include "cf_device.cuh"
#include "cf_kernels.h"

namespace cf {

// obs rows are 136 B (34 floats) or 168 B (42 floats): 8-byte aligned -> 64-bit stores
template <int DIM>
__device__ __forceinline__ void store_row(float* __restrict__ dst, const float* v) {
    float2* p = reinterpret_cast<float2*>(dst);
#pragma unroll
    for (int k = 0; k < DIM / 2; ++k) p[k] = make_float2(v[2 * k], v[2 * k + 1]);
}

__global__ void k_init(float4* __restrict__ S, int n, DevCfg cfg) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Env e;
    init_env(e, cfg);
    store_env(e, S, n, i);
}

// constructor-time compute_observation() (base.py:141): gyro_bias <- sigma_b * N(0,1), event tag 3
__device__ __forceinline__ void consume_fresh(Env& e, const DevCfg& cfg, long long gid, bool noise) {
    if (!misc_fresh(e.misc)) return;
    if (noise && !cfg.injected) {
        Rng rng = make_rng(cfg, gid, 0u, 3u);
        float4 z = rng.normals(0);
        e.bias[0] = cfg.gyro_sb * z.x; e.bias[1] = cfg.gyro_sb * z.y; e.bias[2] = cfg.gyro_sb * z.z;
    }
    e.misc &= 0x7fffffffu;
}

template <bool NOISE>
__global__ void __launch_bounds__(128) k_reset(float4* __restrict__ S, int n, DevCfg cfg, const uint8_t* __restrict__ mask,
                                              float* __restrict__ obs, const float* inj_z, const float* inj_u) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (mask && !mask[i]) return;
    constexpr int OD = NOISE ? 34 : 42;
    Env e;
    load_env(e, S, n, i);
    long long gid = cfg.gid_base + i;
    consume_fresh(e, cfg, gid, NOISE);
    Draws<CF_RESET_NORMALS, CF_RESET_UNIFORMS> dw;
    if (cfg.injected) dw.load(inj_z, inj_u, n, i);
    else { dw.generate(make_rng(cfg, gid, e.rng_ctr, 2u)); e.rng_ctr += 1u; }
    float o[OD];
    reset_env<NOISE>(e, cfg, dw, o);
    store_env(e, S, n, i);
    store_row<OD>(obs + (size_t)i * OD, o);
}

// env.step for all envs, `k` control steps per launch with the state held in registers.
template <bool NOISE, int HJ_MODE>
__global__ void __launch_bounds__(128) k_step(float4* __restrict__ S, int n, int k, DevCfg cfg, DevTables tab, StepIO io) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    constexpr int OD = NOISE ? 34 : 42;
    Env e;
    load_env(e, S, n, i);
    long long gid = cfg.gid_base + i;
    consume_fresh(e, cfg, gid, NOISE);
    for (int t = 0; t < k; ++t) {
        size_t row = (size_t)t * n + i;
        float4 av = *reinterpret_cast<const float4*>(io.act + row * 4);
        float a[4] = {av.x, av.y, av.z, av.w};
        float hist[OD];
        StepOut so;
        {
            Draws<CF_STEP_NORMALS, CF_STEP_UNIFORMS> dw;
            if (cfg.injected) dw.load(io.inj_step_z, io.inj_step_u, n, i);
            else dw.generate(make_rng(cfg, gid, e.rng_ctr, 0u));
            so = step_env<NOISE, HJ_MODE>(e, cfg, tab, a, dw, hist);
        }
        io.rew[row] = so.rew;
        io.done[row] = (uint8_t)so.done;
        if (io.cost) io.cost[row] = so.cost;
        if (io.truncated) io.truncated[i] = (uint8_t)so.truncated;
        if (so.done) {
            io.fin_ret[i] = e.ep_ret;
            io.fin_len[i] = e.ep_len;
            atomicAdd(io.agg + 0, 1.0);
            atomicAdd(io.agg + 1, (double)e.ep_ret);
            atomicAdd(io.agg + 2, (double)e.ep_len);
        }
        if (so.done && cfg.auto_reset) {
            if (io.term_obs) store_row<OD>(io.term_obs + (size_t)i * OD, hist);
            Draws<CF_RESET_NORMALS, CF_RESET_UNIFORMS> dr;
            if (cfg.injected) dr.load(io.inj_reset_z, io.inj_reset_u, n, i);
            else dr.generate(make_rng(cfg, gid, e.rng_ctr, 1u));
            reset_env<NOISE>(e, cfg, dr, hist);
        }
        store_row<OD>(io.obs + row * OD, hist);
    }
    store_env(e, S, n, i);
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

// Grid.get_index + sign lookup for given float64 (roll,pitch,yaw,p,q,r) states
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

// env-major host record <-> device SoA groups
__global__ void k_unpack_state(const float4* __restrict__ S, int n, float4* __restrict__ rec) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
#pragma unroll
    for (int g = 0; g < kGroups; ++g) rec[(size_t)i * kGroups + g] = S[(size_t)g * n + i];
}
__global__ void k_pack_state(float4* __restrict__ S, int n, const float4* __restrict__ rec) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
#pragma unroll
    for (int g = 0; g < kGroups; ++g) S[(size_t)g * n + i] = rec[(size_t)i * kGroups + g];
}
__global__ void k_levels(const float4* __restrict__ S, int n, float* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 g = S[(size_t)(CF_S_MISC / 4) * n + i];
    out[i] = (float)misc_level(__float_as_uint(g.w)) / 10.0f;
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

// ---------------------------------------------------------------------------------- launchers
static inline int blocks_for(int n, int threads) { return (n + threads - 1) / threads; }

cudaError_t upload_constants(const double grid[6][CF_HJ_PTS], const double dx[6], const double cdf[21]) {
    cudaError_t err = cudaMemcpyToSymbol(c_grid, grid, sizeof(double) * 6 * CF_HJ_PTS);
    if (err != cudaSuccess) return err;
    err = cudaMemcpyToSymbol(c_dx, dx, sizeof(double) * 6);
    if (err != cudaSuccess) return err;
    return cudaMemcpyToSymbol(c_boltz_cdf, cdf, sizeof(double) * 21);
}

void launch_init(float4* S, int n, const DevCfg& cfg, cudaStream_t st) {
    k_init<<<blocks_for(n, 128), 128, 0, st>>>(S, n, cfg);
}
void launch_reset(float4* S, int n, const DevCfg& cfg, const uint8_t* mask, float* obs, const float* inj_z,
                  const float* inj_u, cudaStream_t st) {
    if (cfg.obs_noise) k_reset<true><<<blocks_for(n, 128), 128, 0, st>>>(S, n, cfg, mask, obs, inj_z, inj_u);
    else k_reset<false><<<blocks_for(n, 128), 128, 0, st>>>(S, n, cfg, mask, obs, inj_z, inj_u);
}
void launch_step(float4* S, int n, int k, const DevCfg& cfg, const DevTables& tab, const StepIO& io, cudaStream_t st) {
    dim3 g(blocks_for(n, 128)), b(128);
    const bool sign = cfg.hj_mode == CF_HJ_SIGN_TABLE;
    if (cfg.obs_noise) {
        if (sign) k_step<true, CF_HJ_SIGN_TABLE><<<g, b, 0, st>>>(S, n, k, cfg, tab, io);
        else k_step<true, CF_HJ_VALUE_TABLE><<<g, b, 0, st>>>(S, n, k, cfg, tab, io);
    } else {
        if (sign) k_step<false, CF_HJ_SIGN_TABLE><<<g, b, 0, st>>>(S, n, k, cfg, tab, io);
        else k_step<false, CF_HJ_VALUE_TABLE><<<g, b, 0, st>>>(S, n, k, cfg, tab, io);
    }
}
void launch_hj_signs(const float* V, uint8_t* P, cudaStream_t st) {
    int nb = (CF_HJ_CELLS + 3) / 4;
    k_hj_signs<<<blocks_for(nb, 256), 256, 0, st>>>(V, P, nb);
}
void launch_hj_lookup(const double* states, int n, const float* V, const uint8_t* P, int* idx, float* sign,
                      cudaStream_t st) {
    k_hj_lookup<<<blocks_for(n, 128), 128, 0, st>>>(states, n, V, P, idx, sign);
}
void launch_unpack_state(const float4* S, int n, float4* rec, cudaStream_t st) {
    k_unpack_state<<<blocks_for(n, 128), 128, 0, st>>>(S, n, rec);
}
void launch_pack_state(float4* S, int n, const float4* rec, cudaStream_t st) {
    k_pack_state<<<blocks_for(n, 128), 128, 0, st>>>(S, n, rec);
}
void launch_levels(const float4* S, int n, float* out, cudaStream_t st) {
    k_levels<<<blocks_for(n, 128), 128, 0, st>>>(S, n, out);
}
void launch_gae(const float* rew, const float* val, const uint8_t* done, const float* end_val, const float* last_val,
                float gamma, float lam, float rew_scale, int T, int N, float* adv, float* tgt, float* disc_ret,
                cudaStream_t st) {
    k_gae<<<blocks_for(N, 128), 128, 0, st>>>(rew, val, done, end_val, last_val, gamma, lam, rew_scale, T, N, adv, tgt,
                                              disc_ret);
}

}  // namespace cf
