// cf_policy.cu - the rollout's two networks as ONE kernel (sm_100a).
//
// What it replaces: ActorCritic.step of the reference (algs/core.py:371-394) as IWPGAlgorithm.roll_out calls it once per
// control step (algs/iwpg/iwpg.py:372-410): obs_oms(obs) (utils/online_mean_std.py:32-48) -> value MLP 34 -> 64 -> 64 -> 1
// with tanh (core.py:298-312) -> policy MLP 34 -> 50 -> 50 -> 4 with relu (core.py:227-260) -> a = mu + exp(log_std) * N(0,1)
// and its log-probability summed over the action dimensions (core.py:257-266); eval mode returns mu (core.py:284-290).
// float32 FFMA like the reference's torch modules (no TF32: the layers are far too small for tensor cores to matter and the
// parity bound against torch is 1e-5).
//
// One CTA = 128 threads = one tile of 128 envs (the env kernel's tiling, so the sampling noise can be keyed by the env's own
// event counter).  Activations live in shared memory k-major ([feature][env]); every layer is a register-tiled product:
// a thread owns 8 envs x 8 neurons (64 accumulators), per input feature two 128-bit loads of activations and two of weights
// feed 64 FFMAs.  The layer's weights are staged transposed ([in][out], padded to 64 neurons) from the torch layout
// ([out][in], row-major) that the caller passes - no packed copy to keep in step with the optimiser.
#include "cf_device.cuh"
#include "cf_kernels.h"

namespace cf {

constexpr int kPolTile = CF_TILE;        // envs per CTA
constexpr int kPolMaxH = 64;             // hidden width limit (reference defaults: 50 and 64)
constexpr int kWS = 68;                  // row stride of the staged weights (floats): 16-byte aligned rows, spreads banks

struct PolicyArgs {
    const float* obs;                    // [n][obs_dim]
    const float *obs_mean, *obs_std;     // [obs_dim] or null (no standardisation)
    const float *pi_w[3], *pi_b[3];      // torch Linear layout [out][in]
    const float *v_w[3], *v_b[3];
    const float* log_std;                // [4]
    const float* S;                      // env state tiles (event counters) or null
    float *act, *val, *logp;             // [n][4], [n], [n]
    int n, pi_h, v_h, deterministic;
    float eps;
};

__device__ __forceinline__ float tanh_fast(float x) {
    // 1 - 2 / (1 + e^(2x)): exact limits at +-inf, absolute error < 5e-7 (ex2.approx + rcp.approx)
    return 1.0f - 2.0f * frcp(1.0f + __expf(2.0f * x));
}

// weights of one layer, torch layout W[out][in] -> Wt[k][j] (k < in rows, 64 columns, columns >= out zero), bias -> bs[64].
// Asynchronous 4-byte copies (LDGSTS): a thread issues its ~32 elements back to back and waits once - with ordinary loads
// each element's store waited for its own load (ncu round 2, first build: long scoreboard 22 % of the stall samples, 70 us per
// launch of which ~45 us were six serialised stagings).  Call stage_wait() + a barrier before the weights are read.
__device__ __forceinline__ void stage_weights(const float* __restrict__ W, const float* __restrict__ b, int out, int in,
                                              float* Wt, float* bs, int tid) {
    for (int q = tid; q < out * in; q += kPolTile) {
        const int j = q / in, k = q - j * in;
        cp_async_4_nocommit(Wt + k * kWS + j, W + q);
    }
    if (tid < out) cp_async_4_nocommit(bs + tid, b + tid);
    cp_async_commit();
    for (int q = tid; q < in * (kPolMaxH - out); q += kPolTile) {         // zero padding columns
        const int k = q / (kPolMaxH - out), j = out + q % (kPolMaxH - out);
        Wt[k * kWS + j] = 0.0f;
    }
    if (tid >= out && tid < kPolMaxH) bs[tid] = 0.0f;
}
__device__ __forceinline__ void stage_wait() { cp_async_wait_all(); }

// out[j][e] = act(sum_k A[k][e] * Wt[k][j] + b[j]) for the CTA's 128 envs and 64 (padded) neurons.  ACT: 0 relu, 1 tanh.
template <int ACT>
__device__ __forceinline__ void dense_layer(const float* __restrict__ A, int K, const float* __restrict__ Wt,
                                            const float* __restrict__ bs, float* __restrict__ O, int n_out, int tid) {
    const int te = tid & 15, tn = tid >> 4;              // 16 env groups x 8 neuron groups
    if (tn * 8 >= n_out) return;                         // neuron group entirely in the padding
    float acc[8][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float bj = bs[tn * 8 + j];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[j][e] = bj;
    }
    const float4* a4 = reinterpret_cast<const float4*>(A + te * 8);
    const float4* w4 = reinterpret_cast<const float4*>(Wt + tn * 8);
#pragma unroll 2
    for (int k = 0; k < K; ++k) {
        const float4 a0 = a4[k * (kPolTile / 4)], a1 = a4[k * (kPolTile / 4) + 1];
        const float4 w0 = w4[k * (kWS / 4)], w1 = w4[k * (kWS / 4) + 1];
        const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[j][e] = fmaf(a[e], w[j], acc[j][e]);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = ACT == 0 ? fmaxf(acc[j][e], 0.0f) : tanh_fast(acc[j][e]);
        float4* dst = reinterpret_cast<float4*>(O + (tn * 8 + j) * kPolTile + te * 8);
        dst[0] = make_float4(o[0], o[1], o[2], o[3]);
        dst[1] = make_float4(o[4], o[5], o[6], o[7]);
    }
}

template <int OD>
__global__ void __launch_bounds__(kPolTile, 2) k_policy(const __grid_constant__ PolicyArgs pa, const __grid_constant__ DevCfg cfg) {
    extern __shared__ __align__(16) float sm[];
    float* A0 = sm;                                   // [OD][128] standardised observation
    float* A1 = A0 + OD * kPolTile;                   // [64][128]
    float* A2 = A1 + kPolMaxH * kPolTile;             // [64][128]
    float* Wt = A2 + kPolMaxH * kPolTile;             // [64][kWS]
    float* bs = Wt + kPolMaxH * kWS;                  // [64]
    float* nrm = bs + kPolMaxH;                       // [2][OD] mean, 1 / (std + eps)
    const int tid = threadIdx.x, tile = blockIdx.x, i = tile * kPolTile + tid;
    const bool valid = i < pa.n;
    for (int k = tid; k < OD; k += kPolTile) {
        nrm[k] = pa.obs_mean ? __ldg(pa.obs_mean + k) : 0.0f;
        nrm[OD + k] = pa.obs_std ? __fdiv_rn(1.0f, __ldg(pa.obs_std + k) + pa.eps) : 1.0f;
    }
    stage_weights(pa.v_w[0], pa.v_b[0], pa.v_h, OD, Wt, bs, tid);
    __syncthreads();                                  // nrm[] in place (the weights are waited for below)
    {   // observation row of this thread's env -> k-major, standardised (online_mean_std.py:32-48)
        const float2* row = reinterpret_cast<const float2*>(pa.obs + (size_t)i * OD);
#pragma unroll
        for (int k2 = 0; k2 < OD / 2; ++k2) {
            const float2 v = valid ? __ldg(row + k2) : make_float2(0.f, 0.f);
            A0[(2 * k2) * kPolTile + tid] = (v.x - nrm[2 * k2]) * nrm[OD + 2 * k2];
            A0[(2 * k2 + 1) * kPolTile + tid] = (v.y - nrm[2 * k2 + 1]) * nrm[OD + 2 * k2 + 1];
        }
    }
    stage_wait();
    __syncthreads();
    // ---- value net: tanh, tanh, linear (core.py:298-312)
    dense_layer<1>(A0, OD, Wt, bs, A1, pa.v_h, tid);
    __syncthreads();
    stage_weights(pa.v_w[1], pa.v_b[1], pa.v_h, pa.v_h, Wt, bs, tid);
    stage_wait();
    __syncthreads();
    dense_layer<1>(A1, pa.v_h, Wt, bs, A2, pa.v_h, tid);
    __syncthreads();
    {
        float v = __ldg(pa.v_b[2]);
        for (int k = 0; k < pa.v_h; ++k) v = fmaf(A2[k * kPolTile + tid], __ldg(pa.v_w[2] + k), v);
        if (valid) pa.val[i] = v;
    }
    // ---- policy net: relu, relu, linear (core.py:227-260)
    stage_weights(pa.pi_w[0], pa.pi_b[0], pa.pi_h, OD, Wt, bs, tid);
    stage_wait();
    __syncthreads();
    dense_layer<0>(A0, OD, Wt, bs, A1, pa.pi_h, tid);
    __syncthreads();
    stage_weights(pa.pi_w[1], pa.pi_b[1], pa.pi_h, pa.pi_h, Wt, bs, tid);
    stage_wait();
    __syncthreads();
    dense_layer<0>(A1, pa.pi_h, Wt, bs, A2, pa.pi_h, tid);
    __syncthreads();
    float mu[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) mu[j] = __ldg(pa.pi_b[2] + j);
    for (int k = 0; k < pa.pi_h; ++k) {
        const float a = A2[k * kPolTile + tid];
#pragma unroll
        for (int j = 0; j < 4; ++j) mu[j] = fmaf(a, __ldg(pa.pi_w[2] + j * pa.pi_h + k), mu[j]);
    }
    if (!valid) return;
    float act[4] = {mu[0], mu[1], mu[2], mu[3]};
    float logp = 1.0f;                                  // eval mode: predict() returns ones (core.py:284-290)
    if (!pa.deterministic) {
        // exploration noise of this env's NEXT step event: Philox key = the env handle's seed, counter = (global env id, the
        // env's event counter, block 0, tag 4) - independent of how envs are sharded, distinct from the env's own tags 0..3
        const uint32_t ctr = __float_as_uint(__ldcg(pa.S + (size_t)tile * (CF_STATE_WORDS * CF_TILE) + cf_dev_word(CF_S_RNG_CTR) * CF_TILE + tid));
        const float4 z4 = make_rng(cfg.gid_base + i, ctr, 4u).normals(cfg, 0);
        const float z[4] = {z4.x, z4.y, z4.z, z4.w};
        logp = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float ls = __ldg(pa.log_std + j), sd = expf(ls);
            act[j] = mu[j] + sd * z[j];
            const float d = act[j] - mu[j];
            // Normal(mu, std).log_prob(a) summed over the action dimensions (core.py:257-259)
            logp += -(d * d) / (2.0f * sd * sd) - ls - 0.91893853320467274f;
        }
    }
    reinterpret_cast<float4*>(pa.act)[i] = make_float4(act[0], act[1], act[2], act[3]);
    pa.logp[i] = logp;
}

size_t policy_smem_bytes(int obs_dim) {
    return (size_t)(obs_dim * kPolTile + 2 * kPolMaxH * kPolTile + kPolMaxH * kWS + kPolMaxH + 2 * obs_dim) * sizeof(float);
}

cudaError_t launch_policy(const PolicyHostArgs& h, const DevCfg& cfg, cudaStream_t st) {
    PolicyArgs pa;
    pa.obs = h.obs; pa.obs_mean = h.obs_mean; pa.obs_std = h.obs_std;
    for (int l = 0; l < 3; ++l) { pa.pi_w[l] = h.pi_w[l]; pa.pi_b[l] = h.pi_b[l]; pa.v_w[l] = h.v_w[l]; pa.v_b[l] = h.v_b[l]; }
    pa.log_std = h.log_std; pa.S = h.S; pa.act = h.act; pa.val = h.val; pa.logp = h.logp;
    pa.n = h.n; pa.pi_h = h.pi_h; pa.v_h = h.v_h; pa.deterministic = h.deterministic; pa.eps = h.eps;
    static bool configured[64][2] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    const int which = h.obs_dim == 34 ? 0 : 1;
    const size_t smem = policy_smem_bytes(h.obs_dim);
    if (dev >= 0 && dev < 64 && !configured[dev][which]) {
        cudaError_t err = which == 0 ? cudaFuncSetAttribute(k_policy<34>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                                     : cudaFuncSetAttribute(k_policy<42>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return err;
        configured[dev][which] = true;
    }
    const int grid = (h.n + kPolTile - 1) / kPolTile;
    if (which == 0) k_policy<34><<<grid, kPolTile, smem, st>>>(pa, cfg);
    else k_policy<42><<<grid, kPolTile, smem, st>>>(pa, cfg);
    return cudaSuccess;
}

}  // namespace cf
