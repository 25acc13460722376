// cf_ppo.cu - the PPO updates' forward + loss + backward as ONE kernel per optimiser step (sm_100a).
//
// What it replaces: one iteration of IWPGAlgorithm.update_policy_net (algs/iwpg/iwpg.py:448-459) up to the optimiser step:
// compute_loss_pi (algs/ppo/ppo.py:22-40: Gaussian log-prob of the stored actions under the current policy MLP 34 -> 50 -> 50 -> 4
// relu (core.py:227-260), ratio = exp(logp - logp_old), clipped surrogate -mean(min(ratio adv, clip(ratio) adv))) and
// loss.backward() for the six parameter tensors of pi.net.  The reference runs this as ~40 eager torch kernels over the full
// batch, 80 times per epoch; at 2.1 M samples each of them moves hundreds of MB of activations through HBM.  Here a CTA takes a
// tile of 128 samples through the whole chain with the activations in shared memory: three register-tiled products forward,
// the loss gradient per sample, two products for the hidden-layer gradients and three sample-reductions (A B^T over the tile)
// for the weight gradients, which stay in registers across all tiles of the (persistent) CTA; a second small kernel adds the
// per-CTA partials in a fixed order (deterministic).  float32 FFMA like torch (no TF32).
//
// Gradient of the surrogate exactly as autograd gives it: d min(s1, s2) goes to s1 = ratio adv where s1 < s2, to both halves
// where they are equal - which is whenever the ratio is inside the clip range, and both halves then lead to the ratio -, and is
// dropped where the clipped branch is the smaller one (clamp has no gradient outside its range).
#include "cf_device.cuh"
#include "cf_kernels.h"

namespace cf {

constexpr int kT = 128;          // samples per tile = threads per CTA
constexpr int kAS = 132;         // row stride of the activation buffers (floats): 16-byte aligned, rows 8 apart fall on distinct banks
constexpr int kH = 64;           // padded hidden width
constexpr int kWSt = 68;         // row stride of staged weights

struct PpoArgs {
    const float* obs;            // [B][OD] standardised observations (pre_process_data, iwpg.py:350-370)
    const float* act;            // [B][4]
    const float* adv;            // [B]
    const float* logp_old;       // [B]
    const float *w[3], *b[3];    // pi.net.{0,2,4}: torch layout [out][in]
    const float* log_std;        // [4]
    const float* tgt;            // value net: [B_all] targets
    const long long* idx;        // value net: [B] sample indices of this mini-batch into obs / tgt (torch int64), or null
    float* partial;              // [gridDim.x][kPartial] per-CTA gradient sums (+ loss, ratio sums)
    int B, h, n_tiles;
    float clip, inv_B;
};

// transposed staging for the forward products: W[out][in] -> Wt[k][j] (as cf_policy.cu)
__device__ __forceinline__ void stage_T(const float* __restrict__ W, const float* __restrict__ b, int out, int in, float* Wt,
                                        float* bs, int tid) {
    for (int q = tid; q < out * in; q += kT) {
        const int j = q / in, k = q - j * in;
        cp_async_4_nocommit(Wt + k * kWSt + j, W + q);
    }
    if (b && tid < out) cp_async_4_nocommit(bs + tid, b + tid);
    cp_async_commit();
    for (int q = tid; q < in * (kH - out); q += kT) {
        const int k = q / (kH - out), j = out + q % (kH - out);
        Wt[k * kWSt + j] = 0.0f;
    }
    if (b && tid >= out && tid < kH) bs[tid] = 0.0f;
}
// natural staging for the backward products: W[out][in] -> Wn[j][k] (reduce over the OUTPUT index j), columns >= in zero
__device__ __forceinline__ void stage_N(const float* __restrict__ W, int out, int in, float* Wn, int tid) {
    for (int q = tid; q < out * in; q += kT) {
        const int j = q / in, k = q - j * in;
        cp_async_4_nocommit(Wn + j * kWSt + k, W + q);
    }
    cp_async_commit();
    for (int q = tid; q < out * (kH - in); q += kT) {
        const int j = q / (kH - in), k = in + q % (kH - in);
        Wn[j * kWSt + k] = 0.0f;
    }
}

// O[j][e] = f(sum_k A[k][e] Wt[k][j] + bias[j]) for 128 samples x 64 (padded) outputs; thread tile 8 samples x 8 outputs.
// MODE 0: relu(x), MODE 2: tanh(x) (forward, with bias);  MODE 1: x * (M[j][e] > 0), MODE 3: x * (1 - M[j][e]^2) with M = O
// (backward through relu / tanh, in place: the stored activation is read before the tile is written; no bias).
template <int MODE>
__device__ __forceinline__ void dense128(const float* __restrict__ A, int K, const float* __restrict__ Wt, const float* __restrict__ bs,
                                         float* __restrict__ O, int n_out, int tid) {
    const int te = tid & 15, tn = tid >> 4;
    if (tn * 8 >= n_out) return;
    float acc[8][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float bj = (MODE == 0 || MODE == 2) ? bs[tn * 8 + j] : 0.0f;
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[j][e] = bj;
    }
    const float4* a4 = reinterpret_cast<const float4*>(A + te * 8);
    const float4* w4 = reinterpret_cast<const float4*>(Wt + tn * 8);
#pragma unroll 2
    for (int k = 0; k < K; ++k) {
        const float4 a0 = a4[k * (kAS / 4)], a1 = a4[k * (kAS / 4) + 1];
        const float4 w0 = w4[k * (kWSt / 4)], w1 = w4[k * (kWSt / 4) + 1];
        const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[j][e] = fmaf(a[e], w[j], acc[j][e]);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float4* dst = reinterpret_cast<float4*>(O + (tn * 8 + j) * kAS + te * 8);
        float o[8];
        if (MODE == 0) {
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] = fmaxf(acc[j][e], 0.0f);
        } else if (MODE == 2) {
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] = tanhf(acc[j][e]);
        } else {
            const float4 m0 = dst[0], m1 = dst[1];
            const float m[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] = MODE == 1 ? (m[e] > 0.0f ? acc[j][e] : 0.0f) : acc[j][e] * (1.0f - m[e] * m[e]);
        }
        dst[0] = make_float4(o[0], o[1], o[2], o[3]);
        dst[1] = make_float4(o[4], o[5], o[6], o[7]);
    }
}

// acc[r][c] += sum_e P[ga + 8 r][e] Q[gb + 16 c][e] over the tile's 128 samples: the weight gradient dW[p][q] of a layer with
// output-gradient rows P and input rows Q.  Thread (ga = tid & 7, gb = tid >> 3) owns rows ga, ga + 8, ... of P (RA of them) and
// rows gb, gb + 16, ... of Q (RB of them): with the row stride 132 the eight ga of a quarter-warp read distinct banks.
template <int RA, int RB>
__device__ __forceinline__ void outer128(const float* __restrict__ P, const float* __restrict__ Q, float (&acc)[RA][RB], int tid) {
    const int ga = tid & 7, gb = tid >> 3;
#pragma unroll 2
    for (int e = 0; e < kT; e += 4) {
        float4 p[RA], q[RB];
#pragma unroll
        for (int r = 0; r < RA; ++r) p[r] = *reinterpret_cast<const float4*>(P + (ga + 8 * r) * kAS + e);
#pragma unroll
        for (int c = 0; c < RB; ++c) q[c] = *reinterpret_cast<const float4*>(Q + (gb + 16 * c) * kAS + e);
#pragma unroll
        for (int r = 0; r < RA; ++r)
#pragma unroll
            for (int c = 0; c < RB; ++c) {
                acc[r][c] = fmaf(p[r].x, q[c].x, acc[r][c]); acc[r][c] = fmaf(p[r].y, q[c].y, acc[r][c]);
                acc[r][c] = fmaf(p[r].z, q[c].z, acc[r][c]); acc[r][c] = fmaf(p[r].w, q[c].w, acc[r][c]);
            }
    }
}

// per-CTA partial sums: dW1 [64][40] (k, i: rows of X padded to 40), dW2 [64][64] (j, k), dW3 [4][64], db1 [64], db2 [64], db3 [4],
// loss sum, ratio sum

constexpr int kOffW1 = 0, kOffW2 = kOffW1 + kH * 48, kOffW3 = kOffW2 + kH * kH, kOffB1 = kOffW3 + 4 * kH, kOffB2 = kOffB1 + kH,
              kOffB3 = kOffB2 + kH, kOffLoss = kOffB3 + 4, kPartial = kOffLoss + 4;

// NET 0: the policy net (relu, 4 outputs, clipped surrogate);  NET 1: the value net (tanh, 1 output, mean squared error against
// tgt over the mini-batch idx[0..B) - update_value_net, iwpg.py:478-510 with compute_loss_v, iwpg.py:250-255).
template <int OD, int NET>
__global__ void __launch_bounds__(kT, 2) k_ppo_grad(const __grid_constant__ PpoArgs pa) {
    constexpr int XR = 48;                       // observation rows padded to 3 x 16 (rows >= OD stay zero)
    constexpr int NO = NET == 0 ? 4 : 1;         // outputs of the last layer
    constexpr int FWD = NET == 0 ? 0 : 2, BWD = NET == 0 ? 1 : 3;
    static_assert(OD <= XR, "observation width");
    extern __shared__ __align__(16) float sm[];
    float* X = sm;                               // [XR][kAS]
    float* H1 = X + XR * kAS;                    // [64][kAS] hidden 1, later its gradient
    float* H2 = H1 + kH * kAS;                   // [64][kAS] hidden 2, later its gradient
    float* DM = H2 + kH * kAS;                   // [4][kAS]  dL/dmu (value net: row 0 = dL/dv)
    float* Wb = DM + 4 * kAS;                    // [64][kWSt] staged weights
    float* bs = Wb + kH * kWSt;                  // [64]
    float* W3s = bs + kH;                        // [4][64] last layer, natural layout
    float* red = W3s + 4 * kH;                   // [kT] scratch for the two scalar sums
    const int tid = threadIdx.x, h = pa.h;
    // weight-gradient accumulators, kept in registers over all tiles of this CTA
    float g1[8][3], g2[8][4];                    // dW1 rows ga + 8 r (k), cols gb + 16 c (i < 48);  dW2 rows ga + 8 r (j), cols gb + 16 c (k)
    float g3[2], gb1 = 0.f, gb2 = 0.f, gb3 = 0.f, loss_sum = 0.f, ratio_sum = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c) g1[r][c] = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) g2[r][c] = 0.f;
    }
    g3[0] = g3[1] = 0.f;
    for (int q = tid; q < (XR - OD) * kAS; q += kT) X[OD * kAS + q] = 0.0f;            // padding rows of X
    for (int q = tid; q < kH * kAS; q += kT) { H1[q] = 0.0f; H2[q] = 0.0f; }            // rows the products never write stay zero
    for (int q = tid; q < 4 * kH; q += kT) W3s[q] = ((q & 63) < h && (q >> 6) < NO) ? __ldg(pa.w[2] + (q >> 6) * h + (q & 63)) : 0.0f;
    float sd[4], ls[4];
    if (NET == 0) {
#pragma unroll
        for (int a = 0; a < 4; ++a) { ls[a] = __ldg(pa.log_std + a); sd[a] = expf(ls[a]); }
    }
    for (int tile = blockIdx.x; tile < pa.n_tiles; tile += gridDim.x) {
        const int s = tile * kT + tid;               // position in the (mini-)batch
        const bool valid = s < pa.B;
        const long long i = !valid ? 0 : (NET == 1 && pa.idx ? __ldg(pa.idx + s) : (long long)s);     // sample row
        __syncthreads();                             // previous tile's products are done with the buffers
        stage_T(pa.w[0], pa.b[0], h, OD, Wb, bs, tid);
        {
            const float2* row = reinterpret_cast<const float2*>(pa.obs + (size_t)i * OD);
#pragma unroll
            for (int k2 = 0; k2 < OD / 2; ++k2) {
                const float2 v = valid ? __ldg(row + k2) : make_float2(0.f, 0.f);
                X[(2 * k2) * kAS + tid] = v.x; X[(2 * k2 + 1) * kAS + tid] = v.y;
            }
        }
        cp_async_wait_all();
        __syncthreads();
        dense128<FWD>(X, OD, Wb, bs, H1, h, tid);    // h1 = f(W1 x + b1)
        __syncthreads();
        stage_T(pa.w[1], pa.b[1], h, h, Wb, bs, tid);
        cp_async_wait_all();
        __syncthreads();
        dense128<FWD>(H1, h, Wb, bs, H2, h, tid);    // h2 = f(W2 h1 + b2)
        __syncthreads();
        stage_N(pa.w[1], h, h, Wb, tid);             // W2 in natural layout for the backward product (lands under the loss)
        // ---- outputs, loss and dL/d(output) of this thread's sample
        float mu[NO];
#pragma unroll
        for (int a = 0; a < NO; ++a) mu[a] = __ldg(pa.b[2] + a);
        for (int k = 0; k < h; ++k) {
            const float x = H2[k * kAS + tid];
#pragma unroll
            for (int a = 0; a < NO; ++a) mu[a] = fmaf(x, W3s[a * kH + k], mu[a]);
        }
        float dmu[NO];
#pragma unroll
        for (int a = 0; a < NO; ++a) dmu[a] = 0.f;
        if (valid) {
            if constexpr (NET == 0) {
                const float4 av = __ldg(reinterpret_cast<const float4*>(pa.act) + i);
                const float act[4] = {av.x, av.y, av.z, av.w};
                const float adv = __ldg(pa.adv + i);
                float logp = 0.f;
#pragma unroll
                for (int a = 0; a < 4; ++a) {
                    const float d = act[a] - mu[a];
                    logp += -(d * d) / (2.0f * sd[a] * sd[a]) - ls[a] - 0.91893853320467274f;
                }
                const float ratio = expf(logp - __ldg(pa.logp_old + i));
                const float s1 = ratio * adv, s2 = fminf(fmaxf(ratio, 1.0f - pa.clip), 1.0f + pa.clip) * adv;
                loss_sum += -fminf(s1, s2);
                ratio_sum += ratio;
                const bool inside = ratio >= 1.0f - pa.clip && ratio <= 1.0f + pa.clip;
                const float w = (inside || s1 < s2) ? 1.0f : (s1 == s2 ? 0.5f : 0.0f);
                const float glp = -adv * ratio * w * pa.inv_B;                     // dL / dlogp
#pragma unroll
                for (int a = 0; a < 4; ++a) dmu[a] = glp * (act[a] - mu[a]) / (sd[a] * sd[a]);
            } else {
                const float d = mu[0] - __ldg(pa.tgt + i);
                loss_sum += d * d;
                dmu[0] = 2.0f * d * pa.inv_B;
            }
        }
#pragma unroll
        for (int a = 0; a < NO; ++a) DM[a * kAS + tid] = dmu[a];
        cp_async_wait_all();
        __syncthreads();
        // ---- last layer: dW3[a][k] += sum_e dmu[a][e] h2[k][e]  (NO x 64 outputs, up to two per thread), db3
        {
#pragma unroll
            for (int o = 0; o < 2; ++o) {
                const int a = (tid + o * kT) >> 6, k = (tid + o * kT) & 63;
                float sum = 0.f;
                if (k < h && a < NO)
                    for (int e = 0; e < kT; e += 4) {
                        const float4 p = *reinterpret_cast<const float4*>(DM + a * kAS + e), q = *reinterpret_cast<const float4*>(H2 + k * kAS + e);
                        sum = fmaf(p.x, q.x, sum); sum = fmaf(p.y, q.y, sum); sum = fmaf(p.z, q.z, sum); sum = fmaf(p.w, q.w, sum);
                    }
                g3[o] += sum;
            }
            if (tid < NO) { float sum = 0.f; for (int e = 0; e < kT; ++e) sum += DM[tid * kAS + e]; gb3 += sum; }
        }
        __syncthreads();
        // ---- dh2[k][e] = f'(h2) sum_a dmu[a][e] W3[a][k], in place over H2 (this thread's sample column)
        for (int k = 0; k < h; ++k) {
            const float x = H2[k * kAS + tid];
            float sum = 0.f;
#pragma unroll
            for (int a = 0; a < NO; ++a) sum = fmaf(dmu[a], W3s[a * kH + k], sum);
            H2[k * kAS + tid] = NET == 0 ? (x > 0.f ? sum : 0.f) : sum * (1.0f - x * x);
        }
        __syncthreads();
        outer128<8, 4>(H2, H1, g2, tid);             // dW2[j][k] += sum_e dh2[j][e] h1[k][e]
        if (tid < kH) { float sum = 0.f; for (int e = 0; e < kT; e += 4) { const float4 p = *reinterpret_cast<const float4*>(H2 + tid * kAS + e); sum += (p.x + p.y) + (p.z + p.w); } gb2 += sum; }
        __syncthreads();
        dense128<BWD>(H2, h, Wb, nullptr, H1, h, tid); // dh1[k][e] = f'(h1) sum_j dh2[j][e] W2[j][k], in place over H1
        __syncthreads();
        outer128<8, 3>(H1, X, g1, tid);              // dW1[k][i] += sum_e dh1[k][e] x[i][e]
        if (tid < kH) { float sum = 0.f; for (int e = 0; e < kT; e += 4) { const float4 p = *reinterpret_cast<const float4*>(H1 + tid * kAS + e); sum += (p.x + p.y) + (p.z + p.w); } gb1 += sum; }
    }
    // ---- this CTA's partial sums
    float* out = pa.partial + (size_t)blockIdx.x * kPartial;
    const int ga = tid & 7, gb = tid >> 3;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c) out[kOffW1 + (ga + 8 * r) * 48 + gb + 16 * c] = g1[r][c];
#pragma unroll
        for (int c = 0; c < 4; ++c) out[kOffW2 + (ga + 8 * r) * kH + gb + 16 * c] = g2[r][c];
    }
    out[kOffW3 + tid] = g3[0]; out[kOffW3 + kT + tid] = g3[1];
    if (tid < kH) { out[kOffB1 + tid] = gb1; out[kOffB2 + tid] = gb2; }
    if (tid < 4) out[kOffB3 + tid] = tid < NO ? gb3 : 0.f;
    __syncthreads();
    red[tid] = loss_sum;
    __syncthreads();
    if (tid == 0) { float sum = 0.f; for (int e = 0; e < kT; ++e) sum += red[e]; out[kOffLoss] = sum; }
    __syncthreads();
    red[tid] = ratio_sum;
    __syncthreads();
    if (tid == 0) { float sum = 0.f; for (int e = 0; e < kT; ++e) sum += red[e]; out[kOffLoss + 1] = sum; }
}

// partial[c][.] summed over the CTAs in index order (deterministic) into the six gradient tensors (torch layouts) + loss / ratio
__global__ void __launch_bounds__(256) k_ppo_reduce(const float* __restrict__ partial, int n_cta, int od, int h, int n_out, float inv_B,
                                                    float* __restrict__ gw1, float* __restrict__ gb1, float* __restrict__ gw2,
                                                    float* __restrict__ gb2, float* __restrict__ gw3, float* __restrict__ gb3,
                                                    float* __restrict__ loss_ratio) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= kPartial) return;
    float s = 0.f;
    for (int c = 0; c < n_cta; ++c) s += partial[(size_t)c * kPartial + q];
    if (q < kOffW2) { const int k = q / 48, i = q % 48; if (k < h && i < od) gw1[k * od + i] = s; }
    else if (q < kOffW3) { const int j = (q - kOffW2) / kH, k = (q - kOffW2) % kH; if (j < h && k < h) gw2[j * h + k] = s; }
    else if (q < kOffB1) { const int a = (q - kOffW3) / kH, k = (q - kOffW3) % kH; if (k < h && a < n_out) gw3[a * h + k] = s; }
    else if (q < kOffB2) { if (q - kOffB1 < h) gb1[q - kOffB1] = s; }
    else if (q < kOffB3) { if (q - kOffB2 < h) gb2[q - kOffB2] = s; }
    else if (q < kOffLoss) { if (q - kOffB3 < n_out) gb3[q - kOffB3] = s; }
    else if (q < kOffLoss + 2) loss_ratio[q - kOffLoss] = s * inv_B;       // mean surrogate loss, mean ratio
}

size_t ppo_smem_bytes() { return (size_t)(48 * kAS + 2 * kH * kAS + 4 * kAS + kH * kWSt + kH + 4 * kH + kT) * sizeof(float); }
int ppo_partial_floats() { return kPartial; }

template <int OD, int NET>
static cudaError_t launch_one(const PpoArgs& pa, int grid, size_t smem, cudaStream_t st) {
    static bool configured[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        cudaError_t err = cudaFuncSetAttribute(k_ppo_grad<OD, NET>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return err;
        configured[dev] = true;
    }
    k_ppo_grad<OD, NET><<<grid, kT, smem, st>>>(pa);
    return cudaSuccess;
}

// a.net 0: policy (act, adv, logp_old, log_std, clip used);  1: value (tgt, idx used).  Tiles are dealt round-robin to
// ceil(n_tiles / waves) persistent CTAs with waves = ceil(n_tiles / n_cta): every CTA gets waves or waves - 1 tiles.
cudaError_t launch_ppo_grad(const PpoHostArgs& a, cudaStream_t st) {
    PpoArgs pa;
    pa.obs = a.obs; pa.act = a.act; pa.adv = a.adv; pa.logp_old = a.logp_old; pa.tgt = a.tgt; pa.idx = a.idx;
    for (int l = 0; l < 3; ++l) { pa.w[l] = a.w[l]; pa.b[l] = a.b[l]; }
    pa.log_std = a.log_std; pa.partial = a.partial; pa.B = a.B; pa.h = a.h;
    pa.n_tiles = (a.B + kT - 1) / kT; pa.clip = a.clip; pa.inv_B = 1.0f / (float)a.B;
    const size_t smem = ppo_smem_bytes();
    const int waves = (pa.n_tiles + a.n_cta - 1) / a.n_cta;
    const int grid = (pa.n_tiles + waves - 1) / waves;
    cudaError_t err;
    if (a.net == 0) err = a.obs_dim == 34 ? launch_one<34, 0>(pa, grid, smem, st) : launch_one<42, 0>(pa, grid, smem, st);
    else err = a.obs_dim == 34 ? launch_one<34, 1>(pa, grid, smem, st) : launch_one<42, 1>(pa, grid, smem, st);
    if (err != cudaSuccess) return err;
    k_ppo_reduce<<<(kPartial + 255) / 256, 256, 0, st>>>(a.partial, grid, a.obs_dim, a.h, a.net == 0 ? 4 : 1, pa.inv_B, a.gw[0], a.gb[0],
                                                         a.gw[1], a.gb[1], a.gw[2], a.gb[2], a.loss_ratio);
    return cudaSuccess;
}

// ---- the allocation-free variant (cf_ppo_pi_grad_static_dev) ----------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot allocate (code under
// CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The per-CTA partial-sum records live in this
// file-scope array instead of a caller workspace: kStaticCtas x kPartial floats = 148 x 7560 x 4 B = 4.27 MiB, held by every
// GPU that loads the library until the process exits.  Every launch overwrites each record its reduction later reads (all
// kPartial words of records 0 .. grid - 1), so nothing has to start at zero and a call does not see the previous call's
// contents; all calls share the one array, so calls that overlap on the GPU (different streams, other threads) race.
constexpr int kStaticCtas = 148;                          // one persistent CTA per SM
constexpr long long kStaticMaxB = 1ll << 22;              // 4 194 304 samples: 222 tiles per CTA, twice the PPO bench's batch
static __device__ __align__(256) float g_ppo_static_scratch[kStaticCtas * kPartial];

int ppo_static_max_ctas() { return kStaticCtas; }
long long ppo_static_max_samples() { return kStaticMaxB; }

// In host code g_ppo_static_scratch is a stand-in for the symbol, not a device address: the address is resolved once per
// device, with that device current, and cached.  a.partial and a.n_cta are ignored.
cudaError_t launch_ppo_pi_grad_static(const PpoHostArgs& a, cudaStream_t st) {
    static float* scratch[64] = {};
    int dev = -1;
    cudaError_t err = cudaGetDevice(&dev);
    if (err != cudaSuccess) return err;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!scratch[dev]) {
        float* ptr = nullptr;
        err = cudaGetSymbolAddress((void**)&ptr, g_ppo_static_scratch);
        if (err != cudaSuccess) return err;
        if (!ptr) return cudaErrorInvalidSymbol;
        scratch[dev] = ptr;
    }
    PpoHostArgs b = a;
    b.net = 0; b.partial = scratch[dev]; b.n_cta = kStaticCtas;
    return launch_ppo_grad(b, st);
}

}  // namespace cf
