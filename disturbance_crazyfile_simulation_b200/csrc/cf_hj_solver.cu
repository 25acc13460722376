// cf_hj_solver.cu - FasTrack HJ value-table generator on the GPU (sm_100a).
//
// What it replaces: the HeteroCL-built CPU kernel of optimized_dp that writes fastrack_{level}_15x15.npy
// (adversarial_generation/FasTrack_data/Fastrack.py:281-320 -> odp/solver.py:83-233 -> odp/computeGraphs/graph_6D.py,
// derivatives odp/spatialDerivatives/first_orderENO6D.py).  One update of the reference is two sweeps over the
// [pts]^6 grid with two global reductions between them; here it is
//   k_minmax  : one-sided derivatives of every cell (13-point stencil), block min / max per dimension
//   k_choose  : (one CTA) global min / max -> the optimal (u, d) at the extreme derivatives, the twelve torque terms
//               (u + d) / I of the 4 corner combinations, the global max alpha, the step bound, t0 += dt - all on
//               the device, no host round trip inside an update
//   k_update  : derivatives again, Hamiltonian, dissipation, V' = max(V + rate * dt, l0) into the other buffer
// Hoisted without changing a bit of the result: alpha of the three angle dimensions does not depend on (u, d) or on
// V, so its global maximum is taken once at creation; alpha of the three rate dimensions is |(c x) y + (u + d) / I|
// with (u + d) / I a per-update scalar, so its global maximum runs over pts^2 pairs inside k_choose.
// V and V' ping-pong; at 15^6 each is 45.6 MB, so both stay in the 126 MB L2 and the stencil's neighbour reads are
// L2 / L1 hits.  l0 is recomputed per cell from per-dimension float64 squares instead of being stored and re-read.
// Arithmetic: float32 in the reference's expression order through the non-contracting intrinsics (__fmul_rn ...),
// so the result can be compared bit for bit with the NumPy restatement in oracle/hj_solver_oracle.py.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <new>

#include "../../include/cf_abi.h"
#include "cf_kernels.h"

namespace {

__constant__ cf_hj_problem c_p;
// ((0 + sq0[i]) + sq1[j]) + sq2[k] of the target sum in float64, pre-added on the host in the reference's order
// (Fastrack.py:296-306): the per-cell sum continues with dims 3, 4, 5
__constant__ double c_sq012[CF_HJ_MAX_PTS * CF_HJ_MAX_PTS * CF_HJ_MAX_PTS];

// global scalars of an update, in device memory
struct HjGlobals {
    float mn[6], mx[6];          // global min / max of the one-sided derivatives (init 1e9 / -1e9, graph_6D.py:22-33)
    float uL[3], uU[3], dL[3], dU[3];
    float tq[4][3];              // (u + d) / I of the corner combinations (uU,dL), (uL,dL), (uL,dU), (uU,dU)
    float hq[3][3];              // (u + d) / I of a cell's own choice: [i][0] derivative > 0, [1] < 0, [2] == 0
    float a_static[3];           // global max alpha of dims 0..2 (independent of V, u, d): set once at creation
    float ma[6];                 // global max alpha
    float t[2];                  // (t0, t1)
    float dt;
};

constexpr int kThreads = 256;
#ifndef CF_HJ_UPDATE_MIN_BLOCKS
#define CF_HJ_UPDATE_MIN_BLOCKS 6        /* resident CTAs per SM the update kernel is compiled for (40 registers, 32 B of spills: 6 % faster than 5 CTAs at 48 registers) */
#endif

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float dvd(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float sgn(float v) { return v > 0.f ? 1.f : (v < 0.f ? -1.f : 0.f); }   // my_sign
// `if d > 0: x = pos; elif d < 0: x = neg` on an hcl.scalar initialised to 0 (Fastrack.py:54-128)
__device__ __forceinline__ float pick(float d, float neg, float pos) { return d > 0.f ? pos : (d < 0.f ? neg : 0.f); }

// PTS > 0: compile-time grid size (the reference's 15: divisions become multiplies); PTS == 0: c_p.pts
template <int PTS>
__device__ __forceinline__ int npts() { return PTS > 0 ? PTS : c_p.pts; }
template <int PTS>
__device__ __forceinline__ void coords(unsigned c, int idx[6]) {
    const unsigned pts = (unsigned)npts<PTS>();
#pragma unroll
    for (int d = 5; d >= 0; --d) { idx[d] = (int)(c % pts); c /= pts; }
}

// UAV6D.dynamics (Fastrack.py:133-167) at grid cell idx: f[0..2] (no torque enters) and the torque-free part
// base[i] = (c_i x) y of f[3 + i]; f[3 + i] = base[i] + (u_i + d_i) / I_i
__device__ __forceinline__ void dynamics_parts(const int idx[6], float f[3], float base[3]) {
    const float sphi = c_p.sin_phi[idx[0]], cphi = c_p.cos_phi[idx[0]], tth = c_p.tan_theta[idx[1]], cth = c_p.cos_theta[idx[1]];
    const float p = c_p.x[3][idx[3]], q = c_p.x[4][idx[4]], r = c_p.x[5][idx[5]];
    f[0] = add(add(p, mul(mul(r, cphi), tth)), mul(mul(q, sphi), tth));
    f[1] = sub(mul(q, cphi), mul(r, sphi));
    f[2] = add(dvd(mul(r, cphi), cth), dvd(mul(q, sphi), cth));
    base[0] = mul(mul(c_p.c[0], r), q);
    base[1] = mul(mul(c_p.c[1], p), r);
    base[2] = mul(mul(c_p.c[2], p), q);
}
// The per-cell table lookups (grid values, trig, target squares) indexed by l, m, n differ from lane to lane, which the
// constant cache serialises: k_update stages them in shared memory once per CTA.
struct CellTab {
    float x3[CF_HJ_MAX_PTS], x4[CF_HJ_MAX_PTS], x5[CF_HJ_MAX_PTS];
    float sphi[CF_HJ_MAX_PTS], cphi[CF_HJ_MAX_PTS], tth[CF_HJ_MAX_PTS], cth[CF_HJ_MAX_PTS];
    double sq3[CF_HJ_MAX_PTS], sq4[CF_HJ_MAX_PTS], sq5[CF_HJ_MAX_PTS];
    float c[3];
};
__device__ __forceinline__ void fill_tab(CellTab& t) {
    const int i = threadIdx.x;
    if (i < CF_HJ_MAX_PTS) {
        t.x3[i] = c_p.x[3][i]; t.x4[i] = c_p.x[4][i]; t.x5[i] = c_p.x[5][i];
        t.sphi[i] = c_p.sin_phi[i]; t.cphi[i] = c_p.cos_phi[i]; t.tth[i] = c_p.tan_theta[i]; t.cth[i] = c_p.cos_theta[i];
        t.sq3[i] = c_p.target_sq[3][i]; t.sq4[i] = c_p.target_sq[4][i]; t.sq5[i] = c_p.target_sq[5][i];
        if (i < 3) t.c[i] = c_p.c[i];
    }
}
__device__ __forceinline__ void dynamics_parts(const CellTab& t, const int idx[6], float f[3], float base[3]) {
    const float sphi = t.sphi[idx[0]], cphi = t.cphi[idx[0]], tth = t.tth[idx[1]], cth = t.cth[idx[1]];
    const float p = t.x3[idx[3]], q = t.x4[idx[4]], r = t.x5[idx[5]];
    f[0] = add(add(p, mul(mul(r, cphi), tth)), mul(mul(q, sphi), tth));
    f[1] = sub(mul(q, cphi), mul(r, sphi));
    f[2] = add(dvd(mul(r, cphi), cth), dvd(mul(q, sphi), cth));
    base[0] = mul(mul(t.c[0], r), q);
    base[1] = mul(mul(t.c[1], p), r);
    base[2] = mul(mul(t.c[2], p), q);
}
// max over the 4 (u, d) corner combinations of |base + (u + d) / I| (graph_6D.py:232-290), same order of maxima
__device__ __forceinline__ float corner_alpha(float base, const float (*tq)[3], int i) {
    float a = fabsf(add(base, tq[0][i]));
    a = fmaxf(a, fabsf(add(base, tq[1][i])));
    a = fmaxf(a, fabsf(add(base, tq[2][i])));
    return fmaxf(a, fabsf(add(base, tq[3][i])));
}

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// block-level min (j < 6) / max (j >= 6) of 12 per-thread values -> out[12]
__device__ __forceinline__ void block_minmax12(float v[12], float* out) {
    __shared__ float s[kThreads / 32][12];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int j = 0; j < 12; ++j) {
        float r = j < 6 ? warp_min(v[j]) : warp_max(v[j]);
        if (lane == 0) s[warp][j] = r;
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        float r = s[0][threadIdx.x];
        for (int w = 1; w < kThreads / 32; ++w) r = threadIdx.x < 6 ? fminf(r, s[w][threadIdx.x]) : fmaxf(r, s[w][threadIdx.x]);
        out[threadIdx.x] = r;
    }
}

// derivatives of one cell: L[d], R[d] (graph_6D.py:89-101 with accuracy "low").  All 12 neighbour loads are issued
// before the first use (clamped addresses at the ends: the value loaded there is never selected), and the three
// cases of first_orderENO6D.py are selects over the same numerators, so there is no branch and no serialised wait.
template <int PTS>
__device__ __forceinline__ void cell_derivs(const float* __restrict__ V, unsigned c, const int idx[6], float vc, float L[6], float R[6]) {
    const int pts = npts<PTS>();
    float vm[6], vp[6];
    unsigned stride = 1;
#pragma unroll
    for (int d = 5; d >= 0; --d) {
        vp[d] = __ldg(V + c + (idx[d] < pts - 1 ? stride : 0u));
        vm[d] = __ldg(V + c - (idx[d] > 0 ? stride : 0u));
        stride *= (unsigned)pts;
    }
    const float sg = sgn(vc);
#pragma unroll
    for (int d = 0; d < 6; ++d) {
        const float dn = sub(vp[d], vc), dp = sub(vc, vm[d]);               // forward / backward difference
        const float lb = add(vc, mul(fabsf(dn), sg));                       // left_boundary  (used at index 0)
        const float rb = add(vc, mul(fabsf(dp), sg));                       // right_boundary (used at index pts-1)
        const float numL = idx[d] == 0 ? sub(vc, lb) : dp;
        const float numR = idx[d] == pts - 1 ? sub(rb, vc) : dn;
        L[d] = dvd(numL, c_p.dx[d]);
        R[d] = dvd(numR, c_p.dx[d]);
    }
}

// Global min / max of the one-sided derivatives, reduced as NUMERATORS: round-to-nearest division by dx > 0 is
// monotone, so min / max over cells of (num / dx) equals (min / max num) / dx bit for bit and the division happens once
// in k_choose.  Every interior difference is the right derivative of one cell and the left derivative of the next, so a
// cell contributes its forward difference plus, at the two ends, the extrapolated ghost term (first_orderENO6D.py).
// Grid-stride: a thread accumulates many cells in registers before the one block reduction.
template <int PTS>
__global__ void __launch_bounds__(kThreads) k_minmax(const float* __restrict__ V, unsigned n, float* __restrict__ partial) {
    const int pts = npts<PTS>();
    float v[12];
#pragma unroll
    for (int j = 0; j < 6; ++j) { v[j] = 3.0e38f; v[6 + j] = -3.0e38f; }
    for (unsigned c = blockIdx.x * kThreads + threadIdx.x; c < n; c += gridDim.x * kThreads) {
        int idx[6];
        coords<PTS>(c, idx);
        const float vc = __ldg(V + c);
        unsigned stride = 1;
#pragma unroll
        for (int d = 5; d >= 0; --d) {
            const int i = idx[d];
            if (i < pts - 1) {
                const float diff = sub(__ldg(V + c + stride), vc);                    // right of this cell == left of the next
                v[d] = fminf(v[d], diff); v[6 + d] = fmaxf(v[6 + d], diff);
                if (i == 0) {
                    const float num = sub(vc, add(vc, mul(fabsf(diff), sgn(vc))));   // V - left_boundary
                    v[d] = fminf(v[d], num); v[6 + d] = fmaxf(v[6 + d], num);
                }
            } else {
                const float dl = sub(vc, __ldg(V + c - stride));
                const float num = sub(add(vc, mul(fabsf(dl), sgn(vc))), vc);          // right_boundary - V
                v[d] = fminf(v[d], num); v[6 + d] = fmaxf(v[6 + d], num);
            }
            stride *= (unsigned)pts;
        }
    }
    block_minmax12(v, partial + 12 * (size_t)blockIdx.x);
}

// once per solver: global max of alpha_0..2 = |f_0..2| over all cells (they depend on neither V nor (u, d))
template <int PTS>
__global__ void __launch_bounds__(kThreads) k_static_alpha(unsigned n, float* __restrict__ partial) {
    const unsigned c = blockIdx.x * kThreads + threadIdx.x;
    float v[12];
#pragma unroll
    for (int j = 0; j < 6; ++j) { v[j] = 1e9f; v[6 + j] = -1e9f; }
    if (c < n) {
        int idx[6];
        coords<PTS>(c, idx);
        float f[3], base[3];
        dynamics_parts(idx, f, base);
#pragma unroll
        for (int d = 0; d < 3; ++d) v[6 + d] = fabsf(f[d]);
    }
    block_minmax12(v, partial + 12 * (size_t)blockIdx.x);
}
__global__ void __launch_bounds__(kThreads) k_static_alpha_reduce(const float* __restrict__ partial, int nblk, HjGlobals* __restrict__ g) {
    float v[12];
#pragma unroll
    for (int j = 0; j < 6; ++j) { v[j] = 1e9f; v[6 + j] = -1e9f; }
    for (int b = threadIdx.x; b < nblk; b += kThreads)
#pragma unroll
        for (int j = 0; j < 3; ++j) v[6 + j] = fmaxf(v[6 + j], partial[12 * (size_t)b + 6 + j]);
    __shared__ float out[12];
    block_minmax12(v, out);
    __syncthreads();
    if (threadIdx.x < 3) g->a_static[threadIdx.x] = out[6 + threadIdx.x];
}

// One CTA: global min / max over the block partials, opt_ctrl / opt_dstb at the extreme derivatives
// (graph_6D.py:203-230), the torque terms, the global max alpha, the step bound and the time advance (:38-48, :301-318)
__global__ void __launch_bounds__(kThreads) k_choose(const float* __restrict__ partial, int nblk, HjGlobals* __restrict__ g) {
    float v[12];
#pragma unroll
    for (int j = 0; j < 6; ++j) { v[j] = 3.0e38f; v[6 + j] = -3.0e38f; }
    for (int b = threadIdx.x; b < nblk; b += kThreads) {
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            v[j] = fminf(v[j], partial[12 * (size_t)b + j]);
            v[6 + j] = fmaxf(v[6 + j], partial[12 * (size_t)b + 6 + j]);
        }
    }
    __shared__ float out[12];
    __shared__ float tq[4][3];
    block_minmax12(v, out);
    __syncthreads();
    if (threadIdx.x == 0) {
        float uL[3], uU[3], dL[3], dU[3];
        // numerators -> derivatives, then the reference's running min / max that start at 1e9 / -1e9 (graph_6D.py:22-33)
#pragma unroll
        for (int d = 0; d < 6; ++d) {
            out[d] = fminf(1e9f, dvd(out[d], c_p.dx[d]));
            out[6 + d] = fmaxf(-1e9f, dvd(out[6 + d], c_p.dx[d]));
            g->mn[d] = out[d]; g->mx[d] = out[6 + d];
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            uL[i] = pick(out[3 + i], c_p.u_max[i], c_p.u_min[i]);             // uMode "min"
            uU[i] = pick(out[9 + i], c_p.u_max[i], c_p.u_min[i]);
            dL[i] = pick(out[3 + i], c_p.d_max[i], c_p.d_min[i]);             // dMode "max"
            dU[i] = pick(out[9 + i], c_p.d_max[i], c_p.d_min[i]);
            g->uL[i] = uL[i]; g->uU[i] = uU[i]; g->dL[i] = dL[i]; g->dU[i] = dU[i];
            const float I = c_p.inertia[i];
            tq[0][i] = dvd(add(uU[i], dL[i]), I); tq[1][i] = dvd(add(uL[i], dL[i]), I);
            tq[2][i] = dvd(add(uL[i], dU[i]), I); tq[3][i] = dvd(add(uU[i], dU[i]), I);
#pragma unroll
            for (int k = 0; k < 4; ++k) g->tq[k][i] = tq[k][i];
            g->hq[i][0] = dvd(add(c_p.u_min[i], c_p.d_min[i]), I);            // a cell's own choice, derivative > 0
            g->hq[i][1] = dvd(add(c_p.u_max[i], c_p.d_max[i]), I);            // < 0
            g->hq[i][2] = dvd(add(0.0f, 0.0f), I);                            // == 0: both hcl.scalars stay 0
        }
    }
    __syncthreads();
    // alpha_3..5 depend on two grid coordinates each: (r, q), (p, r), (p, q)
    const int pts = c_p.pts;
#pragma unroll
    for (int j = 0; j < 6; ++j) { v[j] = 1e9f; v[6 + j] = -1e9f; }
    for (int e = threadIdx.x; e < pts * pts; e += kThreads) {
        const int ia = e / pts, ib = e - ia * pts;
        const float b0 = mul(mul(c_p.c[0], c_p.x[5][ib]), c_p.x[4][ia]);
        const float b1 = mul(mul(c_p.c[1], c_p.x[3][ia]), c_p.x[5][ib]);
        const float b2 = mul(mul(c_p.c[2], c_p.x[3][ia]), c_p.x[4][ib]);
        v[9] = fmaxf(v[9], corner_alpha(b0, tq, 0));
        v[10] = fmaxf(v[10], corner_alpha(b1, tq, 1));
        v[11] = fmaxf(v[11], corner_alpha(b2, tq, 2));
    }
    __shared__ float out2[12];
    block_minmax12(v, out2);
    __syncthreads();
    if (threadIdx.x == 0) {
        float ma[6] = {g->a_static[0], g->a_static[1], g->a_static[2], out2[9], out2[10], out2[11]};
        float sbi = dvd(ma[0], c_p.dx[0]);
#pragma unroll
        for (int d = 1; d < 6; ++d) sbi = add(sbi, dvd(ma[d], c_p.dx[d]));
#pragma unroll
        for (int d = 0; d < 6; ++d) g->ma[d] = ma[d];
        float sb = dvd(0.8f, sbi);
        const float room = sub(g->t[1], g->t[0]);
        if (sb > room) sb = room;
        g->t[0] = add(g->t[0], sb);
        g->dt = sb;
    }
}

template <int PTS>
__device__ __forceinline__ float target_l0(const int idx[6]) {
    // l0 = FasTrackTarget: float64 sum of the per-dimension squares, rounded once (Fastrack.py:296-306)
    double t = c_sq012[(idx[0] * CF_HJ_MAX_PTS + idx[1]) * CF_HJ_MAX_PTS + idx[2]];
#pragma unroll
    for (int d = 3; d < 6; ++d) t = __dadd_rn(t, c_p.target_sq[d][idx[d]]);
    return (float)t;
}

// Hamiltonian + dissipation + time step + maxVWithV0 of one cell (graph_6D.py:64-330)
template <int PTS>
__global__ void __launch_bounds__(kThreads, CF_HJ_UPDATE_MIN_BLOCKS) k_update(const float* __restrict__ V, float* __restrict__ Vout, unsigned n,
                                                     const HjGlobals* __restrict__ g) {
    __shared__ float tq[4][3], hq[3][3];
    __shared__ float dt_s;
    __shared__ CellTab tab;
    fill_tab(tab);
    if (threadIdx.x < 12) tq[threadIdx.x / 3][threadIdx.x % 3] = g->tq[threadIdx.x / 3][threadIdx.x % 3];
    else if (threadIdx.x < 21) hq[(threadIdx.x - 12) / 3][(threadIdx.x - 12) % 3] = g->hq[(threadIdx.x - 12) / 3][(threadIdx.x - 12) % 3];
    else if (threadIdx.x == 21) dt_s = g->dt;
    __syncthreads();
    const unsigned c = blockIdx.x * kThreads + threadIdx.x;
    if (c >= n) return;
    int idx[6];
    coords<PTS>(c, idx);
    const float vc = __ldg(V + c);
    float L[6], R[6], dV[6];
    cell_derivs<PTS>(V, c, idx, vc, L, R);
#pragma unroll
    for (int d = 0; d < 6; ++d) dV[d] = mul(add(L[d], R[d]), 0.5f);           // (L + R) / 2: exact either way
    float f[6], base[3], a[6];
    dynamics_parts(tab, idx, f, base);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float own = dV[3 + i] > 0.f ? hq[i][0] : (dV[3 + i] < 0.f ? hq[i][1] : hq[i][2]);
        f[3 + i] = add(base[i], own);
        a[i] = fabsf(f[i]);
        a[3 + i] = corner_alpha(base[i], tq, i);
    }
    float h = mul(f[0], dV[0]);
#pragma unroll
    for (int d = 1; d < 6; ++d) h = add(h, mul(f[d], dV[d]));
    float vn = -h;
    float ds = mul(sub(R[0], L[0]), a[0]);
#pragma unroll
    for (int d = 1; d < 6; ++d) ds = add(ds, mul(sub(R[d], L[d]), a[d]));
    ds = mul(0.5f, ds);
    vn = -sub(vn, ds);
    float out = add(vc, mul(vn, dt_s));
    // l0 = FasTrackTarget, float64 sum in the reference's order, rounded once (Fastrack.py:296-306)
    double t = c_sq012[(idx[0] * CF_HJ_MAX_PTS + idx[1]) * CF_HJ_MAX_PTS + idx[2]];
    t = __dadd_rn(t, tab.sq3[idx[3]]);
    t = __dadd_rn(t, tab.sq4[idx[4]]);
    t = __dadd_rn(t, tab.sq5[idx[5]]);
    const float l0 = (float)t;
    if (out < l0) out = l0;
    Vout[c] = out;
}

template <int PTS>
__global__ void __launch_bounds__(kThreads) k_init_l0(float* __restrict__ V, unsigned n) {
    const unsigned c = blockIdx.x * kThreads + threadIdx.x;
    if (c >= n) return;
    int idx[6];
    coords<PTS>(c, idx);
    V[c] = target_l0<PTS>(idx);
}

}  // namespace

// ---------------------------------------------------------------------------------- host side / C ABI
struct cf_hj_solver {
    cf_hj_problem prob;
    int device = 0;
    long n = 0;
    int nblk = 0;                        // blocks of the one-thread-per-cell kernels
    int nred = 0;                        // blocks of the grid-stride reduction kernel = rows of `partial` it writes
    float* V[2] = {nullptr, nullptr};
    int cur = 0;
    float* partial = nullptr;            // [nblk][12]
    HjGlobals* g = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    double ms_sum = 0.0;
    int64_t updates = 0;
};

namespace {

const cf_hj_solver* g_const_owner = nullptr;   // whose problem sits in c_p (one device constant per module)

int hj_fail(const char* where, const char* what) { cf::set_error(where, what); return CF_ERR_INVALID; }
int hj_cuda(cudaError_t e, const char* where) { cf::set_error(where, cudaGetErrorString(e)); return CF_ERR_CUDA; }
#define HJ_CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return hj_cuda(e_, #call); } while (0)

int bind(cf_hj_solver* s) {
    HJ_CU(cudaSetDevice(s->device));
    if (g_const_owner != s) {
        HJ_CU(cudaMemcpyToSymbol(c_p, &s->prob, sizeof(cf_hj_problem)));
        static double sq012[CF_HJ_MAX_PTS * CF_HJ_MAX_PTS * CF_HJ_MAX_PTS];
        for (int i = 0; i < CF_HJ_MAX_PTS; ++i)
            for (int j = 0; j < CF_HJ_MAX_PTS; ++j)
                for (int k = 0; k < CF_HJ_MAX_PTS; ++k) {
                    volatile double t = 0.0;                 // one rounding per addition, as NumPy does it
                    t = t + s->prob.target_sq[0][i];
                    t = t + s->prob.target_sq[1][j];
                    t = t + s->prob.target_sq[2][k];
                    sq012[(i * CF_HJ_MAX_PTS + j) * CF_HJ_MAX_PTS + k] = t;
                }
        HJ_CU(cudaMemcpyToSymbol(c_sq012, sq012, sizeof(sq012)));
        g_const_owner = s;
    }
    return CF_OK;
}

// the three launches of one update, all on the default stream; t = (t0, t1) must already be in s->g
int launch_update(cf_hj_solver* s) {
    const float* V = s->V[s->cur];
    float* Vout = s->V[s->cur ^ 1];
    const unsigned n = (unsigned)s->n;
    if (s->prob.pts == 15) k_minmax<15><<<s->nred, kThreads>>>(V, n, s->partial);
    else k_minmax<0><<<s->nred, kThreads>>>(V, n, s->partial);
    k_choose<<<1, kThreads>>>(s->partial, s->nred, s->g);
    if (s->prob.pts == 15) k_update<15><<<s->nblk, kThreads>>>(V, Vout, n, s->g);
    else k_update<0><<<s->nblk, kThreads>>>(V, Vout, n, s->g);
    HJ_CU(cudaGetLastError());
    s->cur ^= 1;
    return CF_OK;
}

int timed_update(cf_hj_solver* s, float t0, float t1, float* t0_new) {
    const float t[2] = {t0, t1};
    HJ_CU(cudaMemcpy(reinterpret_cast<char*>(s->g) + offsetof(HjGlobals, t), t, sizeof(t), cudaMemcpyHostToDevice));
    HJ_CU(cudaEventRecord(s->ev0));
    int rc = launch_update(s);
    if (rc) return rc;
    HJ_CU(cudaEventRecord(s->ev1));
    HJ_CU(cudaMemcpy(t0_new, reinterpret_cast<char*>(s->g) + offsetof(HjGlobals, t), sizeof(float), cudaMemcpyDeviceToHost));
    float ms = 0.f;
    HJ_CU(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
    s->ms_sum += ms;
    s->updates += 1;
    return CF_OK;
}

}  // namespace

extern "C" {

int cf_hj_solver_create(const cf_hj_problem* prob, int32_t device, cf_hj_solver** out) {
    if (!prob || !out) return hj_fail("cf_hj_solver_create", "null argument");
    if (prob->pts < 3 || prob->pts > CF_HJ_MAX_PTS) return hj_fail("cf_hj_solver_create", "pts must be in 3..16");
    for (int d = 0; d < 6; ++d)
        if (!(prob->dx[d] > 0.f)) return hj_fail("cf_hj_solver_create", "dx must be positive");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1)
        return hj_cuda(cudaErrorNoDevice, "cf_hj_solver_create: no CUDA device (there is no CPU fallback)");
    if (device < 0 || device >= ndev) return hj_fail("cf_hj_solver_create", "bad device index");
    cf_hj_solver* s = new (std::nothrow) cf_hj_solver();
    if (!s) return hj_fail("cf_hj_solver_create", "out of host memory");
    s->prob = *prob;
    s->device = device;
    s->n = 1;
    for (int d = 0; d < 6; ++d) s->n *= prob->pts;
    s->nblk = (int)((s->n + kThreads - 1) / kThreads);
    cudaDeviceProp dp;
    const int sms = cudaGetDeviceProperties(&dp, device) == cudaSuccess ? dp.multiProcessorCount : 148;
    s->nred = s->nblk < sms * 8 ? s->nblk : sms * 8;                 // 8 resident CTAs of 256 threads per SM
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaMalloc((void**)&s->V[0], sizeof(float) * s->n);
    if (e == cudaSuccess) e = cudaMalloc((void**)&s->V[1], sizeof(float) * s->n);
    if (e == cudaSuccess) e = cudaMalloc((void**)&s->partial, sizeof(float) * 12 * (size_t)s->nblk);
    if (e == cudaSuccess) e = cudaMalloc((void**)&s->g, sizeof(HjGlobals));
    if (e == cudaSuccess) e = cudaMemset(s->g, 0, sizeof(HjGlobals));
    if (e == cudaSuccess) e = cudaEventCreate(&s->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&s->ev1);
    if (e != cudaSuccess) { cf_hj_solver_destroy(s); return hj_cuda(e, "cf_hj_solver_create"); }
    int rc = bind(s);
    if (rc) { cf_hj_solver_destroy(s); return rc; }
    if (prob->pts == 15) {
        k_init_l0<15><<<s->nblk, kThreads>>>(s->V[0], (unsigned)s->n);   // V_0 = init_value = target (solver.py:97-108)
        k_static_alpha<15><<<s->nblk, kThreads>>>((unsigned)s->n, s->partial);
    } else {
        k_init_l0<0><<<s->nblk, kThreads>>>(s->V[0], (unsigned)s->n);
        k_static_alpha<0><<<s->nblk, kThreads>>>((unsigned)s->n, s->partial);
    }
    k_static_alpha_reduce<<<1, kThreads>>>(s->partial, s->nblk, s->g);
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { cf_hj_solver_destroy(s); return hj_cuda(e, "k_init_l0"); }
    *out = s;
    return CF_OK;
}

int cf_hj_solver_destroy(cf_hj_solver* s) {
    if (!s) return CF_OK;
    cudaSetDevice(s->device);
    cudaFree(s->V[0]); cudaFree(s->V[1]); cudaFree(s->partial); cudaFree(s->g);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (g_const_owner == s) g_const_owner = nullptr;
    delete s;
    return CF_OK;
}

int cf_hj_solver_set_value(cf_hj_solver* s, const float* V_host) {
    if (!s || !V_host) return hj_fail("cf_hj_solver_set_value", "null argument");
    HJ_CU(cudaSetDevice(s->device));
    HJ_CU(cudaMemcpy(s->V[s->cur], V_host, sizeof(float) * s->n, cudaMemcpyHostToDevice));
    return CF_OK;
}

int cf_hj_solver_get_value(cf_hj_solver* s, float* V_host) {
    if (!s || !V_host) return hj_fail("cf_hj_solver_get_value", "null argument");
    HJ_CU(cudaSetDevice(s->device));
    HJ_CU(cudaMemcpy(V_host, s->V[s->cur], sizeof(float) * s->n, cudaMemcpyDeviceToHost));
    return CF_OK;
}

int cf_hj_solver_update(cf_hj_solver* s, float t0, float t1, float* t0_new, float info[19]) {
    if (!s || !t0_new) return hj_fail("cf_hj_solver_update", "null argument");
    int rc = bind(s);
    if (rc) return rc;
    rc = timed_update(s, t0, t1, t0_new);
    if (rc) return rc;
    if (info) {
        HjGlobals h;
        HJ_CU(cudaMemcpy(&h, s->g, sizeof(h), cudaMemcpyDeviceToHost));
        memcpy(info, h.mn, sizeof(float) * 6);
        memcpy(info + 6, h.mx, sizeof(float) * 6);
        memcpy(info + 12, h.ma, sizeof(float) * 6);
        info[18] = h.dt;
    }
    return CF_OK;
}

// for i in 1..len(tau)-1: t = (tNow, tau[i]); while tNow <= tau[i] - 1e-4: update; tNow = t[0]   (solver.py:152-206)
int cf_hj_solver_run(cf_hj_solver* s, double lookback, double t_step, int64_t max_updates, int64_t* updates_done,
                     double* t_final) {
    if (!s || !(t_step > 0.0) || !(lookback > 0.0)) return hj_fail("cf_hj_solver_run", "bad argument");
    int rc = bind(s);
    if (rc) return rc;
    const int64_t len = (int64_t)ceil((lookback + 1e-5) / t_step);   // len(np.arange(0, lookback + 1e-5, t_step))
    double t_now = 0.0;
    int64_t done = 0;
    bool capped = false;
    for (int64_t i = 1; i < len && !capped; ++i) {
        const double tau_i = 0.0 + (double)i * t_step;
        float t0 = (float)t_now;
        const float t1 = (float)tau_i;
        while (t_now <= tau_i - 1e-4) {
            float t0_new = t0;
            rc = timed_update(s, t0, t1, &t0_new);
            if (rc) return rc;
            t0 = t0_new;
            t_now = (double)t0;
            done += 1;
            if (max_updates > 0 && done >= max_updates) { capped = true; break; }
        }
    }
    if (updates_done) *updates_done = done;
    if (t_final) *t_final = t_now;
    return CF_OK;
}

int cf_hj_solver_timing(cf_hj_solver* s, double* avg_ms_per_update, int64_t* updates) {
    if (!s) return hj_fail("cf_hj_solver_timing", "null argument");
    if (avg_ms_per_update) *avg_ms_per_update = s->updates ? s->ms_sum / (double)s->updates : 0.0;
    if (updates) *updates = s->updates;
    return CF_OK;
}

}  // extern "C"
