// cf_device.cuh - per-env device code of the batched CrazyFlie hover environment (sm_100a).
//
// One thread owns one env.  The env's 76-word state record lives word-major inside a tile of
// CF_TILE envs ([tile][word][env] floats), so the same accessor (EnvRef: base pointer + stride
// CF_TILE) works on the tile staged in shared memory by the step kernel and on the tile in
// global memory (reset / init kernels); consecutive threads touch consecutive floats either way.
// The control step is written in phases that pull only the words they need into registers: HJ
// disturbance lookup, 2 physics substeps (PT1 motor lag, thrust/torque, drag, disturbance torque,
// rigid-body integration), sensor-noise pipeline with gyro low-pass, observation history, reward,
// done/cost, TimeLimit, and the in-place auto-reset with domain randomisation.  Random draws are
// produced block by block (Philox4x32-10) right before they are consumed.  fp32 arithmetic (the
// reference is float64 NumPy + PyBullet); each routine cites the reference function it follows
// (paths relative to phoenix_drone_simulation/ of the reference checkout).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "cf_types.h"

namespace cf {

constexpr float kPi = 3.14159265358979323846f;

__constant__ double c_grid[6][CF_HJ_PTS];   // odp Grid.grid_points, float64 linspace values
__constant__ double c_dx[6];
__constant__ double c_boltz_cdf[21];

// ---------------------------------------------------------------------------------- fast math
// MUFU-based approximations (<= 2 ulp) for quantities whose error budget is 1e-5 relative.
__device__ __forceinline__ float fsqrt(float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float frcp(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(hi, fmaxf(lo, v)); }

// ---------------------------------------------------------------------------------- RNG
// Philox4x32-10, counter = (gid_lo, event counter, block, tag ^ gid_hi<<8), key = seed.  The ten
// round keys are derived on the host (DevCfg::pk) so they are constant-bank operands here.
#ifndef CF_PHILOX_ROUNDS
#define CF_PHILOX_ROUNDS 10             /* tools/ only: instruction-count sensitivity probes build with fewer rounds */
#endif
#define CF_PRAGMA_(x) _Pragma(#x)
#define CF_UNROLL_N_(n) CF_PRAGMA_(unroll n)
#ifndef CF_PHILOX_UNROLL
#define CF_PHILOX_UNROLL CF_PHILOX_ROUNDS   /* tools/ only: smaller values trade straight-line rounds for code size */
#endif
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, const DevCfg& cfg) {
    CF_UNROLL_N_(CF_PHILOX_UNROLL)
    for (int r = 0; r < CF_PHILOX_ROUNDS; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ cfg.pk[r][0], lo1, hi0 ^ c.w ^ cfg.pk[r][1], lo0);
    }
    return c;
}
// 23 random bits -> (2k+1) / 2^24 in (0,1): mantissa trick, no int->float conversion (XU pipe)
__device__ __forceinline__ float u01(uint32_t w) {
    return __uint_as_float(0x3f800000u | (w >> 9)) - 0.99999994f;   // 1 - 2^-24, exact
}
// Box-Muller on the two word pairs of a Philox block: (r cos, r sin) with r = sqrt(-2 ln u0), angle pi (2 u1 - 1);
// MUFU lg2 / sqrt / sin / cos.  tests/test_rng_bits.py sweeps every one of the 2^23 u01 values through this function
// (cf_debug_box_muller) and states the error bound of the normals against float64.
__device__ __forceinline__ float4 box_muller(uint4 w) {
    float r0 = fsqrt(-1.3862943611f * __log2f(u01(w.x))), r1 = fsqrt(-1.3862943611f * __log2f(u01(w.z)));   // sqrt(-2 ln u)
    float s0, c0_, s1, c1_;
    __sincosf(kPi * (2.0f * u01(w.y) - 1.0f), &s0, &c0_);
    __sincosf(kPi * (2.0f * u01(w.w) - 1.0f), &s1, &c1_);
    return make_float4(r0 * c0_, r0 * s0, r1 * c1_, r1 * s1);
}
struct Rng {
    uint32_t c0, c1, c3;
    __device__ __forceinline__ uint4 raw(const DevCfg& cfg, uint32_t block) const {
        return philox4x32_10(make_uint4(c0, c1, block, c3), cfg);
    }
    // four standard normals of block b (Box-Muller on two pairs, same pairing as the oracle)
    __device__ __forceinline__ float4 normals(const DevCfg& cfg, uint32_t b) const { return box_muller(raw(cfg, b)); }
    __device__ __forceinline__ float4 uniforms(const DevCfg& cfg, uint32_t b) const {
        uint4 w = raw(cfg, 16u + b);
        return make_float4(u01(w.x), u01(w.y), u01(w.z), u01(w.w));
    }
};
__device__ __forceinline__ Rng make_rng(long long gid, uint32_t ctr, uint32_t tag) {
    Rng r;
    r.c0 = (uint32_t)gid; r.c1 = ctr;
    r.c3 = tag ^ ((uint32_t)((unsigned long long)gid >> 32) << 8);
    return r;
}

// Standard draws of one event, fetched four at a time where they are consumed: generated
// (Philox) or read from the injected slot-major buffers [slot][n] (parity mode).
struct DrawSrc {
    Rng rng;
    const float* iz;      // injected normals  [nz][n] or null
    const float* iu;      // injected uniforms [nu][n]
    int n, i, nu;
    __device__ __forceinline__ float4 z(const DevCfg& cfg, int b) const {
        if (iz) {
            const float* p = iz + (size_t)(4 * b) * n + i;
            return make_float4(p[0], p[n], p[2 * (size_t)n], p[3 * (size_t)n]);
        }
        return rng.normals(cfg, b);
    }
    __device__ __forceinline__ float4 u(const DevCfg& cfg, int b) const {
        if (iu) {
            const float* p = iu + (size_t)(4 * b) * n + i;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (4 * b < nu) v.x = p[0];
            if (4 * b + 1 < nu) v.y = p[n];
            if (4 * b + 2 < nu) v.z = p[2 * (size_t)n];
            if (4 * b + 3 < nu) v.w = p[3 * (size_t)n];
            return v;
        }
        return rng.uniforms(cfg, b);
    }
};
__device__ __forceinline__ DrawSrc make_draws(const DevCfg& cfg, long long gid, uint32_t ctr, uint32_t tag,
                                              const float* iz, const float* iu, int n, int i, int nu) {
    DrawSrc d;
    d.rng = make_rng(gid, ctr, tag);
    d.iz = cfg.injected ? iz : nullptr;
    d.iu = cfg.injected ? iu : nullptr;
    d.n = n; d.i = i; d.nu = nu;
    return d;
}

// ---------------------------------------------------------------------------------- state access
// Word w (CF_S_* of cf_abi.h) of this thread's env.  `p` addresses the words the step kernel stages in shared memory
// (device words 0..CF_BULK_WORDS-1) - or the whole tile in HBM for the reset / init kernels -, `h` the tile in HBM for the
// history words (device words CF_BULK_WORDS..CF_DYN_WORDS-1), which never go through shared memory.  Both at stride CF_TILE.
// w is a compile-time constant at every call site (after unrolling), so the choice of pointer and the word permutation fold.
struct EnvRef {
    float* p;
    float* h;
    __device__ __forceinline__ float* at(int w) const {
        const int d = cf_dev_word(w);
        return ((d >= CF_BULK_WORDS && d < CF_DYN_WORDS) ? h : p) + d * CF_TILE;
    }
    __device__ __forceinline__ float f(int w) const { return *at(w); }
    __device__ __forceinline__ void f(int w, float v) const { *at(w) = v; }
    __device__ __forceinline__ uint32_t u(int w) const { return __float_as_uint(*at(w)); }
    __device__ __forceinline__ void u(int w, uint32_t v) const { *at(w) = __uint_as_float(v); }
    __device__ __forceinline__ void get3(int w, float v[3]) const { v[0] = f(w); v[1] = f(w + 1); v[2] = f(w + 2); }
    __device__ __forceinline__ void get4(int w, float v[4]) const { v[0] = f(w); v[1] = f(w + 1); v[2] = f(w + 2); v[3] = f(w + 3); }
    __device__ __forceinline__ void set3(int w, const float v[3]) const { f(w, v[0]); f(w + 1, v[1]); f(w + 2, v[2]); }
    __device__ __forceinline__ void set4(int w, const float v[4]) const { f(w, v[0]); f(w + 1, v[1]); f(w + 2, v[2]); f(w + 3, v[3]); }
    // four consecutive words below CF_S_ACT_PREV1 at a RUN-TIME word index (device order == record order there)
    __device__ __forceinline__ void get4_low(int w, float v[4]) const {
        v[0] = p[w * CF_TILE]; v[1] = p[(w + 1) * CF_TILE]; v[2] = p[(w + 2) * CF_TILE]; v[3] = p[(w + 3) * CF_TILE];
    }
};
// global-memory address of env i's word 0 in the tiled state array
__device__ __forceinline__ float* env_global(float* S, int i) {
    return S + (size_t)(i / CF_TILE) * (CF_STATE_WORDS * CF_TILE) + (i % CF_TILE);
}
__device__ __forceinline__ EnvRef env_ref_global(float* S, int i) { float* g = env_global(S, i); return EnvRef{g, g}; }

// domain-randomised parameters (words CF_S_DT..): read-only in the step, rewritten by reset
struct DrParams {
    float dt, mass, J[3], ftf1, B[4], K[4];
    __device__ __forceinline__ void load(const EnvRef g) {
        dt = g.f(CF_S_DT); mass = g.f(CF_S_MASS); g.get3(CF_S_INERTIA, J); ftf1 = g.f(CF_S_FTF1);
        g.get4(CF_S_MOTOR_B, B); g.get4(CF_S_MOTOR_K, K);
    }
    __device__ __forceinline__ void store(const EnvRef g) const {
        g.f(CF_S_DT, dt); g.f(CF_S_MASS, mass); g.set3(CF_S_INERTIA, J); g.f(CF_S_FTF1, ftf1);
        g.set4(CF_S_MOTOR_B, B); g.set4(CF_S_MOTOR_K, K);
    }
};

// misc word: level_idx | ring_idx<<8 | iteration<<12 | fresh<<31   (cf_abi.h CF_S_MISC)
__device__ __forceinline__ int misc_level(uint32_t m) { return (int)(m & 0xffu); }
__device__ __forceinline__ int misc_ring(uint32_t m) { return (int)((m >> 8) & 0xfu); }
__device__ __forceinline__ bool misc_fresh(uint32_t m) { return (m >> 31) != 0u; }
__device__ __forceinline__ uint32_t misc_pack(int level, int ring, int iter, bool fresh) {
    return (uint32_t)level | ((uint32_t)ring << 8) | ((uint32_t)(iter & 1) << 12) | (fresh ? 0x80000000u : 0u);
}

// 64-bit stores of an 8-byte aligned row (obs rows are 136 B or 168 B)
template <int DIM>
__device__ __forceinline__ void store_row(float* __restrict__ dst, const float* v) {
    float2* p = reinterpret_cast<float2*>(dst);
#pragma unroll
    for (int k = 0; k < DIM / 2; ++k) p[k] = make_float2(v[2 * k], v[2 * k + 1]);
}

// ---------------------------------------------------------------------------------- kinematics
// btMatrix3x3::setRotation == pybullet.getMatrixFromQuaternion (agents.py:452, physics.py:234)
__device__ __forceinline__ void quat_to_R(const float q[4], float R[9]) {
    float x = q[0], y = q[1], z = q[2], w = q[3];
    float d = x * x + y * y + z * z + w * w, s = 2.0f * frcp(d);
    float xs = x * s, ys = y * s, zs = z * s;
    float wx = w * xs, wy = w * ys, wz = w * zs, xx = x * xs, xy = x * ys, xz = x * zs;
    float yy = y * ys, yz = y * zs, zz = z * zs;
    R[0] = 1.0f - (yy + zz); R[1] = xy - wz; R[2] = xz + wy;
    R[3] = xy + wz; R[4] = 1.0f - (xx + zz); R[5] = yz - wx;
    R[6] = xz - wy; R[7] = yz + wx; R[8] = 1.0f - (xx + yy);
}
__device__ __forceinline__ void matvec(const float R[9], const float v[3], float o[3]) {
    o[0] = R[0] * v[0] + R[1] * v[1] + R[2] * v[2];
    o[1] = R[3] * v[0] + R[4] * v[1] + R[5] * v[2];
    o[2] = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
}
__device__ __forceinline__ void matTvec(const float R[9], const float v[3], float o[3]) {
    o[0] = R[0] * v[0] + R[3] * v[1] + R[6] * v[2];
    o[1] = R[1] * v[0] + R[4] * v[1] + R[7] * v[2];
    o[2] = R[2] * v[0] + R[5] * v[1] + R[8] * v[2];
}
// atan2 / asin without libdevice's IEEE division and special-case paths (the two Euler conversions of a step were 306
// warp-instructions of atan2f / asinf).  atan(a) / a on a in [0, 1] as a degree-8 polynomial in a^2 (weighted minimax fit,
// coefficients from a Lawson iteration in float64): max |error| of the assembled atan2 over the circle 1.6e-7 rad = at most
// 2 ulp of the result for |angle| >= pi/8, absolute 1.6e-7 below (tests/test_gpu_parity.py::test_fast_atan2_asin).
// atan2(0, 0) = 0 like libm; NaN in, NaN out.
__device__ __forceinline__ float atan2_fast(float y, float x) {
    const float ax = fabsf(x), ay = fabsf(y);
    const float mx = fmaxf(fmaxf(ax, ay), 1e-30f), mn = fminf(ax, ay);
    const float a = mn * frcp(mx), t = a * a;
    float q = 2.479950796e-03f;
    q = fmaf(q, t, -1.449995057e-02f); q = fmaf(q, t, 3.995352576e-02f); q = fmaf(q, t, -7.250832014e-02f);
    q = fmaf(q, t, 1.050737913e-01f); q = fmaf(q, t, -1.416375388e-01f); q = fmaf(q, t, 1.998630745e-01f);
    q = fmaf(q, t, -3.333262557e-01f); q = fmaf(q, t, 9.999998923e-01f);
    float r = a * q;
    r = ay > ax ? 0.5f * kPi - r : r;
    r = x < 0.0f ? kPi - r : r;
    return copysignf(r, y);
}
// asin(x) = atan2(x, sqrt((1 - x)(1 + x))) for |x| <= 1 (1 - x is exact for x >= 0.5: no cancellation near the ends)
__device__ __forceinline__ float asin_fast(float x) { return atan2_fast(x, fsqrt((1.0f - x) * (1.0f + x))); }

// pybullet.getEulerFromQuaternion (agents.py:446) incl. the gimbal branch at |sin pitch| >= 0.99999
__device__ __forceinline__ void quat_to_rpy_bullet(const float q[4], float rpy[3]) {
    float x = q[0], y = q[1], z = q[2], w = q[3];
    float sqx = x * x, sqy = y * y, sqz = z * z, squ = w * w;
    float sarg = -2.0f * (x * z - w * y);
    // one atan2 body for both cases: in the gimbal branches roll = 0 and yaw = 2 atan2(+-x, -+y)
    const bool lo = sarg <= -0.99999f, hi = sarg >= 0.99999f, gimbal = lo || hi;
    const float ry = gimbal ? (lo ? x : -x) : 2.f * (x * y + w * z);
    const float rx = gimbal ? (lo ? -y : y) : squ + sqx - sqy - sqz;
    const float yaw = atan2_fast(ry, rx);
    rpy[2] = gimbal ? 2.f * yaw : yaw;
    rpy[1] = gimbal ? (lo ? -0.5f * kPi : 0.5f * kPi) : asin_fast(fminf(1.0f, fmaxf(-1.0f, sarg)));
    rpy[0] = gimbal ? 0.f : atan2_fast(2.f * (y * z + w * x), squ - sqx - sqy + sqz);
}
// sin and cos for |x| <= pi/2 (half angles of clipped Euler angles): Taylor polynomials to x^11 / x^12,
// truncation < 6e-8 relative at the interval end; no range reduction, no slow path.
__device__ __forceinline__ void sincos_half(float x, float* sn, float* cs) {
    float x2 = x * x;
    float ps = fmaf(x2, fmaf(x2, fmaf(x2, fmaf(x2, fmaf(x2, -2.5052108e-8f, 2.7557319e-6f), -1.9841270e-4f), 8.3333333e-3f),
                             -1.6666667e-1f), 1.0f);
    float pc = fmaf(x2, fmaf(x2, fmaf(x2, fmaf(x2, fmaf(x2, fmaf(x2, 2.0876757e-9f, -2.7557319e-7f), 2.4801587e-5f), -1.3888889e-3f),
                                      4.1666667e-2f), -0.5f), 1.0f);
    *sn = x * ps; *cs = pc;
}
// pybullet.getQuaternionFromEuler (hover_free.py:183,255), normalised.  WIDE: any angles (the reset
// samples yaw in +-2 pi); otherwise |rpy| <= (pi, pi/2, pi) as clipped by compute_observation.
template <bool WIDE>
__device__ __forceinline__ void rpy_to_quat(const float rpy[3], float q[4]) {
    float sp, cp, st, ct, ss, cs;
    if (WIDE) {
        sincosf(0.5f * rpy[0], &sp, &cp);
        sincosf(0.5f * rpy[1], &st, &ct);
        sincosf(0.5f * rpy[2], &ss, &cs);
    } else {
        sincos_half(0.5f * rpy[0], &sp, &cp);
        sincos_half(0.5f * rpy[1], &st, &ct);
        sincos_half(0.5f * rpy[2], &ss, &cs);
    }
    q[0] = sp * ct * cs - cp * st * ss;
    q[1] = cp * st * cs + sp * ct * ss;
    q[2] = cp * ct * ss - sp * st * cs;
    q[3] = cp * ct * cs + sp * st * ss;
    float inv = rsqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
#pragma unroll
    for (int i = 0; i < 4; ++i) q[i] *= inv;
}
// quat2euler of distur_gener.py:186-207
__device__ __forceinline__ void quat2euler_hj(const float q[4], float out[3]) {
    float x = q[0], y = q[1], z = q[2], w = q[3];
    out[0] = atan2_fast(2.0f * (w * x + y * z), 1.0f - 2.0f * (x * x + y * y));
    float t2 = 2.0f * (w * y - z * x);
    t2 = fminf(1.0f, fmaxf(-1.0f, t2));
    out[1] = asin_fast(t2);
    out[2] = atan2_fast(2.0f * (w * z + x * y), 1.0f - 2.0f * (y * y + z * z));
}

// ---------------------------------------------------------------------------------- HJ lookup
// Grid.get_index (odp/Grid/GridProcessing.py:53-72): np.searchsorted(side='left') on the float64
// linspace values, then the strict-< choice of the lower neighbour.  Evaluated in float64 so the
// index of a given float64 state is bit-exact; the search starts from an arithmetic guess.
__device__ __forceinline__ int hj_index_dim(double s, int d) {
    const double* g = c_grid[d];
    if (!(s == s)) return CF_HJ_PTS - 1;               /* NaN sorts last: searchsorted 15 -> 14 */
    double t = (s - g[0]) / c_dx[d];
    int i = t < 0.0 ? 0 : (t > 15.0 ? 15 : (int)t);
    while (i < 15 && g[i] < s) ++i;
    while (i > 0 && !(g[i - 1] < s)) --i;
    if (i > 0 && (i == 15 || fabs(s - g[i - 1]) < fabs(s - g[i]))) i -= 1;
    return i;
}
// The same index for a float32 state, without float64: the rule above is monotone in s, so for
// each cell boundary j the host finds (by bisection over float32 values, with the float64 rule)
// the smallest float32 whose index is >= j+1; index = #{j : s >= thr[d][j]}.  thr is [6][16]
// (14 thresholds, NaN, NaN) in shared or global memory; lo/inv_dx give a guess within +-1.
__device__ __forceinline__ int hj_index_f32(float s, int d, const float* __restrict__ thr, const DevCfg& cfg) {
    const float* t = thr + 16 * d;
    float g = clampf((s - cfg.hj_lo[d]) * cfg.hj_inv_dx[d], 0.0f, 14.0f);
    int i = (int)(g + 0.5f);
    if (s >= t[i]) i += 1;                  /* t[14] = t[15] = NaN: never true */
    else if (i > 0 && s < t[i - 1]) i -= 1;
    return (s != s) ? CF_HJ_PTS - 1 : i;
}
__device__ __forceinline__ float sgnf(float v) { return v > 0.f ? 1.f : (v < 0.f ? -1.f : 0.f); }
// spa_deriv for one non-periodic dim (distur_gener.py:40-99) and the `> 0` test of opt_dstb_non_hcl
// (:106-125).  float32 subtraction exactly where numpy keeps float32, float64 division by dx.
__device__ __forceinline__ bool hj_deriv_positive(const float* __restrict__ V, long c, int i, long stride, double dx) {
    float vc = __ldg(V + c);
    float a, b;
    if (i == 0) {
        float vn = __ldg(V + c + stride);
        float ext = __fmul_rn(fabsf(__fsub_rn(vn, vc)), sgnf(vc));
        float lb = __fadd_rn(vc, ext);
        a = __fsub_rn(vc, lb); b = __fsub_rn(vn, vc);
    } else if (i == CF_HJ_PTS - 1) {
        float vp = __ldg(V + c - stride);
        float ext = __fmul_rn(fabsf(__fsub_rn(vc, vp)), sgnf(vc));
        float rb = __fadd_rn(vc, ext);
        a = __fsub_rn(vc, vp); b = __fsub_rn(rb, vc);
    } else {
        a = __fsub_rn(vc, __ldg(V + c - stride)); b = __fsub_rn(__ldg(V + c + stride), vc);
    }
    double left = (double)a / dx, right = (double)b / dx;
    return (left + right) / 2.0 > 0.0;
}
__device__ __forceinline__ int hj_cell(const int idx[6]) {
    return ((((idx[0] * 15 + idx[1]) * 15 + idx[2]) * 15 + idx[3]) * 15 + idx[4]) * 15 + idx[5];
}
// bit0: dV/dx3 > 0, bit1: dV/dx4 > 0
__device__ __forceinline__ int hj_signbits_value(const float* __restrict__ V, const int idx[6]) {
    long c = hj_cell(idx);
    int b0 = hj_deriv_positive(V, c, idx[3], 225, c_dx[3]) ? 1 : 0;
    int b1 = hj_deriv_positive(V, c, idx[4], 15, c_dx[4]) ? 2 : 0;
    return b0 | b1;
}
__device__ __forceinline__ int hj_signbits_packed(const uint8_t* __restrict__ P, const int idx[6]) {
    int c = hj_cell(idx);
    return (__ldg(P + (c >> 2)) >> (2 * (c & 3))) & 3;
}
template <int HJ_MODE>
__device__ __forceinline__ int hj_signbits(const DevTables& tab, int level_idx, const int idx[6]) {
    if (HJ_MODE == CF_HJ_SIGN_TABLE) return hj_signbits_packed(tab.signs + (size_t)level_idx * CF_SIGN_STRIDE, idx);
    return hj_signbits_value(tab.value[level_idx], idx);
}
// distur_gener (distur_gener.py:19-183) reduced to the two torque components the physics uses: the sign bits
// of the level's table at the nearest grid cell of (rpy, body rates).  Split in two so that the gather is in
// flight while the motor model runs: `issue` returns the raw table byte (packed mode) or the decoded bits
// (value-table mode) plus the shift, `hj_bits` extracts the two bits where the torque is needed.
// Packed mode fetches the aligned 32-bit word that holds the cell's two bits with an asynchronous copy into a
// per-thread shared-memory slot (cp.async, SASS LDGSTS): no register scoreboard is tied up while it is in flight,
// so nothing issued between `issue` and `hj_bits` (the second half of the step's draws, the motor model) can
// end up waiting for the gather by scoreboard aliasing.
struct HjFetch { uint32_t raw; int shift; uint32_t* slot; };
__device__ __forceinline__ void cp_async_4(uint32_t* dst_smem, const void* src_gmem) {
#ifdef CF_GATHER_EVICT_LAST
    // the few MB of table cells the envs actually visit should outlive the ~46 MB of state and I/O streamed per step
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 4, %2;"
                 ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src_gmem), "l"(pol) : "memory");
#else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src_gmem) : "memory");
#endif
    asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int HJ_MODE>
__device__ __forceinline__ HjFetch hj_disturbance_issue(const float q[4], const float wb[3], const DevCfg& cfg,
                                                        const DevTables& tab, const float* thr, int level_idx, uint32_t* slot) {
    float rpy[3];
    quat2euler_hj(q, rpy);
    int idx[6];
#pragma unroll
    for (int k = 0; k < 3; ++k) { idx[k] = hj_index_f32(rpy[k], k, thr, cfg); idx[3 + k] = hj_index_f32(wb[k], 3 + k, thr, cfg); }
    HjFetch f;
    f.slot = nullptr;
    if (HJ_MODE == CF_HJ_SIGN_TABLE) {
        const int c = hj_cell(idx);
        const uint8_t* base = tab.signs + (size_t)level_idx * CF_SIGN_STRIDE;      // slab and stride are 256-byte multiples
        if (slot) {
#ifdef CF_PROBE_NOGATHER
            *slot = (uint32_t)c;                               // tools/ only: the step without the table gather's latency
#else
            cp_async_4(slot, base + 4 * (c >> 4));
#endif
            f.raw = 0u; f.slot = slot; f.shift = 2 * (c & 15);
        } else {
            f.raw = __ldg(base + (c >> 2));
            f.shift = 2 * (c & 3);
        }
    } else {
        f.raw = (uint32_t)hj_signbits_value(tab.value[level_idx], idx);
        f.shift = 0;
    }
    return f;
}
__device__ __forceinline__ int hj_bits(const HjFetch& f) {
    uint32_t raw = f.raw;
    if (f.slot) {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        raw = *reinterpret_cast<volatile uint32_t*>(f.slot);
    }
    return (int)((raw >> f.shift) & 3u);
}

// ---------------------------------------------------------------------------------- sensors
// SensorNoise.add_noise_to_omega (envs/sensors.py:121-134) at dt = 1/SIM_FREQ, followed by
// LowPassFilter.apply (envs/utils.py:102-105, ratio T_s/T = 0.5; base.py:107).
// zb, zr, zt: bias-walk, random-walk, turn-on standard normals.
__device__ __forceinline__ void gyro_update(float bias[3], float lpf[3], const DevCfg& cfg, const float wb[3],
                                            const float zb[3], const float zr[3], const float zt[3]) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        bias[i] = cfg.gyro_pi * bias[i] + cfg.gyro_sb * zb[i];
        float om = ((wb[i] + bias[i]) + cfg.gyro_rw * zr[i]) + cfg.gyro_ton * zt[i];
        lpf[i] = 0.5f * lpf[i] + 0.5f * om;
    }
}
// The 100 Hz branch of compute_observation (hover_free.py:168-200) with SensorNoise.add_noise
// (sensors.py:75-119).  Normal blocks zb0.. cover 18 slots starting at slot offset `off` inside
// block zb0: pos3, vel3, gyro9 (bias, random walk, turn-on), rot3; uniforms: pos3, rot3.
// the part of it that does not involve the gyro: noisy position, velocity and attitude quaternion -> obs[0:10]
__device__ __forceinline__ void observe_nongyro(const float pos[3], const float vel[3], const float rpy[3], const DevCfg& cfg,
                                                const float z[18], const float u[6], float obs[13]) {
    const float kPosStd = 0.002f, kPosUnif = 0.001f, kVelStd = 0.01f;     // sensors.py:18-25
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        obs[i] = pos[i] + (kPosStd * z[i] + (-kPosUnif + (2.0f * kPosUnif) * u[i]));
        obs[7 + i] = vel[i] + kVelStd * z[3 + i];                          // vel_unif_range = 0
    }
    const float lo[3] = {-kPi, -0.5f * kPi, -kPi}, hi[3] = {kPi, 0.5f * kPi, kPi};
    float rot[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float th = cfg.quat_std * z[15 + i] + (-cfg.quat_unif + (2.0f * cfg.quat_unif) * u[3 + i]);
        rot[i] = fminf(hi[i], fmaxf(lo[i], rpy[i] + th));
    }
    rpy_to_quat<false>(rot, obs + 3);
}
__device__ __forceinline__ void observe_full(const float pos[3], const float vel[3], const float rpy[3], const float wb[3],
                                             float bias[3], float lpf[3], const DevCfg& cfg, const float z[18],
                                             const float u[6], float obs[13]) {
    observe_nongyro(pos, vel, rpy, cfg, z, u, obs);
    gyro_update(bias, lpf, cfg, wb, z + 6, z + 9, z + 12);
#pragma unroll
    for (int i = 0; i < 3; ++i) obs[10 + i] = lpf[i];
}

// ---------------------------------------------------------------------------------- physics substep
// Rigid-body + motor state held in registers across the two substeps of a control step.
struct Body {
    float pos[3], quat[4], vel[3], ow[3], x[4], ou[4];
};
// PybulletPhysicsWithAdversary.step_forward (envs/physics.py:213-250) with
// CrazyFlieAgent.apply_action (agents.py:259-298), PWM.act (control.py:94-100), OUNoise.noise
// (envs/utils.py:128-133) and Bullet's stepSimulation as restated by the oracle (B1-B9:
// semi-implicit Euler, exponential-map quaternion update, gyroscopic term, 0.04 damping).
// `delayed` is the action leaving the latency ring in this substep (agents.py:268-276);
// invJ / invm are reciprocals of the (domain-randomised) inertia and mass.
// R, wb: rotation matrix and body rates of the state before the substep (frame_of).
// d0, d1: adversary roll / pitch torque; with DECODE they are extracted here from the table fetch issued at
// the top of the step (distur_gener.py:106-125: -dmax where dV/dx > 0, else +dmax), as late as the data flow allows.
template <bool DECODE>
__device__ __forceinline__ void physics_substep(Body& b, const float R[9], const float wb[3], const DrParams& p,
                                                const float invJ[3], float invm, const DevCfg& cfg, const float delayed[4],
                                                float& d0, float& d1, const float zou[4], const HjFetch& hjf, bool hj, float dmax) {
    const float kFtf0 = 1.56e-5f, kArm = 0.028f, kPropZ = 0.0108f, kG = 9.81f;
    const float kDrag[3] = {9.1785e-7f, 9.1785e-7f, 10.311e-7f};
    float f[4], t[4];
    float rpm_sum = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float c = clampf(delayed[i], -1.0f, 1.0f);
        float pwm = 30000.0f + c * 30000.0f;
        b.ou[i] = b.ou[i] + (0.15f * (0.0f - b.ou[i]) + cfg.ou_sigma * zou[i]);
        float rot = fsqrt(pwm * (1.0f / 60000.0f));
        b.x[i] = (1.0f - p.B[i]) * b.x[i] + p.B[i] * rot;
        float xx = b.x[i] * b.x[i];
        float nn = clampf((1.0f + b.ou[i]) * xx, 0.0f, 1.0f);
        f[i] = p.K[i] * nn;
        t[i] = p.ftf1 * f[i] + kFtf0;
        rpm_sum += xx;
    }
    rpm_sum *= 2.0f * kPi * 25000.0f / 60.0f;                 // sum_i 2 pi (25000 x_i^2) / 60, physics.py:236
    float zt = ((-t[0] + t[1]) - t[2]) + t[3];
    float vb[3];
    matTvec(R, b.vel, vb);
    // drag (physics.py:232-241): R @ (k * v_world) handed to Bullet as a LINK_FRAME force
    float kv[3], drag_link[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) kv[i] = (-1.0f * kDrag[i] * rpm_sum) * b.vel[i];
    matvec(R, kv, drag_link);
    float fe[4] = {f[0], f[1], f[2], f[3]};
    if (cfg.ground_effect) {                               // physics.py:27-58; reference leaves it off
        const float kPx[4] = {0.028f, -0.028f, -0.028f, 0.028f}, kPy[4] = {-0.028f, -0.028f, 0.028f, 0.028f};
        const float kGnd = 11.36859f, kPropR = 2.31348e-2f;
        float rpy[3];
        quat_to_rpy_bullet(b.quat, rpy);
        if (fabsf(rpy[0]) < 0.5f * kPi && fabsf(rpy[1]) < 0.5f * kPi) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float pz = b.pos[2] + R[6] * kPx[i] + R[7] * kPy[i] + R[8] * kPropZ;
                pz = fmaxf(pz, cfg.gnd_hclip);
                float q = kPropR / (4.0f * pz);
                fe[i] += f[i] * kGnd * (q * q);
            }
        }
    }
    // body-frame wrench: thrust at the prop positions (agents.py:311-321, urdf:59,85,111,136),
    // yaw torque (:330-337), adversary roll/pitch torques (:517-533)
    float tau[3] = {kArm * (-fe[0] - fe[1] + fe[2] + fe[3]), kArm * (-fe[0] + fe[1] + fe[2] - fe[3]), zt};
    float fb[3] = {drag_link[0], drag_link[1], drag_link[2] + (fe[0] + fe[1] + fe[2] + fe[3])};
    float gz = -kG * p.mass;
    float gb[3] = {R[6] * gz, R[7] * gz, R[8] * gz};        // R^T (0,0,-g m)
    float Jw[3] = {p.J[0] * wb[0], p.J[1] * wb[1], p.J[2] * wb[2]};
    if (cfg.gyroscopic) {
        tau[0] -= wb[1] * Jw[2] - wb[2] * Jw[1];
        tau[1] -= wb[2] * Jw[0] - wb[0] * Jw[2];
        tau[2] -= wb[0] * Jw[1] - wb[1] * Jw[0];
    }
    float wn = fsqrt(wb[0] * wb[0] + wb[1] * wb[1] + wb[2] * wb[2]);
    float vn = fsqrt(vb[0] * vb[0] + vb[1] * vb[1] + vb[2] * vb[2]);
    float ka = cfg.ang_damp, kl = cfg.lin_damp;
    float alpha_b[3], acc_b[3], alpha_w[3], acc_w[3];
    if (DECODE && hj) {
        const int bits = hj_bits(hjf);
        d0 = (bits & 1) ? -dmax : dmax;
        d1 = (bits & 2) ? -dmax : dmax;
    }
    tau[0] += d0; tau[1] += d1;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        alpha_b[i] = (tau[i] - Jw[i] * (ka + ka * wn)) * invJ[i];
        acc_b[i] = ((fb[i] + gb[i]) - p.mass * vb[i] * (kl + kl * vn)) * invm;
    }
    matvec(R, alpha_b, alpha_w);
    matvec(R, acc_b, acc_w);
    float dt = p.dt;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        b.ow[i] += dt * alpha_w[i];
        b.vel[i] += dt * acc_w[i];
        b.pos[i] += dt * b.vel[i];
    }
    // exponential-map quaternion update (btMultiBody::stepPositionsMultiDof): dq = (w sin(h)/|w|, cos(h)),
    // h = |w| dt / 2 clamped to pi/8.  sin(h)/h and cos(h) as even polynomials in h (|h| <= pi/8: truncation
    // < 2e-9), which is also Bullet's own small-angle branch; no range reduction, no division.
    float w0 = b.ow[0], w1 = b.ow[1], w2 = b.ow[2];
    float h = fminf(0.5f * dt * fsqrt(w0 * w0 + w1 * w1 + w2 * w2), 0.125f * kPi);   // Bullet clamps the angle only
    float h2 = h * h;
    float sinc = 1.0f + h2 * (-1.0f / 6.0f + h2 * (1.0f / 120.0f + h2 * (-1.0f / 5040.0f + h2 * (1.0f / 362880.0f))));
    float cw = 1.0f + h2 * (-0.5f + h2 * (1.0f / 24.0f + h2 * (-1.0f / 720.0f + h2 * (1.0f / 40320.0f))));
    float s = (0.5f * dt) * sinc;                           // sin(h) / (2h/dt): Bullet divides by the clamped rate
    float dq[4] = {w0 * s, w1 * s, w2 * s, cw}, q[4];
    const float* bq = b.quat;
    q[0] = dq[3] * bq[0] + dq[0] * bq[3] + dq[1] * bq[2] - dq[2] * bq[1];
    q[1] = dq[3] * bq[1] - dq[0] * bq[2] + dq[1] * bq[3] + dq[2] * bq[0];
    q[2] = dq[3] * bq[2] + dq[0] * bq[1] - dq[1] * bq[0] + dq[2] * bq[3];
    q[3] = dq[3] * bq[3] - dq[0] * bq[0] - dq[1] * bq[1] - dq[2] * bq[2];
    float inv = rsqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
#pragma unroll
    for (int i = 0; i < 4; ++i) b.quat[i] = q[i] * inv;
}

// ---------------------------------------------------------------------------------- reset
__device__ __forceinline__ float unif(float lo, float hi, float u) { return lo + (hi - lo) * u; }
__device__ __forceinline__ float dr_draw(float def, float f, float u) { float b = f * def; return unif(def - b, def + b, u); }

// Constructor state of an env (DroneBaseEnv.__init__ + CrazyFlieAgent.__init__): zero state at
// (0,0,1), nominal parameters; `fresh` marks that the constructor-time compute_observation()
// (base.py:141), which advances gyro_bias once, has not been accounted for yet.
__device__ __forceinline__ void init_env(const EnvRef e, const DevCfg& cfg) {
#pragma unroll
    for (int k = 0; k < CF_STATE_WORDS; ++k) e.f(k, 0.0f);
    e.f(CF_S_POS + 2, 1.0f); e.f(CF_S_QUAT + 3, 1.0f);
    e.f(CF_S_DT, cfg.sim_dt); e.f(CF_S_MASS, 0.030f);
    e.f(CF_S_INERTIA, 1.33e-5f); e.f(CF_S_INERTIA + 1, 1.33e-5f); e.f(CF_S_INERTIA + 2, 2.64e-5f);
    e.f(CF_S_FTF1, 5.96e-3f);
#pragma unroll
    for (int i = 0; i < 4; ++i) { e.f(CF_S_MOTOR_B + i, cfg.nomB); e.f(CF_S_MOTOR_K + i, cfg.nomK); }
    e.u(CF_S_EP_LEN, 0u); e.u(CF_S_RNG_CTR, 0u);
    e.u(CF_S_MISC, misc_pack(cfg.fixed_level_idx, 0, 0, true));
}

// constructor-time compute_observation() (base.py:141): gyro_bias <- sigma_b * N(0,1), event tag 3
__device__ __forceinline__ void consume_fresh(const EnvRef e, const DevCfg& cfg, long long gid, bool noise) {
    uint32_t m = e.u(CF_S_MISC);
    if (!misc_fresh(m)) return;
    if (noise && !cfg.injected) {
        float4 z = make_rng(gid, 0u, 3u).normals(cfg, 0);
        e.f(CF_S_GYRO_BIAS, cfg.gyro_sb * z.x); e.f(CF_S_GYRO_BIAS + 1, cfg.gyro_sb * z.y); e.f(CF_S_GYRO_BIAS + 2, cfg.gyro_sb * z.z);
    }
    e.u(CF_S_MISC, m & 0x7fffffffu);
}

// reset(): envs/base.py:420-464, task_specific_reset hover_free.py:237-289 / hover.py:1463-1513,
// apply_domain_randomization base.py:241-298 (+ agents.py:208-224), RandomHJ level draw
// hover_free.py:536.  Draw slots: z[0:4] motor x, z[4:12] latency ring, z[12:30] and z[30:48]
// the two compute_observation calls; u[0:14] initial state, u[14:29] DR, u[29:35], u[35:41]
// observation uniforms, u[41] Boltzmann.
// The reset is written as independent tasks so that the step kernel can run them on different warps
// (one finished env per lane); reset_env below composes the same tasks serially for the reset kernel.

// What the gyro task needs from the previous episode and from the pose task.
struct ResetCtx {
    float stale_wb[3], bias0[3];      // previous state's body rate (base.py:444 runs before :448), gyro bias (never reset)
    float wb[3];
};
// the sampled kinematic state of task_specific_reset (hover_free.py:249-274), a function of the uniforms only
struct ResetKin { float pos[3], rpy[3], vel[3], rate[3], rpy_t[3]; };
__device__ __forceinline__ void reset_sample_kin(const DevCfg& cfg, const float* un, ResetKin& k) {
#pragma unroll
    for (int i = 0; i < 3; ++i) { k.pos[i] = 0.f; k.rpy[i] = 0.f; k.vel[i] = 0.f; k.rate[i] = 0.f; }
    k.pos[2] = 1.f;
    if (cfg.reset_dist) {
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            k.pos[i] += unif(-cfg.init_pos, cfg.init_pos, un[i]);
            k.rpy[i] = unif(-cfg.init_angle, cfg.init_angle, un[3 + i]);
            k.vel[i] = unif(-cfg.init_vel, cfg.init_vel, un[7 + i]);
            k.rate[i] = unif(-cfg.init_rate, cfg.init_rate, un[10 + i]);
        }
        k.rpy[2] = unif(-cfg.init_yaw, cfg.init_yaw, un[6]);
        k.rate[2] = unif(-cfg.init_yaw_rate, cfg.init_yaw_rate, un[13]);
    }
    // getEulerFromQuaternion(getQuaternionFromEuler(rpy)) (agents.py:446): |roll|, |pitch| <= init_angle < pi/2, so it
    // returns the sampled roll and pitch and the yaw wrapped into [-pi, pi]
    k.rpy_t[0] = k.rpy[0]; k.rpy_t[1] = k.rpy[1];
    k.rpy_t[2] = k.rpy[2] > kPi ? k.rpy[2] - 2.0f * kPi : (k.rpy[2] < -kPi ? k.rpy[2] + 2.0f * kPi : k.rpy[2]);
}
// last_action after reset = action_buffer[-1,:] (agents.py:386), a function of the normals only
__device__ __forceinline__ void reset_sample_la(const DevCfg& cfg, const float* nz, float la[4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) la[i] = cfg.reset_dist ? clampf(cfg.hover_a + 0.02f * nz[8 + i], -1.0f, 1.0f) : 0.0f;
}
// task 0 (reads the OLD state): stale rpy_dot and the gyro bias carried over (sensors.py:68)
__device__ __forceinline__ void reset_task_stale(const EnvRef e, ResetCtx& c) {
    float q[4], ow[3], R[9];
    e.get4(CF_S_QUAT, q); e.get3(CF_S_OMEGA_W, ow);
    quat_to_R(q, R);
    matTvec(R, ow, c.stale_wb);
    e.get3(CF_S_GYRO_BIAS, c.bias0);
}
// task 1: drone.reset() (agents.py:377-386) + task_specific_reset: pose, velocities, motor state, latency ring
__device__ __forceinline__ void reset_task_pose(const EnvRef e, const DevCfg& cfg, const float* nz, const float* un, float wb[3]) {
    ResetKin k;
    reset_sample_kin(cfg, un, k);
    float quat[4] = {0.f, 0.f, 0.f, 1.f}, x[4] = {0.f, 0.f, 0.f, 0.f};
    if (cfg.reset_dist) {
#pragma unroll
        for (int i = 0; i < 4; ++i) x[i] = cfg.hover_x + 0.02f * nz[i];
#pragma unroll
        for (int i = 0; i < 8; ++i) e.f(CF_S_RING + i, clampf(cfg.hover_a + 0.02f * nz[4 + i], -1.0f, 1.0f));
        rpy_to_quat<true>(k.rpy, quat);
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) e.f(CF_S_RING + i, 0.0f);
    }
    e.set3(CF_S_POS, k.pos); e.set4(CF_S_QUAT, quat); e.set3(CF_S_VEL, k.vel); e.set4(CF_S_MOTOR_X, x);
    float R[9], ow[3];
    quat_to_R(quat, R);
    matTvec(R, k.rate, ow);                                 // resetBaseVelocity(angularVelocity = R.T @ rpy_dot), hover_free.py:283-289
    e.set3(CF_S_OMEGA_W, ow);
    matTvec(R, ow, wb);                                     // update_information: R.T again (agents.py:452-453)
}
// task 2: apply_domain_randomization (written to the env's words in HBM), episode counters, level draw
__device__ __forceinline__ void reset_task_params(const EnvRef e, const EnvRef g, const DevCfg& cfg, const float* un) {
    int level = misc_level(e.u(CF_S_MISC));
    e.u(CF_S_EP_LEN, 0u); e.f(CF_S_EP_RET, 0.0f);
    if (cfg.dr_on) {
        // un[14..28]: dt, mass, Jxx, Jyy, Jzz, (ftf0: cancels in the yaw torque), ftf1, motor tc[4], t2w[4]
        float f = cfg.dr;
        DrParams p;
        p.dt = dr_draw(cfg.sim_dt, f, un[14]);
        p.mass = dr_draw(0.030f, f, un[15]);
        p.J[0] = dr_draw(1.33e-5f, f, un[16]);
        p.J[1] = dr_draw(1.33e-5f, f, un[17]);
        p.J[2] = dr_draw(2.64e-5f, f, un[18]);
        p.ftf1 = dr_draw(5.96e-3f, f, un[20]);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float mtc = dr_draw(cfg.motor_tc, f, un[21 + i]);
            float t2w = dr_draw(1.8f, f, un[25 + i]);
            float T = fmaxf(mtc, p.dt);
            p.B[i] = p.dt / T;
            p.K[i] = 0.028f * 9.81f * t2w / 4.0f;           // agents.py:224 (0.028 kg, not the URDF mass)
        }
        p.store(g);
    }
    // Boltzmann() at the end of reset, hover_free.py:536: cdf.searchsorted(u, side='right') =
    // #{k : cdf[k] <= u}; for a float32 u that is #{k : u >= boltz_thr[k]} with boltz_thr[k] the smallest
    // float32 >= cdf[k] (derived on the host)
    if (cfg.dstb_kind == CF_DSTB_HJ_BOLTZMANN) {
        const float u = un[41];
        int k = 0;
#pragma unroll
        for (int j = 0; j < 20; ++j) k += (u >= cfg.boltz_thr[j]) ? 1 : 0;
        level = k;
    }
    e.u(CF_S_MISC, misc_pack(level, 0, 0, false));
}
// tasks 3 / 4 (noise on): the gyro-free part of the two compute_observation calls of reset (base.py:453 and,
// via compute_history, :308): WHICH = 0 -> obs(k-1) = row[0:10], WHICH = 1 -> obs(k) = row[17:27] and the
// env's last-emitted observation; task 4 also writes the action history (base.py:459-460)
template <int WHICH>
__device__ __forceinline__ void reset_task_obs_nongyro(const EnvRef e, const DevCfg& cfg, const float* nz, const float* un, float* row) {
    ResetKin k;
    reset_sample_kin(cfg, un, k);
    float ob[13];
    observe_nongyro(k.pos, k.vel, k.rpy_t, cfg, nz + 12 + 18 * WHICH, un + 29 + 6 * WHICH, ob);
#pragma unroll
    for (int i = 0; i < 10; ++i) row[17 * WHICH + i] = ob[i];
    if (WHICH == 1) {
#pragma unroll
        for (int i = 0; i < 10; ++i) e.f(CF_S_OBS_PREV + i, ob[i]);
        float la[4];
        reset_sample_la(cfg, nz, la);
        e.set4(CF_S_ACT_PREV1, la); e.set4(CF_S_ACT_PREV2, la);
#pragma unroll
        for (int i = 0; i < 4; ++i) { row[13 + i] = la[i]; row[30 + i] = la[i]; }
    }
}
// task 5 (noise on): gyro_lpf.set(stale rpy_dot), base.py:444, then the gyro samples of the two observations
__device__ __forceinline__ void reset_task_gyro(const EnvRef e, const ResetCtx& c, const DevCfg& cfg, const float* nz, float* row) {
    float bias[3], lpf[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) { bias[i] = c.bias0[i]; lpf[i] = c.stale_wb[i]; }
    gyro_update(bias, lpf, cfg, c.wb, nz + 12 + 6, nz + 12 + 9, nz + 12 + 12);
#pragma unroll
    for (int i = 0; i < 3; ++i) row[10 + i] = lpf[i];
    gyro_update(bias, lpf, cfg, c.wb, nz + 30 + 6, nz + 30 + 9, nz + 30 + 12);
#pragma unroll
    for (int i = 0; i < 3; ++i) { row[27 + i] = lpf[i]; e.f(CF_S_OBS_PREV + 10 + i, lpf[i]); }
    e.set3(CF_S_GYRO_BIAS, bias); e.set3(CF_S_LPF, lpf);
}

// All standard draws of a reset event, slot order of the oracle: z[48], u[44] (42 used).
__device__ __forceinline__ void reset_draws(const DrawSrc& dw, const DevCfg& cfg, float* z, float* u) {
#pragma unroll
    for (int b = 0; b < CF_RESET_NORMALS / 4; ++b) { float4 v = dw.z(cfg, b); z[4 * b] = v.x; z[4 * b + 1] = v.y; z[4 * b + 2] = v.z; z[4 * b + 3] = v.w; }
#pragma unroll
    for (int b = 0; b < (CF_RESET_UNIFORMS + 3) / 4; ++b) { float4 v = dw.u(cfg, b); u[4 * b] = v.x; u[4 * b + 1] = v.y; u[4 * b + 2] = v.z; u[4 * b + 3] = v.w; }
}

// The whole reset of one env by one thread (reset kernel; in-place resets of the noise-free variant).
// `e` holds the dynamic words (shared or global), `g` the env in global memory (the domain-randomised
// words are written there); nz / un: the event's standard draws; row: obs_dim floats.
template <bool NOISE>
__device__ __forceinline__ void reset_env(const EnvRef e, const EnvRef g, const DevCfg& cfg, const float* nz, const float* un,
                                          float* row) {
    constexpr int OD = NOISE ? 34 : 42;
    ResetCtx c;
    reset_task_stale(e, c);
    reset_task_pose(e, cfg, nz, un, c.wb);
    reset_task_params(e, g, cfg, un);
    if (NOISE) {
        reset_task_obs_nongyro<0>(e, cfg, nz, un, row);
        reset_task_obs_nongyro<1>(e, cfg, nz, un, row);
        reset_task_gyro(e, c, cfg, nz, row);
    } else {
        ResetKin k;
        reset_sample_kin(cfg, un, k);
        float la[4], quat[4], hist[OD], ob[13];
        reset_sample_la(cfg, nz, la);
        e.set3(CF_S_LPF, c.stale_wb);
        e.get4(CF_S_QUAT, quat);
#pragma unroll
        for (int i = 0; i < 3; ++i) { ob[i] = k.pos[i]; ob[7 + i] = k.vel[i]; ob[10 + i] = c.wb[i]; }
#pragma unroll
        for (int i = 0; i < 4; ++i) ob[3 + i] = quat[i];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
#pragma unroll
            for (int i = 0; i < 13; ++i) hist[21 * h + i] = ob[i];
#pragma unroll
            for (int i = 0; i < 4; ++i) { hist[21 * h + 13 + i] = la[i]; hist[21 * h + 17 + i] = la[i]; }
        }
#pragma unroll
        for (int i = 0; i < 13; ++i) e.f(CF_S_OBS_PREV + i, ob[i]);
        e.set4(CF_S_ACT_PREV1, la); e.set4(CF_S_ACT_PREV2, la);
        store_row<OD>(row, hist);
    }
}

// The 44 standard normals of a step event, generated ahead of their use into shared-memory scratch (they depend on the env
// id and the event counter only): slot k of this lane = normal k.  Slots 0..31 live in the warp's own 32 observation rows,
// word-major ([slot][lane], conflict-free, lane-private; the rows are written only after the last slot has been read), slots
// 32..43 in the CTA's extra scratch ([slot][thread]).  Slots 0..20 are reused late in the step: once the physics and the gyro
// samples have consumed z[0:26], the env's 21 history words are copied into them asynchronously from HBM (step_env).
// The step's six live uniforms are generated where the observation is assembled; they never touch shared memory.
#define CF_DRAW_BLOCKS (CF_STEP_NORMALS / 4)   /* 11 Philox blocks */
#define CF_ROW_SLOTS 32
#define CF_EXTRA_SLOTS (4 * CF_DRAW_BLOCKS - CF_ROW_SLOTS)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    }
}
struct StepDraws {
    float* s;             // s[k * 32]: slot k < 32
    float* x;             // x[(k - 32) * CF_TILE]: slot k >= 32
    // warp-specialised kernel: the normals are written by producer warps in four stages (blocks 0-1 | 2-4 | 5-6 | 7-10), each
    // announced on an mbarrier; null in the single-role kernels (the waits fold away)
    uint64_t* stage_bar;
    uint32_t stage_parity;
    __device__ __forceinline__ void wait_stage(int st) const { if (stage_bar) mbar_wait(stage_bar + st, stage_parity); }
    __device__ __forceinline__ float* slot(int k) const { return k < CF_ROW_SLOTS ? s + k * 32 : x + (k - CF_ROW_SLOTS) * CF_TILE; }
    __device__ __forceinline__ void put(int b, float4 v) const {     // b: run-time block index
        float* p = slot(4 * b);
        const int st = 4 * b < CF_ROW_SLOTS ? 32 : CF_TILE;
        p[0] = v.x; p[st] = v.y; p[2 * st] = v.z; p[3 * st] = v.w;
    }
    template <int B>
    __device__ __forceinline__ float4 z() const {
        const float* p = slot(4 * B);
        constexpr int st = 4 * B < CF_ROW_SLOTS ? 32 : CF_TILE;
        return make_float4(p[0], p[st], p[2 * st], p[3 * st]);
    }
};
// which normal blocks of a step event are live: all 11 with sensor noise; without it the OU draws z[0:4] (block 0) and z[13:17]
// (blocks 3, 4)
template <bool NOISE>
__device__ __forceinline__ bool step_block_live(int b) { return NOISE || b == 0 || b == 3 || b == 4; }

// 4-byte asynchronous copy global -> shared without committing a group (LDGSTS); cp_async_commit closes the group
__device__ __forceinline__ void cp_async_4_nocommit(float* dst_smem, const float* src_gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src_gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// The HJ table fetch of a step, issued from the tile as soon as it has landed (step_env consumes it).
template <int HJ_MODE>
__device__ __forceinline__ HjFetch step_issue_hj(const EnvRef e, const DevCfg& cfg, const DevTables& tab, const float* thr,
                                                 uint32_t* slot) {
    HjFetch f;
    f.raw = 0u; f.shift = 0; f.slot = nullptr;
    const int level = misc_level(e.u(CF_S_MISC));
    if ((cfg.dstb_kind == CF_DSTB_HJ_FIXED || cfg.dstb_kind == CF_DSTB_HJ_BOLTZMANN) && level > 0) {
        float q[4], ow[3], R[9], wb[3];
        e.get4(CF_S_QUAT, q); e.get3(CF_S_OMEGA_W, ow);
        quat_to_R(q, R); matTvec(R, ow, wb);
        f = hj_disturbance_issue<HJ_MODE>(q, wb, cfg, tab, thr, level, slot);
    }
    return f;
}

struct StepOut {
    float rew, cost, ep_ret;
    int done, truncated, ep_len;
};

// step(): hover_free.py:391-444 / 608-665 / 778-833 (hover.py twins) under gym TimeLimit.
// Draw slots: z[0:4] OU substep 1, z[4:13] gyro after substep 1, z[13:17] OU substep 2,
// z[17:26] gyro after substep 2, z[26:44] the full compute_observation of compute_history;
// u[0:6] its position / attitude uniforms, u[6:8] the random-box disturbance.
// row: obs_dim floats = [o(k-1), u(k-2), o(k), u(k-1)] (base.py:305-321).  No reset here.
// e: dynamic words (shared memory in the step kernel); p: the env's domain-randomised parameters.
template <bool NOISE, int HJ_MODE>
__device__ __forceinline__ StepOut step_env(const EnvRef e, const DrParams& p, const DevCfg& cfg, const DevTables& tab,
                                            const HjFetch& hjf, const float a[4], const StepDraws& dw, const DrawSrc& src, float* row,
                                            unsigned wmask) {
    constexpr int OD = NOISE ? 34 : 42;
    const uint32_t misc = e.u(CF_S_MISC);
    const int ep_len = (int)e.u(CF_S_EP_LEN);
    Body b;
    e.get4(CF_S_QUAT, b.quat); e.get3(CF_S_OMEGA_W, b.ow);
    // disturbance, once per control step: start the table gather first, the motor model hides its latency
    const int level = misc_level(misc);
    const bool hj = (cfg.dstb_kind == CF_DSTB_HJ_FIXED || cfg.dstb_kind == CF_DSTB_HJ_BOLTZMANN) && level > 0;
    float R[9], wb[3];
    quat_to_R(b.quat, R); matTvec(R, b.ow, wb);             // frame of the pre-step state: HJ lookup and substep 1
    float d0 = 0.0f, d1 = 0.0f;
    if (cfg.dstb_kind == CF_DSTB_RANDOM_BOX) {                // dstb_space.sample(), hover_free.py:868-878
        const float4 ub = src.u(cfg, 1);                    // u[6], u[7] of the event
        d0 = unif(-1e-3f, 1e-3f, ub.z);
        d1 = unif(-1e-3f, 1e-3f, ub.w);
    }
    e.get3(CF_S_POS, b.pos); e.get3(CF_S_VEL, b.vel); e.get4(CF_S_MOTOR_X, b.x); e.get4(CF_S_OU, b.ou);
    // latency ring (agents.py:268-276): substep 1 reads slot ring_idx, substep 2 the other one; both
    // are overwritten with the new action and ring_idx returns to its value
    const int ridx = misc_ring(misc) & 1;
    float del1[4], del2[4];
    e.get4_low(CF_S_RING + 4 * ridx, del1); e.get4_low(CF_S_RING + 4 * (ridx ^ 1), del2);
    const float invJ[3] = {frcp(p.J[0]), frcp(p.J[1]), frcp(p.J[2])};
    const float invm = frcp(p.mass);
    float bias[3], lpf[3];
    if (NOISE) { e.get3(CF_S_GYRO_BIAS, bias); e.get3(CF_S_LPF, lpf); }
    dw.wait_stage(0);
    float4 zq = dw.z<0>();
    const float dmax = ((float)level / 10.0f) * 5.3e-3f;     // distur_gener.py:152,160
    // substeps; the reference calls compute_observation() after each (base.py:494-501): only its
    // gyro sample survives (bias walk + low-pass), the rest of that call is overwritten below
    {
        const float zou[4] = {zq.x, zq.y, zq.z, zq.w};
        physics_substep<true>(b, R, wb, p, invJ, invm, cfg, del1, d0, d1, zou, hjf, hj, dmax);
    }
    quat_to_R(b.quat, R); matTvec(R, b.ow, wb);             // frame after substep 1: gyro sample and substep 2
    float4 z3;
    dw.wait_stage(1);
    if (NOISE) {
        float4 z1 = dw.z<1>(), z2 = dw.z<2>();
        z3 = dw.z<3>();
        const float zb[3] = {z1.x, z1.y, z1.z}, zr[3] = {z1.w, z2.x, z2.y}, zt[3] = {z2.z, z2.w, z3.x};
        gyro_update(bias, lpf, cfg, wb, zb, zr, zt);
    } else {
        z3 = dw.z<3>();
    }
    float4 z4 = dw.z<4>();
    {
        const float zou[4] = {z3.y, z3.z, z3.w, z4.x};
        physics_substep<false>(b, R, wb, p, invJ, invm, cfg, del2, d0, d1, zou, hjf, hj, dmax);
    }
    // final rigid-body state back to the tile; true attitude and body rates (agents.py:434-453)
    e.set3(CF_S_POS, b.pos); e.set4(CF_S_QUAT, b.quat); e.set3(CF_S_VEL, b.vel); e.set3(CF_S_OMEGA_W, b.ow);
    e.set4(CF_S_MOTOR_X, b.x); e.set4(CF_S_OU, b.ou);
    e.set4(CF_S_RING, a); e.set4(CF_S_RING + 4, a);
    float rpy[3];
    quat_to_rpy_bullet(b.quat, rpy);
    quat_to_R(b.quat, R); matTvec(R, b.ow, wb);
    float4 z5, z6;
    if (NOISE) {
        dw.wait_stage(2);
        z5 = dw.z<5>(); z6 = dw.z<6>();
        const float zb[3] = {z4.y, z4.z, z4.w}, zr[3] = {z5.x, z5.y, z5.z}, zt[3] = {z5.w, z6.x, z6.y};
        gyro_update(bias, lpf, cfg, wb, zb, zr, zt);
    }
    // The 21 history words (u(k-1), u(k-2), last emitted obs13: device words 39..59 of the tile in HBM, L2-prefetched at CTA
    // start) are only copied through into the row.  Draw slots 0..25 of this lane are dead by now (z[0:26] consumed): the words
    // land there by asynchronous copy - no register is held and nothing waits while the observation below is computed.
    float* const hs = dw.s;
#ifndef CF_PROBE_NOHIST                                       /* tools/ only (wrong results): the step without its history loads */
#pragma unroll
    for (int k = 0; k < CF_HIST_WORDS; ++k) cp_async_4_nocommit(hs + k * 32, e.h + (CF_BULK_WORDS + k) * CF_TILE);
    cp_async_commit();
#endif
    // compute_history, base.py:305-321.  Action history with the reference's aliased view of
    // action_buffer[-1,:] in the first two steps of an episode (agents.py:386, base.py:459-460):
    // [a,a] then [a,u(k-1)] then normal
    const bool first = ep_len == 0, second = ep_len == 1;
    float hist[OD], on[13];
    if (NOISE) {
        dw.wait_stage(3);
        float4 z7 = dw.z<7>(), z8 = dw.z<8>(), z9 = dw.z<9>(), z10 = dw.z<10>();
        const float z[18] = {z6.z, z6.w, z7.x, z7.y, z7.z, z7.w, z8.x, z8.y, z8.z, z8.w, z9.x, z9.y, z9.z, z9.w,
                             z10.x, z10.y, z10.z, z10.w};
        const float4 u0 = src.u(cfg, 0), u1 = src.u(cfg, 1);   // u[0:6]: position and attitude uniforms of the observation
        const float u[6] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y};
        observe_full(b.pos, b.vel, rpy, wb, bias, lpf, cfg, z, u, on);
        e.set3(CF_S_GYRO_BIAS, bias); e.set3(CF_S_LPF, lpf);
    } else {
#pragma unroll
        for (int i = 0; i < 3; ++i) { on[i] = b.pos[i]; on[7 + i] = b.vel[i]; on[10 + i] = wb[i]; }
#pragma unroll
        for (int i = 0; i < 4; ++i) on[3 + i] = b.quat[i];
    }
    cp_async_wait_all();
    float ap1[4], ap2[4], h0[4], h1[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        ap1[i] = hs[(cf_dev_word(CF_S_ACT_PREV1) - CF_BULK_WORDS + i) * 32];
        ap2[i] = hs[(cf_dev_word(CF_S_ACT_PREV2) - CF_BULK_WORDS + i) * 32];
    }
#pragma unroll
    for (int i = 0; i < 13; ++i) hist[i] = hs[(cf_dev_word(CF_S_OBS_PREV) - CF_BULK_WORDS + i) * 32];
#ifdef CF_PROBE_NOHIST
#pragma unroll
    for (int i = 0; i < 4; ++i) { ap1[i] = a[i]; ap2[i] = a[i]; }
#endif
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        h0[i] = (first || second) ? a[i] : ap2[i];
        h1[i] = first ? a[i] : ap1[i];
    }
    e.set4(CF_S_ACT_PREV2, ap1); e.set4(CF_S_ACT_PREV1, a);
#pragma unroll
    for (int i = 0; i < 13; ++i) e.f(CF_S_OBS_PREV + i, on[i]);
    if (NOISE) {
#pragma unroll
        for (int i = 0; i < 13; ++i) hist[17 + i] = on[i];
#pragma unroll
        for (int i = 0; i < 4; ++i) { hist[13 + i] = h0[i]; hist[30 + i] = h1[i]; }
    } else {
        // 42-d row without sensor noise: [o(k-1) 13, u(k-1)?..] as the reference's compute_history lays it out for obs 17 + 4
#pragma unroll
        for (int i = 0; i < 4; ++i) { hist[13 + i] = ap1[i]; hist[17 + i] = h0[i]; hist[38 + i] = h1[i]; hist[34 + i] = a[i]; }
#pragma unroll
        for (int i = 0; i < 13; ++i) hist[21 + i] = on[i];
    }
    __syncwarp(wmask);                                      // every lane has read its draws: the rows may overwrite the scratch
    store_row<OD>(row, hist);
    // compute_done (hover_free.py:449-461), reward (hover_free.py:206-235 / hover.py:184-202),
    // cost (hover_free.py:138-166) on the TRUE state
    bool dn = b.pos[2] < 0.2f;
    dn |= fabsf(rpy[0]) > cfg.done_rp_rad || fabsf(rpy[1]) > cfg.done_rp_rad;
#pragma unroll
    for (int i = 0; i < 3; ++i) dn |= (fabsf(wb[i]) * (180.0f / kPi)) > cfg.done_rate_deg;
    float nca = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; ++i) { float c = 0.5f * (clampf(a[i], -1.0f, 1.0f) + 1.0f); nca += c * c; }
    float p_action = cfg.pen_action * fsqrt(nca);
    float p_rpy = cfg.pen_angle * fsqrt(rpy[0] * rpy[0] + rpy[1] * rpy[1] + rpy[2] * rpy[2]);
    float p_spin = cfg.pen_spin * fsqrt(wb[0] * wb[0] + wb[1] * wb[1] + wb[2] * wb[2]);
    float p_term = dn ? cfg.pen_terminal : 0.0f;
    float p_vel = cfg.pen_velocity * fsqrt(b.vel[0] * b.vel[0] + b.vel[1] * b.vel[1] + b.vel[2] * b.vel[2]);
    float penalties = (((p_rpy + p_spin) + p_vel) + p_action) + p_term;
    float r;
    if (cfg.reward_kind == CF_REWARD_FREE) {
        r = -penalties - cfg.pen_z * fabsf(b.pos[2] - 1.0f);
    } else {
        float dx = b.pos[0], dy = b.pos[1], dz = b.pos[2] - 1.0f;
        r = -fsqrt(dx * dx + dy * dy + dz * dz) - penalties;
    }
    const float kDeg10 = 10.0f * kPi / 180.0f;
    bool c = fabsf(b.pos[0]) > 0.10f || fabsf(b.pos[1]) > 0.10f || b.pos[2] > 1.20f;
    c |= fabsf(rpy[0]) > kDeg10 || fabsf(rpy[1]) > kDeg10;
#pragma unroll
    for (int i = 0; i < 3; ++i) c |= fabsf(wb[i]) > 0.25f || fabsf(a[i]) > cfg.cost_rate_limit;   // mislabelled slices, :155-160
    StepOut o;
    o.ep_len = ep_len + 1;
    o.ep_ret = e.f(CF_S_EP_RET) + r;
    e.u(CF_S_EP_LEN, (uint32_t)o.ep_len); e.f(CF_S_EP_RET, o.ep_ret);
    o.truncated = 0;
    if (cfg.max_ep_steps > 0 && o.ep_len >= cfg.max_ep_steps) { o.truncated = dn ? 0 : 1; dn = true; }
    o.rew = r; o.cost = c ? 1.0f : 0.0f; o.done = dn ? 1 : 0;
    return o;
}

}  // namespace cf
