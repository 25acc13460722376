// cf_device.cuh - per-env device code of the batched CrazyFlie hover environment (sm_100a).
//
// One thread owns one env and keeps its whole 76-word state in registers across the control
// step: HJ disturbance lookup, 2 physics substeps (PT1 motor lag, thrust/torque, drag,
// disturbance torque, rigid-body integration), sensor-noise pipeline with gyro low-pass,
// observation history, reward, done/cost, TimeLimit and the in-place auto-reset with domain
// randomisation.  fp32 arithmetic (the reference is float64 NumPy + PyBullet); the operation
// order follows the reference functions cited at each routine (paths relative to
// phoenix_drone_simulation/ of the reference checkout).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "cf_types.h"

namespace cf {

constexpr float kPi = 3.14159265358979323846f;

__constant__ double c_grid[6][CF_HJ_PTS];   // odp Grid.grid_points, float64 linspace values
__constant__ double c_dx[6];
__constant__ double c_boltz_cdf[21];

// ---------------------------------------------------------------------------------- RNG
// Philox4x32-10, counter = (gid_lo, event counter, block, tag ^ gid_hi<<8), key = seed.
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return c;
}
__device__ __forceinline__ float u01(uint32_t w) {
    return fmaf((float)(w >> 8), 1.0f / 16777216.0f, 0.5f / 16777216.0f);
}
struct Rng {
    uint32_t k0, k1, c0, c1, c3;
    __device__ __forceinline__ uint4 raw(uint32_t block) const {
        return philox4x32_10(make_uint4(c0, c1, block, c3), k0, k1);
    }
    // four standard normals of block b (Box-Muller on two pairs, same pairing as the oracle)
    __device__ __forceinline__ float4 normals(uint32_t b) const {
        uint4 w = raw(b);
        float r0 = sqrtf(-2.0f * __logf(u01(w.x))), r1 = sqrtf(-2.0f * __logf(u01(w.z)));
        float s0, c0_, s1, c1_;
        __sincosf(kPi * (2.0f * u01(w.y) - 1.0f), &s0, &c0_);
        __sincosf(kPi * (2.0f * u01(w.w) - 1.0f), &s1, &c1_);
        return make_float4(r0 * c0_, r0 * s0, r1 * c1_, r1 * s1);
    }
    __device__ __forceinline__ float4 uniforms(uint32_t b) const {
        uint4 w = raw(16u + b);
        return make_float4(u01(w.x), u01(w.y), u01(w.z), u01(w.w));
    }
};
__device__ __forceinline__ Rng make_rng(const DevCfg& cfg, long long gid, uint32_t ctr, uint32_t tag) {
    Rng r;
    r.k0 = cfg.seed_lo; r.k1 = cfg.seed_hi;
    r.c0 = (uint32_t)gid; r.c1 = ctr;
    r.c3 = tag ^ ((uint32_t)((unsigned long long)gid >> 32) << 8);
    return r;
}

// Standard draws of one event: either generated (Philox) or read from the injected slot-major
// buffers [slot][n] (parity mode).  NZ/NU are compile-time slot counts.
template <int NZ, int NU>
struct Draws {
    float z[NZ];
    float u[NU];
    __device__ __forceinline__ void generate(const Rng& rng) {
#pragma unroll
        for (int b = 0; 4 * b < NZ; ++b) {
            float4 v = rng.normals(b);
            z[4 * b] = v.x;
            if (4 * b + 1 < NZ) z[4 * b + 1] = v.y;
            if (4 * b + 2 < NZ) z[4 * b + 2] = v.z;
            if (4 * b + 3 < NZ) z[4 * b + 3] = v.w;
        }
#pragma unroll
        for (int b = 0; 4 * b < NU; ++b) {
            float4 v = rng.uniforms(b);
            u[4 * b] = v.x;
            if (4 * b + 1 < NU) u[4 * b + 1] = v.y;
            if (4 * b + 2 < NU) u[4 * b + 2] = v.z;
            if (4 * b + 3 < NU) u[4 * b + 3] = v.w;
        }
    }
    __device__ __forceinline__ void load(const float* zbuf, const float* ubuf, int n, int i) {
#pragma unroll
        for (int s = 0; s < NZ; ++s) z[s] = zbuf[(size_t)s * n + i];
#pragma unroll
        for (int s = 0; s < NU; ++s) u[s] = ubuf[(size_t)s * n + i];
    }
};

// ---------------------------------------------------------------------------------- state
struct alignas(16) Env {
    float pos[3], ep_ret, quat[4], vel[3], ow[3], lpf[3], bias[3];
    float x[4], ou[4], ring[8], ap1[4], ap2[4], obs_prev[13];
    int ep_len; uint32_t rng_ctr, misc;
    float dt, mass, J[3], ftf1, B[4], K[4];
};
static_assert(sizeof(Env) == CF_STATE_WORDS * 4, "Env must mirror the CF_S_* record");

// misc word: level_idx | ring_idx<<8 | iteration<<12 | fresh<<31   (cf_abi.h CF_S_MISC)
__device__ __forceinline__ int misc_level(uint32_t m) { return (int)(m & 0xffu); }
__device__ __forceinline__ int misc_ring(uint32_t m) { return (int)((m >> 8) & 0xfu); }
__device__ __forceinline__ bool misc_fresh(uint32_t m) { return (m >> 31) != 0u; }
__device__ __forceinline__ uint32_t misc_pack(int level, int ring, int iter, bool fresh) {
    return (uint32_t)level | ((uint32_t)ring << 8) | ((uint32_t)(iter & 1) << 12) | (fresh ? 0x80000000u : 0u);
}

// HBM layout: float4 S[kGroups][n_envs]; group g of env i holds record words 4g..4g+3.
// Every load/store below is one 128-bit access per thread, 512 contiguous bytes per warp.
__device__ __forceinline__ void load_env(Env& e, const float4* __restrict__ S, int n, int i) {
    float4* w = reinterpret_cast<float4*>(&e);
#pragma unroll
    for (int g = 0; g < kGroups; ++g) w[g] = S[(size_t)g * n + i];
}
__device__ __forceinline__ void store_env(const Env& e, float4* __restrict__ S, int n, int i) {
    const float4* w = reinterpret_cast<const float4*>(&e);
#pragma unroll
    for (int g = 0; g < kGroups; ++g) S[(size_t)g * n + i] = w[g];
}

// ---------------------------------------------------------------------------------- kinematics
// btMatrix3x3::setRotation == pybullet.getMatrixFromQuaternion (agents.py:452, physics.py:234)
__device__ __forceinline__ void quat_to_R(const float q[4], float R[9]) {
    float x = q[0], y = q[1], z = q[2], w = q[3];
    float d = x * x + y * y + z * z + w * w, s = 2.0f / d;
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
// pybullet.getEulerFromQuaternion (agents.py:446) incl. the gimbal branch at |sin pitch| >= 0.99999
__device__ __forceinline__ void quat_to_rpy_bullet(const float q[4], float rpy[3]) {
    float x = q[0], y = q[1], z = q[2], w = q[3];
    float sqx = x * x, sqy = y * y, sqz = z * z, squ = w * w;
    float sarg = -2.0f * (x * z - w * y);
    if (sarg <= -0.99999f) { rpy[0] = 0.f; rpy[1] = -0.5f * kPi; rpy[2] = 2.f * atan2f(x, -y); }
    else if (sarg >= 0.99999f) { rpy[0] = 0.f; rpy[1] = 0.5f * kPi; rpy[2] = 2.f * atan2f(-x, y); }
    else {
        rpy[0] = atan2f(2.f * (y * z + w * x), squ - sqx - sqy + sqz);
        rpy[1] = asinf(sarg);
        rpy[2] = atan2f(2.f * (x * y + w * z), squ + sqx - sqy - sqz);
    }
}
// pybullet.getQuaternionFromEuler (hover_free.py:183,255), normalised
__device__ __forceinline__ void rpy_to_quat(const float rpy[3], float q[4]) {
    float sp, cp, st, ct, ss, cs;
    sincosf(0.5f * rpy[0], &sp, &cp);
    sincosf(0.5f * rpy[1], &st, &ct);
    sincosf(0.5f * rpy[2], &ss, &cs);
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
    out[0] = atan2f(2.0f * (w * x + y * z), 1.0f - 2.0f * (x * x + y * y));
    float t2 = 2.0f * (w * y - z * x);
    t2 = fminf(1.0f, fmaxf(-1.0f, t2));
    out[1] = asinf(t2);
    out[2] = atan2f(2.0f * (w * z + x * y), 1.0f - 2.0f * (y * y + z * z));
}
// true-state attitude and body rates as CrazyFlieAgent.update_information exposes them (agents.py:434-453)
__device__ __forceinline__ void body_rates(const Env& e, float rpy[3], float wb[3]) {
    float R[9];
    quat_to_rpy_bullet(e.quat, rpy);
    quat_to_R(e.quat, R);
    matTvec(R, e.ow, wb);
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
__device__ __forceinline__ long hj_cell(const int idx[6]) {
    return ((((long)idx[0] * 15 + idx[1]) * 15 + idx[2]) * 15 + idx[3]) * 15 * 15 + (long)idx[4] * 15 + idx[5];
}
// bit0: dV/dx3 > 0, bit1: dV/dx4 > 0
__device__ __forceinline__ int hj_signbits_value(const float* __restrict__ V, const int idx[6]) {
    long c = hj_cell(idx);
    int b0 = hj_deriv_positive(V, c, idx[3], 225, c_dx[3]) ? 1 : 0;
    int b1 = hj_deriv_positive(V, c, idx[4], 15, c_dx[4]) ? 2 : 0;
    return b0 | b1;
}
__device__ __forceinline__ int hj_signbits_packed(const uint8_t* __restrict__ P, const int idx[6]) {
    long c = hj_cell(idx);
    return (__ldg(P + (c >> 2)) >> (2 * (int)(c & 3))) & 3;
}
// distur_gener (distur_gener.py:19-183) reduced to the two torque components the physics uses
template <int HJ_MODE>
__device__ __forceinline__ void hj_disturbance(const Env& e, const DevTables& tab, int level_idx, float d[2],
                                               int* idx_out) {
    float rpy[3], R[9], wb[3];
    quat2euler_hj(e.quat, rpy);
    quat_to_R(e.quat, R);
    matTvec(R, e.ow, wb);
    int idx[6];
#pragma unroll
    for (int k = 0; k < 3; ++k) { idx[k] = hj_index_dim((double)rpy[k], k); idx[3 + k] = hj_index_dim((double)wb[k], 3 + k); }
    int bits;
    if (HJ_MODE == CF_HJ_SIGN_TABLE) bits = hj_signbits_packed(tab.signs[level_idx], idx);
    else bits = hj_signbits_value(tab.value[level_idx], idx);
    float dmax = ((float)level_idx / 10.0f) * 5.3e-3f;          // distur_gener.py:152,160
    d[0] = (bits & 1) ? -dmax : dmax;
    d[1] = (bits & 2) ? -dmax : dmax;
    if (idx_out) {
#pragma unroll
        for (int k = 0; k < 6; ++k) idx_out[k] = idx[k];
    }
}

// ---------------------------------------------------------------------------------- sensors
// SensorNoise.add_noise_to_omega (envs/sensors.py:121-134) at dt = 1/SIM_FREQ, followed by
// LowPassFilter.apply (envs/utils.py:102-105, ratio T_s/T = 0.5; base.py:107).
// z: bias-walk 3, random-walk 3, turn-on 3 standard normals.
__device__ __forceinline__ void gyro_update(Env& e, const DevCfg& cfg, const float wb[3],
                                            const float zb[3], const float zr[3], const float zt[3]) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        e.bias[i] = cfg.gyro_pi * e.bias[i] + cfg.gyro_sb * zb[i];
        float om = ((wb[i] + e.bias[i]) + cfg.gyro_rw * zr[i]) + cfg.gyro_ton * zt[i];
        e.lpf[i] = 0.5f * e.lpf[i] + 0.5f * om;
    }
}
// The 100 Hz branch of compute_observation (hover_free.py:168-200) with SensorNoise.add_noise
// (sensors.py:75-119): z = pos3, vel3, gyro9, rot3; u = pos3, rot3.  Writes obs13.
__device__ __forceinline__ void observe_full(Env& e, const DevCfg& cfg, const float rpy[3], const float wb[3],
                                             const float* z, const float* u, float obs[13]) {
    const float kPosStd = 0.002f, kPosUnif = 0.001f, kVelStd = 0.01f;     // sensors.py:18-25
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        obs[i] = e.pos[i] + (kPosStd * z[i] + (-kPosUnif + (2.0f * kPosUnif) * u[i]));
        obs[7 + i] = e.vel[i] + kVelStd * z[3 + i];                        // vel_unif_range = 0
    }
    gyro_update(e, cfg, wb, z + 6, z + 9, z + 12);
    const float lo[3] = {-kPi, -0.5f * kPi, -kPi}, hi[3] = {kPi, 0.5f * kPi, kPi};
    float rot[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float th = cfg.quat_std * z[15 + i] + (-cfg.quat_unif + (2.0f * cfg.quat_unif) * u[3 + i]);
        rot[i] = fminf(hi[i], fmaxf(lo[i], rpy[i] + th));
    }
    rpy_to_quat(rot, obs + 3);
#pragma unroll
    for (int i = 0; i < 3; ++i) obs[10 + i] = e.lpf[i];
}

// ---------------------------------------------------------------------------------- physics substep
// PybulletPhysicsWithAdversary.step_forward (envs/physics.py:213-250) with
// CrazyFlieAgent.apply_action (agents.py:259-298), PWM.act (control.py:94-100), OUNoise.noise
// (envs/utils.py:128-133) and Bullet's stepSimulation as restated by the oracle (B1-B9:
// semi-implicit Euler, exponential-map quaternion update, gyroscopic term, 0.04 damping).
__device__ __forceinline__ void physics_substep(Env& e, const DevCfg& cfg, const float a[4], const float d[2],
                                                const float zou[4]) {
    const float kFtf0 = 1.56e-5f, kArm = 0.028f, kPropZ = 0.0108f, kG = 9.81f;
    const float kDrag[3] = {9.1785e-7f, 9.1785e-7f, 10.311e-7f};
    // latency ring (agents.py:268-276): read the delayed action, then overwrite the slot
    int ridx = misc_ring(e.misc);
    float f[4], t[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float delayed = ridx ? e.ring[4 + i] : e.ring[i];
        if (ridx) e.ring[4 + i] = a[i]; else e.ring[i] = a[i];
        float c = fminf(1.0f, fmaxf(-1.0f, delayed));
        float pwm = 30000.0f + c * 30000.0f;
        e.ou[i] = e.ou[i] + (0.15f * (0.0f - e.ou[i]) + cfg.ou_sigma * zou[i]);
        float rot = sqrtf(pwm / 60000.0f);
        e.x[i] = (1.0f - e.B[i]) * e.x[i] + e.B[i] * rot;
        float noisy = (1.0f + e.ou[i]) * (e.x[i] * e.x[i]);
        float nn = fminf(1.0f, fmaxf(0.0f, noisy));
        f[i] = e.K[i] * nn;
        t[i] = e.ftf1 * f[i] + kFtf0;
    }
    e.misc ^= (1u << 8);                                    // action_idx = (action_idx + 1) % 2
    float zt = ((-t[0] + t[1]) - t[2]) + t[3];
    float R[9], wb[3], vb[3];
    quat_to_R(e.quat, R);
    matTvec(R, e.ow, wb);
    matTvec(R, e.vel, vb);
    // drag (physics.py:232-241): R @ (k * v_world) handed to Bullet as a LINK_FRAME force
    float rpm_sum = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; ++i) rpm_sum += 2.0f * kPi * ((e.x[i] * e.x[i]) * 25000.0f) / 60.0f;
    float kv[3], drag_link[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) kv[i] = (-1.0f * kDrag[i] * rpm_sum) * e.vel[i];
    matvec(R, kv, drag_link);
    float fe[4] = {f[0], f[1], f[2], f[3]};
    if (cfg.ground_effect) {                               // physics.py:27-58; reference leaves it off
        const float kPx[4] = {0.028f, -0.028f, -0.028f, 0.028f}, kPy[4] = {-0.028f, -0.028f, 0.028f, 0.028f};
        const float kGnd = 11.36859f, kPropR = 2.31348e-2f;
        float rpy[3];
        quat_to_rpy_bullet(e.quat, rpy);
        if (fabsf(rpy[0]) < 0.5f * kPi && fabsf(rpy[1]) < 0.5f * kPi) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float pz = e.pos[2] + R[6] * kPx[i] + R[7] * kPy[i] + R[8] * kPropZ;
                pz = fmaxf(pz, cfg.gnd_hclip);
                float q = kPropR / (4.0f * pz);
                fe[i] += f[i] * kGnd * (q * q);
            }
        }
    }
    // body-frame wrench: thrust at the prop positions (agents.py:311-321, urdf:59,85,111,136),
    // yaw torque (:330-337), adversary roll/pitch torques (:517-533)
    float tau[3] = {kArm * (-fe[0] - fe[1] + fe[2] + fe[3]) + d[0],
                    kArm * (-fe[0] + fe[1] + fe[2] - fe[3]) + d[1], zt};
    float fb[3] = {drag_link[0], drag_link[1], drag_link[2] + (fe[0] + fe[1] + fe[2] + fe[3])};
    float gz = -kG * e.mass;
    float gb[3] = {R[6] * gz, R[7] * gz, R[8] * gz};        // R^T (0,0,-g m)
    float Jw[3] = {e.J[0] * wb[0], e.J[1] * wb[1], e.J[2] * wb[2]};
    if (cfg.gyroscopic) {
        tau[0] -= wb[1] * Jw[2] - wb[2] * Jw[1];
        tau[1] -= wb[2] * Jw[0] - wb[0] * Jw[2];
        tau[2] -= wb[0] * Jw[1] - wb[1] * Jw[0];
    }
    float wn = sqrtf(wb[0] * wb[0] + wb[1] * wb[1] + wb[2] * wb[2]);
    float vn = sqrtf(vb[0] * vb[0] + vb[1] * vb[1] + vb[2] * vb[2]);
    float ka = cfg.ang_damp, kl = cfg.lin_damp;
    float alpha_b[3], acc_b[3], alpha_w[3], acc_w[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        alpha_b[i] = (tau[i] - Jw[i] * (ka + ka * wn)) / e.J[i];
        acc_b[i] = ((fb[i] + gb[i]) - e.mass * vb[i] * (kl + kl * vn)) / e.mass;
    }
    matvec(R, alpha_b, alpha_w);
    matvec(R, acc_b, acc_w);
    float dt = e.dt;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        e.ow[i] += dt * alpha_w[i];
        e.vel[i] += dt * acc_w[i];
        e.pos[i] += dt * e.vel[i];
    }
    // exponential-map quaternion update (btMultiBody::stepPositionsMultiDof)
    float w0 = e.ow[0], w1 = e.ow[1], w2 = e.ow[2];
    float ang = sqrtf(w0 * w0 + w1 * w1 + w2 * w2), s, cw;
    if (ang * dt > 0.25f * kPi) ang = 0.25f * kPi / dt;
    if (ang < 0.001f) { s = 0.5f * dt - (dt * dt * dt) * 0.020833333333f * ang * ang; cw = cosf(ang * dt * 0.5f); }
    else { float sn; sincosf(0.5f * ang * dt, &sn, &cw); s = sn / ang; }
    float dq[4] = {w0 * s, w1 * s, w2 * s, cw}, q[4];
    const float* b = e.quat;
    q[0] = dq[3] * b[0] + dq[0] * b[3] + dq[1] * b[2] - dq[2] * b[1];
    q[1] = dq[3] * b[1] - dq[0] * b[2] + dq[1] * b[3] + dq[2] * b[0];
    q[2] = dq[3] * b[2] + dq[0] * b[1] - dq[1] * b[0] + dq[2] * b[3];
    q[3] = dq[3] * b[3] - dq[0] * b[0] - dq[1] * b[1] - dq[2] * b[2];
    float inv = rsqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
#pragma unroll
    for (int i = 0; i < 4; ++i) e.quat[i] = q[i] * inv;
}

// ---------------------------------------------------------------------------------- reset
__device__ __forceinline__ float unif(float lo, float hi, float u) { return lo + (hi - lo) * u; }
__device__ __forceinline__ float dr_draw(float def, float f, float u) { float b = f * def; return unif(def - b, def + b, u); }

// Constructor state of an env (DroneBaseEnv.__init__ + CrazyFlieAgent.__init__): zero state at
// (0,0,1), nominal parameters; `fresh` marks that the constructor-time compute_observation()
// (base.py:141), which advances gyro_bias once, has not been accounted for yet.
__device__ __forceinline__ void init_env(Env& e, const DevCfg& cfg) {
    float* w = reinterpret_cast<float*>(&e);
#pragma unroll
    for (int k = 0; k < CF_STATE_WORDS; ++k) w[k] = 0.0f;
    e.pos[2] = 1.0f; e.quat[3] = 1.0f;
    e.dt = cfg.sim_dt; e.mass = 0.030f;
    e.J[0] = 1.33e-5f; e.J[1] = 1.33e-5f; e.J[2] = 2.64e-5f;
    e.ftf1 = 5.96e-3f;
#pragma unroll
    for (int i = 0; i < 4; ++i) { e.B[i] = cfg.nomB; e.K[i] = cfg.nomK; }
    e.ep_len = 0; e.rng_ctr = 0u;
    e.misc = misc_pack(cfg.fixed_level_idx, 0, 0, true);
}

// reset(): envs/base.py:420-464, task_specific_reset hover_free.py:237-289 / hover.py:1463-1513,
// apply_domain_randomization base.py:241-298 (+ agents.py:208-224), RandomHJ level draw
// hover_free.py:536.  Draw slots: z[0:4] motor x, z[4:12] latency ring, z[12:30] and z[30:48]
// the two compute_observation calls; u[0:14] initial state, u[14:29] DR, u[29:35], u[35:41]
// observation uniforms, u[41] Boltzmann.  obs: obs_dim floats.
template <bool NOISE>
__device__ __forceinline__ void reset_env(Env& e, const DevCfg& cfg, const Draws<CF_RESET_NORMALS, CF_RESET_UNIFORMS>& dw,
                                          float* obs) {
    const float* nz = dw.z;
    const float* un = dw.u;
    // stale body rate of the previous state, read before anything moves (base.py:444 before :448)
    float stale_rpy[3], stale_wb[3];
    body_rates(e, stale_rpy, stale_wb);
    int level = misc_level(e.misc);
    e.ep_len = 0; e.ep_ret = 0.0f;
    // drone.reset(), agents.py:377-386
#pragma unroll
    for (int i = 0; i < 4; ++i) e.x[i] = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) e.ring[i] = 0.0f;
    float pos[3] = {0.f, 0.f, 1.f}, rpy[3] = {0.f, 0.f, 0.f}, vel[3] = {0.f, 0.f, 0.f}, rate[3] = {0.f, 0.f, 0.f};
    if (cfg.reset_dist) {
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            pos[i] += unif(-cfg.init_pos, cfg.init_pos, un[i]);
            rpy[i] = unif(-cfg.init_angle, cfg.init_angle, un[3 + i]);
            vel[i] = unif(-cfg.init_vel, cfg.init_vel, un[7 + i]);
            rate[i] = unif(-cfg.init_rate, cfg.init_rate, un[10 + i]);
        }
        rpy[2] = unif(-cfg.init_yaw, cfg.init_yaw, un[6]);
        rate[2] = unif(-cfg.init_yaw_rate, cfg.init_yaw_rate, un[13]);
#pragma unroll
        for (int i = 0; i < 4; ++i) e.x[i] = cfg.hover_x + 0.02f * nz[i];
#pragma unroll
        for (int i = 0; i < 8; ++i) e.ring[i] = fminf(1.0f, fmaxf(-1.0f, cfg.hover_a + 0.02f * nz[4 + i]));
        rpy_to_quat(rpy, e.quat);
    } else {
        e.quat[0] = e.quat[1] = e.quat[2] = 0.0f; e.quat[3] = 1.0f;
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) { e.pos[i] = pos[i]; e.vel[i] = vel[i]; }
    {   // resetBaseVelocity(angularVelocity = R.T @ rpy_dot), hover_free.py:283-289
        float R[9];
        quat_to_R(e.quat, R);
        matTvec(R, rate, e.ow);
    }
    if (cfg.dr_on) {
        float f = cfg.dr;
        e.dt = dr_draw(cfg.sim_dt, f, un[14]);
        e.mass = dr_draw(0.030f, f, un[15]);
        e.J[0] = dr_draw(1.33e-5f, f, un[16]);
        e.J[1] = dr_draw(1.33e-5f, f, un[17]);
        e.J[2] = dr_draw(2.64e-5f, f, un[18]);
        e.ftf1 = dr_draw(5.96e-3f, f, un[20]);          // un[19]: force_torque_factor_0, cancels in the yaw torque
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float mtc = dr_draw(cfg.motor_tc, f, un[21 + i]);
            float t2w = dr_draw(1.8f, f, un[25 + i]);
            float T = fmaxf(mtc, e.dt);
            e.B[i] = e.dt / T;
            e.K[i] = 0.028f * 9.81f * t2w / 4.0f;       // agents.py:224 (0.028 kg, not the URDF mass)
        }
    }
    // gyro_lpf.set(stale rpy_dot), base.py:444
#pragma unroll
    for (int i = 0; i < 3; ++i) e.lpf[i] = stale_wb[i];
    // two compute_observation() calls: base.py:453 and, via compute_history, :308
    const float* la = e.ring + 4;                           // view of action_buffer[-1,:], agents.py:386
    float rpy_t[3], wb[3];
    body_rates(e, rpy_t, wb);
    if (NOISE) {
        float ob[13];
        observe_full(e, cfg, rpy_t, wb, nz + 12, un + 29, ob);
#pragma unroll
        for (int i = 0; i < 13; ++i) obs[i] = ob[i];
        observe_full(e, cfg, rpy_t, wb, nz + 30, un + 35, ob);
#pragma unroll
        for (int i = 0; i < 13; ++i) { obs[17 + i] = ob[i]; e.obs_prev[i] = ob[i]; }
#pragma unroll
        for (int i = 0; i < 4; ++i) { obs[13 + i] = la[i]; obs[30 + i] = la[i]; }
    } else {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float* o = obs + 21 * h;
#pragma unroll
            for (int i = 0; i < 3; ++i) { o[i] = e.pos[i]; o[7 + i] = e.vel[i]; o[10 + i] = wb[i]; }
#pragma unroll
            for (int i = 0; i < 4; ++i) { o[3 + i] = e.quat[i]; o[13 + i] = la[i]; o[17 + i] = la[i]; }
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) { e.obs_prev[i] = e.pos[i]; e.obs_prev[7 + i] = e.vel[i]; e.obs_prev[10 + i] = wb[i]; }
#pragma unroll
        for (int i = 0; i < 4; ++i) e.obs_prev[3 + i] = e.quat[i];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) { e.ap1[i] = la[i]; e.ap2[i] = la[i]; }
    // Boltzmann() at the end of reset, hover_free.py:536: cdf.searchsorted(u, side='right')
    if (cfg.dstb_kind == CF_DSTB_HJ_BOLTZMANN) {
        int k = 0;
        double u = (double)un[41];
        while (k < 20 && c_boltz_cdf[k] <= u) ++k;
        level = k;
    }
    e.misc = misc_pack(level, 0, 0, false);
}

// ---------------------------------------------------------------------------------- step
struct StepOut {
    float rew, cost;
    int done, truncated;
};

// step(): hover_free.py:391-444 / 608-665 / 778-833 (hover.py twins) under gym TimeLimit.
// Draw slots: z[0:4] OU substep 1, z[4:13] gyro after substep 1, z[13:17] OU substep 2,
// z[17:26] gyro after substep 2, z[26:44] the full compute_observation of compute_history;
// u[0:6] its position / attitude uniforms, u[6:8] the random-box disturbance.
// hist: obs_dim floats = [o(k-1), u(k-2), o(k), u(k-1)] (base.py:305-321).  No reset here.
template <bool NOISE, int HJ_MODE>
__device__ __forceinline__ StepOut step_env(Env& e, const DevCfg& cfg, const DevTables& tab, const float a[4],
                                            const Draws<CF_STEP_NORMALS, CF_STEP_UNIFORMS>& dw, float* hist) {
    const float* nz = dw.z;
    const float* un = dw.u;
    e.rng_ctr += 1u;
    // disturbance, once per control step
    float d[2] = {0.0f, 0.0f};
    if (cfg.dstb_kind == CF_DSTB_HJ_FIXED || cfg.dstb_kind == CF_DSTB_HJ_BOLTZMANN) {
        int level = misc_level(e.misc);
        if (level > 0) hj_disturbance<HJ_MODE>(e, tab, level, d, nullptr);
    } else if (cfg.dstb_kind == CF_DSTB_RANDOM_BOX) {       // dstb_space.sample(), hover_free.py:868-878
        d[0] = unif(-1e-3f, 1e-3f, un[6]);
        d[1] = unif(-1e-3f, 1e-3f, un[7]);
    }
    float rpy[3], wb[3];
    // substeps; the reference calls compute_observation() after each (base.py:494-501): only its
    // gyro sample survives (bias walk + low-pass), the rest of that call is overwritten below
    physics_substep(e, cfg, a, d, nz + 0);
    if (NOISE) { body_rates(e, rpy, wb); gyro_update(e, cfg, wb, nz + 4, nz + 7, nz + 10); }
    physics_substep(e, cfg, a, d, nz + 13);
    body_rates(e, rpy, wb);
    if (NOISE) gyro_update(e, cfg, wb, nz + 17, nz + 20, nz + 23);
    // compute_history, base.py:305-321
    // action history with the reference's aliased view of action_buffer[-1,:] in the first two
    // steps of an episode (agents.py:386, base.py:459-460): [a,a] then [a,u(k-1)] then normal
    const bool first = e.ep_len == 0, second = e.ep_len == 1;
    float h0[4], h1[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        h0[i] = (first || second) ? e.ring[4 + i] : e.ap2[i];
        h1[i] = first ? e.ring[4 + i] : e.ap1[i];
    }
    if (NOISE) {
        float on[13];
        observe_full(e, cfg, rpy, wb, nz + 26, un, on);
#pragma unroll
        for (int i = 0; i < 13; ++i) { hist[i] = e.obs_prev[i]; hist[17 + i] = on[i]; e.obs_prev[i] = on[i]; }
#pragma unroll
        for (int i = 0; i < 4; ++i) { hist[13 + i] = h0[i]; hist[30 + i] = h1[i]; }
    } else {
#pragma unroll
        for (int i = 0; i < 13; ++i) hist[i] = e.obs_prev[i];
#pragma unroll
        for (int i = 0; i < 4; ++i) { hist[13 + i] = e.ap1[i]; hist[17 + i] = h0[i]; hist[38 + i] = h1[i]; hist[34 + i] = a[i]; }
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            hist[21 + i] = e.pos[i]; hist[28 + i] = e.vel[i]; hist[31 + i] = wb[i];
            e.obs_prev[i] = e.pos[i]; e.obs_prev[7 + i] = e.vel[i]; e.obs_prev[10 + i] = wb[i];
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) { hist[24 + i] = e.quat[i]; e.obs_prev[3 + i] = e.quat[i]; }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) { e.ap2[i] = e.ap1[i]; e.ap1[i] = a[i]; }
    // compute_done (hover_free.py:449-461), reward (hover_free.py:206-235 / hover.py:184-202),
    // cost (hover_free.py:138-166) on the TRUE state
    bool dn = e.pos[2] < 0.2f;
    dn |= fabsf(rpy[0]) > cfg.done_rp_rad || fabsf(rpy[1]) > cfg.done_rp_rad;
#pragma unroll
    for (int i = 0; i < 3; ++i) dn |= (180.0f * fabsf(wb[i]) / kPi) > cfg.done_rate_deg;
    float nca = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; ++i) { float c = 0.5f * (fminf(1.0f, fmaxf(-1.0f, a[i])) + 1.0f); nca += c * c; }
    float p_action = cfg.pen_action * sqrtf(nca);
    float p_rpy = cfg.pen_angle * sqrtf(rpy[0] * rpy[0] + rpy[1] * rpy[1] + rpy[2] * rpy[2]);
    float p_spin = cfg.pen_spin * sqrtf(wb[0] * wb[0] + wb[1] * wb[1] + wb[2] * wb[2]);
    float p_term = dn ? cfg.pen_terminal : 0.0f;
    float p_vel = cfg.pen_velocity * sqrtf(e.vel[0] * e.vel[0] + e.vel[1] * e.vel[1] + e.vel[2] * e.vel[2]);
    float penalties = (((p_rpy + p_spin) + p_vel) + p_action) + p_term;
    float r;
    if (cfg.reward_kind == CF_REWARD_FREE) {
        r = -penalties - cfg.pen_z * fabsf(e.pos[2] - 1.0f);
    } else {
        float dx = e.pos[0], dy = e.pos[1], dz = e.pos[2] - 1.0f;
        r = -sqrtf(dx * dx + dy * dy + dz * dz) - penalties;
    }
    const float kDeg10 = 10.0f * kPi / 180.0f;
    bool c = fabsf(e.pos[0]) > 0.10f || fabsf(e.pos[1]) > 0.10f || e.pos[2] > 1.20f;
    c |= fabsf(rpy[0]) > kDeg10 || fabsf(rpy[1]) > kDeg10;
#pragma unroll
    for (int i = 0; i < 3; ++i) c |= fabsf(wb[i]) > 0.25f || fabsf(a[i]) > cfg.cost_rate_limit;   // mislabelled slices, :155-160
    e.ep_len += 1; e.ep_ret += r;
    StepOut o;
    o.truncated = 0;
    if (cfg.max_ep_steps > 0 && e.ep_len >= cfg.max_ep_steps) { o.truncated = dn ? 0 : 1; dn = true; }
    o.rew = r; o.cost = c ? 1.0f : 0.0f; o.done = dn ? 1 : 0;
    return o;
}

}  // namespace cf
