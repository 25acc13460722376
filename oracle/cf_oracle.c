/* cf_oracle.c - oracle-B: float64 CPU restatement of the reference hover-env step.
 *
 * TEST INFRASTRUCTURE ONLY.  Imported/linked only by tests/, __graft_entry__.smoke()
 * and bench.py's CPU-baseline legs, and there only as the checker or the timed CPU
 * baseline - never by the product package.
 *
 * PARITY STATUS: the Python side of the reference (everything except PyBullet's
 * stepSimulation) is pinned: this file is checked against the reference's own Python
 * executed here on oracle/fake_bullet.py (oracle/ref_harness.py, fixtures under
 * tests/golden/ made by oracle/make_golden.py).  The rigid-body integrator itself
 * follows Bullet's btMultiBody as restated in fake_bullet.py (switches B1-B9) and is
 * PARITY UNPINNED against real PyBullet 3.21, which cannot be installed here.
 *
 * Each function cites the reference file:line it follows (paths relative to
 * /root/reference/phoenix_drone_simulation/).
 *
 * Build: see oracle/Makefile  (gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/cf_abi.h"

#ifdef _OPENMP
#include <omp.h>
#endif

#define PI 3.141592653589793

/* ---- constants: envs/assets/cf21x_bullet.urdf:12,21,31 and envs/agents.py:141-156 ---- */
static const double URDF_M = 0.030;
static const double URDF_J[3] = {1.33e-5, 1.33e-5, 2.64e-5};
static const double T2W = 1.8;
static const double ARM_XY = 0.028;            /* prop joint origins, urdf:59,85,111,136 */
static const double PROP_Z = 0.0108;
static const double DRAG[3] = {9.1785e-7, 9.1785e-7, 10.311e-7};
static const double GND_EFF_COEFF = 11.36859, PROP_RADIUS = 2.31348e-2, URDF_KF = 3.16e-10;
static const double FTF0 = 1.56e-5, FTF1 = 5.96e-3;   /* agents.py:142-143 */
static const double G_ACC = 9.81;
static const double OU_THETA = 0.15;                  /* envs/utils.py:111 */
static const double PROP_X[4] = {0.028, -0.028, -0.028, 0.028};
static const double PROP_Y[4] = {-0.028, -0.028, 0.028, 0.028};

/* SensorNoise defaults, envs/sensors.py:18-33 */
static const double POS_STD = 0.002, POS_UNIF = 0.001, VEL_STD = 0.01, VEL_UNIF = 0.0;
static const double GYRO_NOISE_DENSITY = 0.000175, GYRO_RW = 0.0105, GYRO_CORR_T = 1000.0;

typedef struct cfo_env {
    /* rigid body (what Bullet holds) */
    double pos[3], quat[4], vel[3], omega_w[3];
    /* CrazyFlieAgent internals, agents.py:170-206 */
    double x[4], ring[2][4], ou[4];
    int ring_idx;
    /* SensorNoise.gyro_bias (sensors.py:68), DroneBaseEnv.gyro_lpf (base.py:107) */
    double bias[3], lpf[3];
    /* cached noisy 100 Hz state self.state[0:10] (hover_free.py:184-189) */
    double st_pos[3], st_quat[4], st_vel[3];
    /* history deques (base.py:134-135) */
    double obs_prev[17], act_prev1[4], act_prev2[4];
    double ep_ret;
    int ep_len, iteration, level_idx, fresh;
    uint32_t rng_ctr;
    /* domain-randomised parameters (base.py:241-298) */
    double dt, mass, J[3], ftf1, A[4], B[4], K[4];
    /* identity */
    uint64_t seed;
    int64_t gid;
} cfo_env;

static const float* g_tables[CF_MAX_LEVELS];
static double g_grid[6][CF_HJ_PTS];
static double g_dx[6];
static double g_boltz_cdf[21];
static int g_static_ready = 0;

/* ------------------------------------------------------------------ Philox4x32-10 */
static void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
void cfo_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) { philox4x32_10(ctr, key, out); }

/* 23 random bits -> (2k+1)/2^24, the value the CUDA side builds with the mantissa trick */
static double u01(uint32_t w) { return ((double)(w >> 9) * 2.0 + 1.0) * (1.0 / 16777216.0); }

/* standard draws of one RNG event: normals block b -> z[4b..4b+3]; uniform block b -> counter 16+b */
static void draw_event(const cfo_env* e, uint32_t ctr1, uint32_t tag, int n_norm, int n_unif,
                       double* nz, double* un) {
    uint32_t key[2] = {(uint32_t)e->seed, (uint32_t)(e->seed >> 32)};
    uint32_t ctr[4] = {(uint32_t)e->gid, ctr1, 0, tag ^ ((uint32_t)((uint64_t)e->gid >> 32) << 8)};
    uint32_t w[4];
    for (int b = 0; 4 * b < n_norm; ++b) {
        ctr[2] = (uint32_t)b;
        philox4x32_10(ctr, key, w);
        for (int h = 0; h < 2; ++h) {
            double r = sqrt(-2.0 * log(u01(w[2 * h])));
            double th = PI * (2.0 * u01(w[2 * h + 1]) - 1.0);
            if (4 * b + 2 * h < n_norm) nz[4 * b + 2 * h] = r * cos(th);
            if (4 * b + 2 * h + 1 < n_norm) nz[4 * b + 2 * h + 1] = r * sin(th);
        }
    }
    for (int b = 0; 4 * b < n_unif; ++b) {
        ctr[2] = 16u + (uint32_t)b;
        philox4x32_10(ctr, key, w);
        for (int j = 0; j < 4; ++j)
            if (4 * b + j < n_unif) un[4 * b + j] = u01(w[j]);
    }
}
void cfo_draw_event(const cfo_env* e, uint32_t ctr1, uint32_t tag, int n_norm, int n_unif,
                    double* nz, double* un) { draw_event(e, ctr1, tag, n_norm, n_unif, nz, un); }

/* ------------------------------------------------------------------ static tables */
static void init_static(void) {
    if (g_static_ready) return;
    /* odp/Grid/GridProcessing.py:6-51 with distur_gener.py:179 arguments; np.linspace:
     * step = (hi-lo)/14, y = lo + i*step, last point forced to hi. */
    for (int d = 0; d < 6; ++d) {
        double lo = d < 3 ? -PI / 2.4 : -PI, hi = d < 3 ? PI / 2.4 : PI;
        if (d < 3) hi = lo + (hi - lo) * (1.0 - 1.0 / 15.0);
        g_dx[d] = (hi - lo) / (15.0 - 1.0);
        double step = (hi - lo) / 14.0;
        for (int i = 0; i < 15; ++i) g_grid[d][i] = lo + (double)i * step;
        g_grid[d][14] = hi;
    }
    /* Boltzmann(), envs/utils.py:27-39 + legacy np.random.choice(p=) */
    double w[21], s = 0.0, c = 0.0;
    for (int i = 0; i < 21; ++i) { w[i] = exp(-1.0 * (0.0 + (double)i * 0.1)); }
    /* np.sum uses pairwise summation; for n < 128 with 8-way unrolling: reproduce below */
    {
        double r[8];
        for (int j = 0; j < 8; ++j) r[j] = w[j];
        int i;
        for (i = 8; i < 21 - (21 % 8); i += 8)
            for (int j = 0; j < 8; ++j) r[j] += w[i + j];
        s = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < 21; ++i) s += w[i];
    }
    for (int i = 0; i < 21; ++i) { c += w[i] / s; g_boltz_cdf[i] = c; }
    double last = g_boltz_cdf[20];
    for (int i = 0; i < 21; ++i) g_boltz_cdf[i] /= last;
    g_static_ready = 1;
}

void cfo_grid_points(double* out /* [6][15] */) { init_static(); memcpy(out, g_grid, sizeof(g_grid)); }
void cfo_boltzmann_cdf(double* out /* [21] */) { init_static(); memcpy(out, g_boltz_cdf, sizeof(g_boltz_cdf)); }
void cfo_set_table(int level_idx, const float* V) { if (level_idx >= 0 && level_idx < CF_MAX_LEVELS) g_tables[level_idx] = V; }

/* ------------------------------------------------------------------ kinematics (Bullet surface) */
/* btMatrix3x3::setRotation == pybullet.getMatrixFromQuaternion (agents.py:452, physics.py:234) */
static void quat_to_R(const double q[4], double R[9]) {
    double x = q[0], y = q[1], z = q[2], w = q[3];
    double d = x * x + y * y + z * z + w * w, s = 2.0 / d;
    double xs = x * s, ys = y * s, zs = z * s;
    double wx = w * xs, wy = w * ys, wz = w * zs, xx = x * xs, xy = x * ys, xz = x * zs;
    double yy = y * ys, yz = y * zs, zz = z * zs;
    R[0] = 1.0 - (yy + zz); R[1] = xy - wz; R[2] = xz + wy;
    R[3] = xy + wz; R[4] = 1.0 - (xx + zz); R[5] = yz - wx;
    R[6] = xz - wy; R[7] = yz + wx; R[8] = 1.0 - (xx + yy);
}
static void matvec(const double R[9], const double v[3], double o[3]) {
    o[0] = R[0] * v[0] + R[1] * v[1] + R[2] * v[2];
    o[1] = R[3] * v[0] + R[4] * v[1] + R[5] * v[2];
    o[2] = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
}
static void matTvec(const double R[9], const double v[3], double o[3]) {
    o[0] = R[0] * v[0] + R[3] * v[1] + R[6] * v[2];
    o[1] = R[1] * v[0] + R[4] * v[1] + R[7] * v[2];
    o[2] = R[2] * v[0] + R[5] * v[1] + R[8] * v[2];
}
/* pybullet.getEulerFromQuaternion (agents.py:446), switch B9 */
static void quat_to_rpy_bullet(const double q[4], double rpy[3]) {
    double x = q[0], y = q[1], z = q[2], w = q[3];
    double sqx = x * x, sqy = y * y, sqz = z * z, squ = w * w;
    double sarg = -2.0 * (x * z - w * y);
    if (sarg <= -0.99999) { rpy[0] = 0; rpy[1] = -0.5 * PI; rpy[2] = 2 * atan2(x, -y); }
    else if (sarg >= 0.99999) { rpy[0] = 0; rpy[1] = 0.5 * PI; rpy[2] = 2 * atan2(-x, y); }
    else {
        rpy[0] = atan2(2 * (y * z + w * x), squ - sqx - sqy + sqz);
        rpy[1] = asin(sarg);
        rpy[2] = atan2(2 * (x * y + w * z), squ + sqx - sqy - sqz);
    }
}
/* pybullet.getQuaternionFromEuler (hover_free.py:183,255), normalised */
static void rpy_to_quat(const double rpy[3], double q[4]) {
    double phi = rpy[0] / 2.0, the = rpy[1] / 2.0, psi = rpy[2] / 2.0;
    q[0] = sin(phi) * cos(the) * cos(psi) - cos(phi) * sin(the) * sin(psi);
    q[1] = cos(phi) * sin(the) * cos(psi) + sin(phi) * cos(the) * sin(psi);
    q[2] = cos(phi) * cos(the) * sin(psi) - sin(phi) * sin(the) * cos(psi);
    q[3] = cos(phi) * cos(the) * cos(psi) + sin(phi) * sin(the) * sin(psi);
    double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    for (int i = 0; i < 4; ++i) q[i] /= n;
}
/* distur_gener.py:186-207 quat2euler */
static void quat2euler_hj(const double q[4], double out[3]) {
    double x = q[0], y = q[1], z = q[2], w = q[3];
    double t0 = 2.0 * (w * x + y * z), t1 = 1.0 - 2.0 * (x * x + y * y);
    out[0] = atan2(t0, t1);
    double t2 = 2.0 * (w * y - z * x);
    t2 = t2 > 1.0 ? 1.0 : t2; t2 = t2 < -1.0 ? -1.0 : t2;
    out[1] = asin(t2);
    double t3 = 2.0 * (w * z + x * y), t4 = 1.0 - 2.0 * (y * y + z * z);
    out[2] = atan2(t3, t4);
}

/* ------------------------------------------------------------------ HJ lookup */
/* Grid.get_index, odp/Grid/GridProcessing.py:53-72 */
void cfo_hj_index(const double s[6], int idx[6]) {
    init_static();
    for (int d = 0; d < 6; ++d) {
        const double* g = g_grid[d];
        int i = 0;
        if (s[d] != s[d]) i = 15;                               /* np.searchsorted sorts NaN last */
        while (i < 15 && g[i] < s[d]) ++i;                      /* np.searchsorted, side='left' */
        if (i > 0 && (i == 15 || fabs(s[d] - g[i - 1]) < fabs(s[d] - g[i]))) i -= 1;
        idx[d] = i;
    }
}
static float sgnf(float v) { return v > 0.f ? 1.f : (v < 0.f ? -1.f : 0.f); }
/* spa_deriv for one non-periodic dim (distur_gener.py:40-99) followed by the `> 0` test of
 * opt_dstb_non_hcl (:106-125); float32 arithmetic exactly where numpy keeps float32. */
static int deriv_positive(const float* V, const int idx[6], int dim) {
    static const long stride[6] = {759375, 50625, 3375, 225, 15, 1};
    long c = 0;
    for (int d = 0; d < 6; ++d) c += stride[d] * idx[d];
    int i = idx[dim];
    float vc = V[c];
    volatile float a, b;   /* volatile: force float32 rounding of every intermediate */
    if (i == 0) {
        float vn = V[c + stride[dim]];
        volatile float diff = vn - vc;
        volatile float ext = fabsf(diff) * sgnf(vc);
        volatile float lb = vc + ext;
        a = vc - lb; b = vn - vc;
    } else if (i == 14) {
        float vp = V[c - stride[dim]];
        volatile float diff = vc - vp;
        volatile float ext = fabsf(diff) * sgnf(vc);
        volatile float rb = vc + ext;
        a = vc - vp; b = rb - vc;
    } else {
        a = vc - V[c - stride[dim]]; b = V[c + stride[dim]] - vc;
    }
    double left = (double)a / g_dx[dim], right = (double)b / g_dx[dim];
    return (left + right) / 2.0 > 0.0;
}
/* bit0: dV/dx3 > 0 (roll-torque disturbance = -dmax), bit1: dV/dx4 > 0 */
int cfo_hj_signbits(const float* V, const int idx[6]) {
    init_static();
    return deriv_positive(V, idx, 3) | (deriv_positive(V, idx, 4) << 1);
}
/* distur_gener (distur_gener.py:19-183) reduced to the two components the physics uses */
static void hj_disturbance(const cfo_env* e, double level, double d[2]) {
    double rpy[3], R[9], wb[3], s[6];
    int idx[6];
    quat2euler_hj(e->quat, rpy);
    quat_to_R(e->quat, R);
    matTvec(R, e->omega_w, wb);
    s[0] = rpy[0]; s[1] = rpy[1]; s[2] = rpy[2]; s[3] = wb[0]; s[4] = wb[1]; s[5] = wb[2];
    cfo_hj_index(s, idx);
    int li = (int)lround(level * 10.0);
    const float* V = g_tables[li];
    int bits = V ? cfo_hj_signbits(V, idx) : 0;
    double dmax0 = level * 5.3e-3, dmax1 = level * 5.3e-3;      /* distur_gener.py:152,160 */
    d[0] = (bits & 1) ? -dmax0 : dmax0;
    d[1] = (bits & 2) ? -dmax1 : dmax1;
}

/* ------------------------------------------------------------------ sensors */
static double deg2rad(double x) { return PI * x / 180.0; }      /* envs/utils.py:47-49 */
static double rad2deg(double x) { return 180.0 * x / PI; }      /* envs/utils.py:42-44 */

/* SensorNoise.add_noise_to_omega, envs/sensors.py:121-134 (dt = 1/SIM_FREQ, hover_free.py:181,192) */
static void noisy_omega(cfo_env* e, const double wb[3], const double z[9], double out[3]) {
    double dt = 1.0 / 200.0;
    double sigma_g_d = GYRO_NOISE_DENSITY / pow(dt, 0.5);
    double sigma_b_g_d = pow(-(sigma_g_d * sigma_g_d) * (GYRO_CORR_T / 2.0) * (exp(-2.0 * dt / GYRO_CORR_T) - 1.0), 0.5);
    double pi_g_d = exp(-dt / GYRO_CORR_T);
    double turn_on = deg2rad(5.0);
    for (int i = 0; i < 3; ++i) {
        e->bias[i] = pi_g_d * e->bias[i] + sigma_b_g_d * z[i];
        out[i] = ((wb[i] + e->bias[i]) + GYRO_RW * z[3 + i]) + turn_on * z[6 + i];
    }
}
/* LowPassFilter.apply, envs/utils.py:102-105; gyro_lpf = LowPassFilter(1, 2/sim_freq, 1/sim_freq), base.py:107 */
static void lpf_apply(cfo_env* e, const double u[3], double out[3]) {
    double ratio = (1.0 / 200.0) / (2.0 / 200.0);
    for (int i = 0; i < 3; ++i) { e->lpf[i] = (1.0 - ratio) * e->lpf[i] + 1.0 * ratio * u[i]; out[i] = e->lpf[i]; }
}
static void body_rates(const cfo_env* e, double rpy[3], double wb[3]) {   /* agents.py:434-453 */
    double R[9];
    quat_to_rpy_bullet(e->quat, rpy);
    quat_to_R(e->quat, R);
    matTvec(R, e->omega_w, wb);
}
/* compute_observation, hover_free.py:168-200.  full: z[18] = pos3, vel3, gyro9, rot3; u[6] = pos3, rot3.
 * gyro-only: z[9].  Noise off: obs17 = drone.get_state() (agents.py:339-348). Returns obs length. */
static int compute_observation(cfo_env* e, const cf_config* cfg, const double last_action[4],
                               const double* z, const double* u, double* obs) {
    double rpy[3], wb[3];
    body_rates(e, rpy, wb);
    if (cfg->observation_noise <= 0) {
        memcpy(obs, e->pos, 24); memcpy(obs + 3, e->quat, 32); memcpy(obs + 7, e->vel, 24);
        memcpy(obs + 10, wb, 24); memcpy(obs + 13, last_action, 32);
        return 17;
    }
    double om[3];
    if (e->iteration % 2 == 0) {           /* obs_rate = 200 // 100, base.py:106 */
        double quat_std = deg2rad(0.1), quat_unif = deg2rad(0.05), rot[3];
        static const double lo[3] = {-PI, -PI / 2, -PI}, hi[3] = {PI, PI / 2, PI};
        for (int i = 0; i < 3; ++i) {       /* sensors.py:84-95 */
            double off = (0.0 + POS_STD * z[i]) + (-POS_UNIF + (POS_UNIF - (-POS_UNIF)) * u[i]);
            e->st_pos[i] = e->pos[i] + off;
            e->st_vel[i] = (e->vel[i] + (0.0 + VEL_STD * z[3 + i])) + (-VEL_UNIF + (VEL_UNIF - (-VEL_UNIF)) * 0.5);
        }
        noisy_omega(e, wb, z + 6, om);
        for (int i = 0; i < 3; ++i) {       /* sensors.py:101-109 */
            double th = (0.0 + quat_std * z[15 + i]) + (-quat_unif + (quat_unif - (-quat_unif)) * u[3 + i]);
            double r = rpy[i] + th;
            rot[i] = r < lo[i] ? lo[i] : (r > hi[i] ? hi[i] : r);
        }
        rpy_to_quat(rot, e->st_quat);
    } else {
        noisy_omega(e, wb, z, om);
    }
    double of[3];
    lpf_apply(e, om, of);
    memcpy(obs, e->st_pos, 24); memcpy(obs + 3, e->st_quat, 32); memcpy(obs + 7, e->st_vel, 24);
    memcpy(obs + 10, of, 24);
    return 13;
}
int cfo_obs_dim(const cf_config* cfg) { return cfg->observation_noise > 0 ? 34 : 42; }

/* ------------------------------------------------------------------ reset */
static void nominal_params(cfo_env* e, const cf_config* cfg) {   /* agents.py:145-204 */
    e->dt = cfg->sim_time_step; e->mass = URDF_M;
    for (int i = 0; i < 3; ++i) e->J[i] = URDF_J[i];
    e->ftf1 = FTF1;
    for (int i = 0; i < 4; ++i) {
        e->A[i] = 1.0 * (1.0 - cfg->sim_time_step / cfg->motor_time_constant);
        e->B[i] = 1.0 * cfg->sim_time_step / cfg->motor_time_constant;
        e->K[i] = (G_ACC * URDF_M) * T2W / 4.0;
    }
}
static double unif(double lo, double hi, double u) { return lo + (hi - lo) * u; }
static double dr_draw(double def, double f, double u) { double b = f * def; return unif(def - b, def + b, u); }

/* constructor side effects: zero state, nominal parameters, and the one compute_observation()
 * call of DroneBaseEnv.__init__ (base.py:141) that advances gyro_bias once */
void cfo_init(cfo_env* e, const cf_config* cfg, uint64_t seed, int64_t gid) {
    init_static();
    memset(e, 0, sizeof(*e));
    e->seed = seed; e->gid = gid;
    e->pos[2] = 1.0; e->quat[3] = 1.0; e->st_quat[3] = 1.0;
    nominal_params(e, cfg);
    e->level_idx = (int)lround(cfg->disturbance_level * 10.0);
    e->fresh = 1;
}
int cfo_sizeof_env(void) { return (int)sizeof(cfo_env); }

/* reset(): envs/base.py:420-464, hover_free.py:237-289 (task_specific_reset), base.py:241-298 (DR),
 * hover_free.py:492-538 (RandomHJ).  nz[48], un[42]: standard draws, slot layout in DESIGN.md;
 * NULL -> drawn from the env's Philox stream (tag 1 auto-reset / 2 explicit). obs: cfo_obs_dim(). */
void cfo_reset(cfo_env* e, const cf_config* cfg, const double* nz_in, const double* un_in, int explicit_call,
               double* obs) {
    double nzb[CF_RESET_NORMALS], unb[CF_RESET_UNIFORMS];
    const double *nz = nz_in, *un = un_in;
    if (e->fresh) {
        /* constructor-time compute_observation(): bias <- sigma_b * z (tag 3), see cfo_init */
        if (cfg->observation_noise > 0 && !nz_in) {
            double z0[4], dummy[1];
            draw_event(e, 0, 3, 4, 0, z0, dummy);
            double dt = 1.0 / 200.0, sg = GYRO_NOISE_DENSITY / pow(dt, 0.5);
            double sb = pow(-(sg * sg) * (GYRO_CORR_T / 2.0) * (exp(-2.0 * dt / GYRO_CORR_T) - 1.0), 0.5);
            for (int i = 0; i < 3; ++i) e->bias[i] = sb * z0[i];
        }
        e->fresh = 0;
    }
    if (!nz || !un) {
        draw_event(e, e->rng_ctr, explicit_call ? 2u : 1u, CF_RESET_NORMALS, CF_RESET_UNIFORMS, nzb, unb);
        if (!nz) nz = nzb;
        if (!un) un = unb;
        if (explicit_call) e->rng_ctr += 1;
    }
    /* stale body rate of the previous state, read before anything moves (base.py:444 runs before :448) */
    double stale_rpy[3], stale_wb[3];
    body_rates(e, stale_rpy, stale_wb);

    e->iteration = 0; e->ep_len = 0; e->ep_ret = 0.0;
    /* drone.reset(), agents.py:377-386 */
    memset(e->x, 0, sizeof(e->x)); memset(e->ring, 0, sizeof(e->ring)); e->ring_idx = 0;
    /* task_specific_reset, hover_free.py:237-289 / hover.py:1463-1513 */
    double pos[3] = {0, 0, 1}, rpy[3] = {0, 0, 0}, vel[3] = {0, 0, 0}, rate[3] = {0, 0, 0};
    if (cfg->enable_reset_distribution) {
        /* init_xyz is a float32 array and `pos += uniform(...)` adds in place (hover_free.py:245-251):
         * the float64 sum is rounded to float32 */
        for (int i = 0; i < 3; ++i) pos[i] = (double)(float)(pos[i] + unif(-cfg->init_pos_lim, cfg->init_pos_lim, un[i]));
        for (int i = 0; i < 3; ++i) rpy[i] = unif(-cfg->init_angle, cfg->init_angle, un[3 + i]);
        rpy[2] = unif(-cfg->init_yaw, cfg->init_yaw, un[6]);
        for (int i = 0; i < 3; ++i) vel[i] = vel[i] + unif(-cfg->init_vel_lim, cfg->init_vel_lim, un[7 + i]);
        for (int i = 0; i < 3; ++i) rate[i] = rate[i] + unif(-cfg->init_rate_lim, cfg->init_rate_lim, un[10 + i]);
        rate[2] = unif(-cfg->init_yaw_rate_lim, cfg->init_yaw_rate_lim, un[13]);
        double hover_x = sqrt(1.0 / T2W), hover_a = 2.0 * 1.0 / T2W - 1.0;   /* agents.py:152-153 */
        for (int i = 0; i < 4; ++i) e->x[i] = hover_x + 0.02 * nz[i];
        for (int r = 0; r < 2; ++r)
            for (int i = 0; i < 4; ++i) {
                double v = hover_a + 0.02 * nz[4 + 4 * r + i];
                e->ring[r][i] = v < -1.0 ? -1.0 : (v > 1.0 ? 1.0 : v);
            }
        rpy_to_quat(rpy, e->quat);
    } else {
        e->quat[0] = e->quat[1] = e->quat[2] = 0.0; e->quat[3] = 1.0;  /* get_quaternion_from_euler(0), base.py:87 */
    }
    memcpy(e->pos, pos, 24); memcpy(e->vel, vel, 24);
    { /* resetBaseVelocity(angularVelocity = R.T @ rpy_dot), hover_free.py:283-289 (quirk Q3) */
        double R[9]; quat_to_R(e->quat, R); matTvec(R, rate, e->omega_w);
    }
    /* apply_domain_randomization, base.py:241-298; update_motor_dynamics agents.py:208-224 */
    if (cfg->domain_randomization > 0) {
        double f = cfg->domain_randomization;
        e->dt = dr_draw(cfg->sim_time_step, f, un[14]);
        e->mass = dr_draw(URDF_M, f, un[15]);
        for (int i = 0; i < 3; ++i) e->J[i] = dr_draw(URDF_J[i], f, un[16 + i]);
        /* un[19]: force_torque_factor_0 - cancels in the yaw torque, not kept */
        e->ftf1 = dr_draw(FTF1, f, un[20]);
        for (int i = 0; i < 4; ++i) {
            double mtc = dr_draw(cfg->motor_time_constant, f, un[21 + i]);
            double t2w = dr_draw(T2W, f, un[25 + i]);
            double T = mtc < e->dt ? e->dt : mtc;
            e->A[i] = 1.0 - e->dt / T; e->B[i] = e->dt / T;
            e->K[i] = 0.028 * G_ACC * t2w / 4.0;       /* quirk Q8 */
        }
    }
    /* gyro_lpf.set(stale rpy_dot), base.py:444 (quirk Q2) */
    memcpy(e->lpf, stale_wb, 24);
    /* two compute_observation() calls: base.py:453 and, via compute_history, :308 */
    const double* last_action = e->ring[1];              /* view of action_buffer[-1,:], agents.py:386 */
    double oa[17], ob[17];
    int n = compute_observation(e, cfg, last_action, nz + 12, un + 29, oa);
    compute_observation(e, cfg, last_action, nz + 30, un + 35, ob);
    memcpy(obs, oa, 8 * n); memcpy(obs + n, last_action, 32);
    memcpy(obs + n + 4, ob, 8 * n); memcpy(obs + 2 * n + 4, last_action, 32);
    memcpy(e->obs_prev, ob, 8 * n);
    memcpy(e->act_prev1, last_action, 32); memcpy(e->act_prev2, last_action, 32);
    /* Boltzmann() at the end of reset, hover_free.py:536 */
    if (cfg->disturbance_kind == CF_DSTB_HJ_BOLTZMANN) {
        int k = 0;
        while (k < 20 && g_boltz_cdf[k] <= un[41]) ++k;   /* cdf.searchsorted(u, side='right') */
        e->level_idx = k;
    }
}

/* ------------------------------------------------------------------ physics substep */
/* CrazyFlieAgent.apply_action, agents.py:259-298 with control.PWM.act, control.py:94-100 and
 * OUNoise.noise, envs/utils.py:128-133 */
static void apply_action(cfo_env* e, const cf_config* cfg, const double a[4], const double z[4],
                         double f[4], double* z_torque) {
    double delayed[4], t[4];
    memcpy(delayed, e->ring[e->ring_idx], 32);
    memcpy(e->ring[e->ring_idx], a, 32);
    e->ring_idx = (e->ring_idx + 1) % 2;
    double sigma = 0.2 * cfg->motor_thrust_noise;
    for (int i = 0; i < 4; ++i) {
        double c = delayed[i] < -1.0 ? -1.0 : (delayed[i] > 1.0 ? 1.0 : delayed[i]);
        double pwm = 30000.0 + c * 30000.0;
        double dx = OU_THETA * (0.0 - e->ou[i]) + sigma * z[i];
        e->ou[i] = e->ou[i] + dx;
        double thrust_normed = pwm / 60000.0;
        double rot = sqrt(thrust_normed);
        e->x[i] = e->A[i] * e->x[i] + e->B[i] * rot;
        double noisy = (1.0 + e->ou[i]) * (e->x[i] * e->x[i]);
        double n = noisy < 0.0 ? 0.0 : (noisy > 1.0 ? 1.0 : noisy);
        f[i] = e->K[i] * n;
        t[i] = e->ftf1 * f[i] + FTF0;
    }
    *z_torque = ((-t[0] + t[1]) - t[2]) + t[3];
}

/* PybulletPhysicsWithAdversary.step_forward, envs/physics.py:213-250, with Bullet's
 * stepSimulation restated as in oracle/fake_bullet.py (B1-B9) */
static void physics_substep(cfo_env* e, const cf_config* cfg, const double a[4], const double d[2],
                            const double z_ou[4]) {
    double f[4], zt, R[9], wb[3], vb[3];
    apply_action(e, cfg, a, z_ou, f, &zt);
    quat_to_R(e->quat, R);
    matTvec(R, e->omega_w, wb);
    matTvec(R, e->vel, vb);
    /* drag (physics.py:232-241): R @ (k * v_world) handed over as a LINK_FRAME force (quirk Q7) */
    double rpm_sum = 0.0;
    for (int i = 0; i < 4; ++i) rpm_sum += 2.0 * PI * ((e->x[i] * e->x[i]) * 25000.0) / 60.0;
    double kv[3], drag_link[3];
    for (int i = 0; i < 3; ++i) kv[i] = (-1.0 * DRAG[i] * rpm_sum) * e->vel[i];
    matvec(R, kv, drag_link);
    /* ground effect (physics.py:27-58, 244-246): only when the switch is on */
    double fe[4];
    memcpy(fe, f, 32);
    if (cfg->use_ground_effect) {
        double rpy[3]; quat_to_rpy_bullet(e->quat, rpy);
        double max_thrust = (G_ACC * URDF_M) * T2W / 4.0;
        double max_rpm = sqrt((T2W * (G_ACC * URDF_M)) / (4.0 * max_thrust));
        double h_clip = 0.25 * PROP_RADIUS * sqrt((15.0 * max_rpm * max_rpm * URDF_KF * GND_EFF_COEFF) / max_thrust);
        if (fabs(rpy[0]) < PI / 2 && fabs(rpy[1]) < PI / 2)
            for (int i = 0; i < 4; ++i) {
                double pz = e->pos[2] + R[6] * PROP_X[i] + R[7] * PROP_Y[i] + R[8] * PROP_Z;
                if (pz < h_clip) pz = h_clip;
                double q = PROP_RADIUS / (4.0 * pz);
                fe[i] += f[i] * GND_EFF_COEFF * (q * q);
            }
    }
    /* body-frame wrench: thrust at the prop positions (agents.py:311-321), yaw torque (:330-337),
     * adversary roll/pitch torques (:517-533, physics.py:228-229) */
    double tau[3] = {ARM_XY * (-fe[0] - fe[1] + fe[2] + fe[3]) + d[0],
                     ARM_XY * (-fe[0] + fe[1] + fe[2] - fe[3]) + d[1], zt};
    double fb[3] = {drag_link[0], drag_link[1], drag_link[2] + (fe[0] + fe[1] + fe[2] + fe[3])};
    /* gravity, expressed in the body frame */
    double gw[3] = {0.0, 0.0, -G_ACC * e->mass}, gb[3];
    matTvec(R, gw, gb);
    double Jw[3] = {e->J[0] * wb[0], e->J[1] * wb[1], e->J[2] * wb[2]};
    if (cfg->gyroscopic) {                                  /* B2 */
        tau[0] -= wb[1] * Jw[2] - wb[2] * Jw[1];
        tau[1] -= wb[2] * Jw[0] - wb[0] * Jw[2];
        tau[2] -= wb[0] * Jw[1] - wb[1] * Jw[0];
    }
    double wn = sqrt(wb[0] * wb[0] + wb[1] * wb[1] + wb[2] * wb[2]);
    double vn = sqrt(vb[0] * vb[0] + vb[1] * vb[1] + vb[2] * vb[2]);
    double ka = cfg->angular_damping, kl = cfg->linear_damping;   /* B3 */
    double alpha_b[3], acc_b[3], alpha_w[3], acc_w[3];
    for (int i = 0; i < 3; ++i) {
        alpha_b[i] = (tau[i] - Jw[i] * (ka + ka * wn)) / e->J[i];
        acc_b[i] = ((fb[i] + gb[i]) - e->mass * vb[i] * (kl + kl * vn)) / e->mass;
    }
    matvec(R, alpha_b, alpha_w);
    matvec(R, acc_b, acc_w);
    double dt = e->dt;
    for (int i = 0; i < 3; ++i) {                            /* B1 */
        e->omega_w[i] += dt * alpha_w[i];
        e->vel[i] += dt * acc_w[i];
        e->pos[i] += dt * e->vel[i];
    }
    /* exponential-map quaternion update, btMultiBody::stepPositionsMultiDof */
    double w0 = e->omega_w[0], w1 = e->omega_w[1], w2 = e->omega_w[2];
    double ang = sqrt(w0 * w0 + w1 * w1 + w2 * w2), s;
    if (ang * dt > 0.5 * (PI / 2.0)) ang = 0.5 * (PI / 2.0) / dt;
    if (ang < 0.001) s = 0.5 * dt - (dt * dt * dt) * 0.020833333333 * ang * ang;
    else s = sin(0.5 * ang * dt) / ang;
    double dq[4] = {w0 * s, w1 * s, w2 * s, cos(ang * dt * 0.5)}, q[4];
    const double* b = e->quat;
    q[0] = dq[3] * b[0] + dq[0] * b[3] + dq[1] * b[2] - dq[2] * b[1];
    q[1] = dq[3] * b[1] - dq[0] * b[2] + dq[1] * b[3] + dq[2] * b[0];
    q[2] = dq[3] * b[2] + dq[0] * b[1] - dq[1] * b[0] + dq[2] * b[3];
    q[3] = dq[3] * b[3] - dq[0] * b[0] - dq[1] * b[1] - dq[2] * b[2];
    double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    for (int i = 0; i < 4; ++i) e->quat[i] = q[i] / n;
}

/* compute_done overrides, hover_free.py:449-461 (+ base-class limits :124-136 via cfg) */
static int compute_done(const cf_config* cfg, const double pos[3], const double rpy[3], const double wb[3]) {
    int z_limit = pos[2] < 0.2;
    int rp = fabs(rpy[0]) > cfg->done_roll_pitch_rad || fabs(rpy[1]) > cfg->done_roll_pitch_rad;
    int rate = 0;
    for (int i = 0; i < 3; ++i) rate |= rad2deg(fabs(wb[i])) > cfg->done_rate_deg;
    return rp || rate || z_limit;
}

/* step(): hover_free.py:391-444 / 608-665 / 778-833 under gym TimeLimit(500).
 * nz[44], un[8]: standard draws (NULL -> Philox stream, tag 0).  With cfg->auto_reset the env is
 * reset in place when done and `obs` is the first observation of the next episode, `term_obs`
 * (nullable) the last of the finished one. rst_nz/rst_un: injected draws for that reset (nullable). */
void cfo_step(cfo_env* e, const cf_config* cfg, const double a[4], const double* nz_in, const double* un_in,
              const double* rst_nz, const double* rst_un,
              double* obs, double* rew, int* done, double* cost, int* truncated, double* term_obs) {
    double nzb[CF_STEP_NORMALS], unb[CF_STEP_UNIFORMS];
    const double *nz = nz_in, *un = un_in;
    if (!nz || !un) {
        draw_event(e, e->rng_ctr, 0u, CF_STEP_NORMALS, CF_STEP_UNIFORMS, nzb, unb);
        if (!nz) nz = nzb;
        if (!un) un = unb;
    }
    e->rng_ctr += 1;
    /* disturbance, once per control step */
    double d[2] = {0.0, 0.0};
    double level = 0.1 * 0.0;
    if (cfg->disturbance_kind == CF_DSTB_HJ_FIXED || cfg->disturbance_kind == CF_DSTB_HJ_BOLTZMANN) {
        level = (double)e->level_idx / 10.0;
        hj_disturbance(e, level, d);
    } else if (cfg->disturbance_kind == CF_DSTB_RANDOM_BOX) {   /* dstb_space.sample(), hover_free.py:868-878 */
        double lo = (double)(float)(-1e-3), hi = (double)(float)(1e-3);
        d[0] = (double)(float)unif(lo, hi, un[6]);
        d[1] = (double)(float)unif(lo, hi, un[7]);
    }
    double scratch[17];
    int nobs = cfg->observation_noise > 0 ? 13 : 17;
    /* substeps: normals [0:4] OU, [4:13] gyro(call 1, full call whose pos/vel/rot noise is dead),
     * [13:17] OU, [17:26] gyro (call 2).  Generic in aggregate_phy_steps only for 2. */
    for (int s = 0; s < cfg->aggregate_phy_steps; ++s) {
        physics_substep(e, cfg, a, d, nz + 13 * s);
        if (cfg->observation_noise > 0) {
            if (e->iteration % 2 == 0) {
                /* full call; its pos/vel/rot part is overwritten before any obs is emitted: feed zeros */
                double zf[18] = {0}, uf[6] = {0.5, 0.5, 0.5, 0.5, 0.5, 0.5};
                memcpy(zf + 6, nz + 13 * s + 4, 72);
                compute_observation(e, cfg, a, zf, uf, scratch);
            } else {
                compute_observation(e, cfg, a, nz + 13 * s + 4, NULL, scratch);
            }
        }
        e->iteration += 1;
    }
    /* compute_history, base.py:305-321: third compute_observation + ring concat (quirk Q1) */
    double on[17];
    compute_observation(e, cfg, a, nz + 26, un, on);
    const double *ah0, *ah1;
    if (e->ep_len == 0) { ah0 = e->ring[1]; ah1 = e->ring[1]; }
    else if (e->ep_len == 1) { ah0 = e->ring[1]; ah1 = e->act_prev1; }
    else { ah0 = e->act_prev2; ah1 = e->act_prev1; }
    double hist[42];
    memcpy(hist, e->obs_prev, 8 * nobs);
    if (nobs == 17 ) memcpy(hist + 13, e->obs_prev + 13, 32);
    memcpy(hist + nobs, ah0, 32);
    memcpy(hist + nobs + 4, on, 8 * nobs); memcpy(hist + 2 * nobs + 4, ah1, 32);
    memcpy(e->obs_prev, on, 8 * nobs);
    memcpy(e->act_prev2, e->act_prev1, 32); memcpy(e->act_prev1, a, 32);
    /* reward / info / done on the TRUE state (hover_free.py:206-235, 138-166, 449-461) */
    double rpy[3], wb[3];
    body_rates(e, rpy, wb);
    int dn = compute_done(cfg, e->pos, rpy, wb);
    double nca = 0.0;
    for (int i = 0; i < 4; ++i) { double c = a[i] < -1 ? -1 : (a[i] > 1 ? 1 : a[i]); c = 0.5 * (c + 1.0); nca += c * c; }
    double p_action = cfg->penalty_action * sqrt(nca);
    double p_rpy = cfg->penalty_angle * sqrt(rpy[0] * rpy[0] + rpy[1] * rpy[1] + rpy[2] * rpy[2]);
    double p_spin = cfg->penalty_spin * sqrt(wb[0] * wb[0] + wb[1] * wb[1] + wb[2] * wb[2]);
    double p_term = dn ? cfg->penalty_terminal : 0.0;
    double p_vel = cfg->penalty_velocity * sqrt(e->vel[0] * e->vel[0] + e->vel[1] * e->vel[1] + e->vel[2] * e->vel[2]);
    double penalties = ((((p_rpy + 0.0) + p_spin) + p_vel) + p_action) + p_term;
    double r;
    if (cfg->reward_kind == CF_REWARD_FREE) {
        double dist = cfg->penalty_z * fabs(e->pos[2] - 1.0);
        r = -penalties - dist;
    } else {
        double dx = e->pos[0], dy = e->pos[1], dz = e->pos[2] - 1.0;
        r = -sqrt(dx * dx + dy * dy + dz * dz) - penalties;
    }
    double c = 0.0;
    if (fabs(e->pos[0]) > 0.10 || fabs(e->pos[1]) > 0.10 || e->pos[2] > 1.20) c = 1.0;
    if (fabs(rpy[0]) > deg2rad(10) || fabs(rpy[1]) > deg2rad(10)) c = 1.0;
    for (int i = 0; i < 3; ++i) if (fabs(wb[i]) > 0.25) c = 1.0;                     /* quirk Q6 */
    for (int i = 0; i < 3; ++i) if (fabs(a[i]) > cfg->cost_rate_limit) c = 1.0;    /* quirk Q6 */
    e->ep_len += 1; e->ep_ret += r;
    int trunc = 0;
    if (cfg->max_episode_steps > 0 && e->ep_len >= cfg->max_episode_steps) { trunc = !dn; dn = 1; }
    *rew = r; *done = dn;
    if (cost) *cost = c;
    if (truncated) *truncated = trunc;
    int od = 2 * nobs + 8;
    if (dn && cfg->auto_reset) {
        if (term_obs) memcpy(term_obs, hist, 8 * od);
        cfo_reset(e, cfg, rst_nz, rst_un, 0, obs);
    } else {
        memcpy(obs, hist, 8 * od);
    }
}

/* ------------------------------------------------------------------ batch drivers */
/* envs: contiguous array of n cfo_env.  act [n][4], obs [n][obs_dim]; injected draws slot-major
 * [slots][n] (nullable).  Parallel over envs with OpenMP when built with it. */
void cfo_init_batch(cfo_env* envs, int n, const cf_config* cfg, uint64_t seed, int64_t gid_base) {
    for (int i = 0; i < n; ++i) cfo_init(&envs[i], cfg, seed, gid_base + i);
}
static void gather_slots(const double* src, int slots, int n, int i, double* dst) {
    for (int s = 0; s < slots; ++s) dst[s] = src[(size_t)s * n + i];
}
void cfo_reset_batch(cfo_env* envs, int n, const cf_config* cfg, const uint8_t* mask,
                     const double* nz, const double* un, double* obs) {
    int od = cfo_obs_dim(cfg);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n; ++i) {
        if (mask && !mask[i]) continue;
        double zb[CF_RESET_NORMALS], ub[CF_RESET_UNIFORMS];
        if (nz) gather_slots(nz, CF_RESET_NORMALS, n, i, zb);
        if (un) gather_slots(un, CF_RESET_UNIFORMS, n, i, ub);
        cfo_reset(&envs[i], cfg, nz ? zb : NULL, un ? ub : NULL, 1, obs + (size_t)i * od);
    }
}
void cfo_step_batch(cfo_env* envs, int n, const cf_config* cfg, const double* act,
                    const double* nz, const double* un, const double* rst_nz, const double* rst_un,
                    double* obs, double* rew, uint8_t* done, double* cost, uint8_t* truncated,
                    double* term_obs) {
    int od = cfo_obs_dim(cfg);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n; ++i) {
        double zb[CF_STEP_NORMALS], ub[CF_STEP_UNIFORMS], rzb[CF_RESET_NORMALS], rub[CF_RESET_UNIFORMS];
        if (nz) gather_slots(nz, CF_STEP_NORMALS, n, i, zb);
        if (un) gather_slots(un, CF_STEP_UNIFORMS, n, i, ub);
        if (rst_nz) gather_slots(rst_nz, CF_RESET_NORMALS, n, i, rzb);
        if (rst_un) gather_slots(rst_un, CF_RESET_UNIFORMS, n, i, rub);
        int dn = 0, tr = 0;
        double c = 0.0;
        cfo_step(&envs[i], cfg, act + (size_t)i * 4, nz ? zb : NULL, un ? ub : NULL,
                 rst_nz ? rzb : NULL, rst_un ? rub : NULL, obs + (size_t)i * od, &rew[i], &dn, &c, &tr,
                 term_obs ? term_obs + (size_t)i * od : NULL);
        done[i] = (uint8_t)dn;
        if (cost) cost[i] = c;
        if (truncated) truncated[i] = (uint8_t)tr;
    }
}
/* bench.py's CPU baseline uses every host core even when the launcher (torchrun) exported OMP_NUM_THREADS=1 */
void cfo_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
int cfo_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* state record exchange in the CF_S_* word layout of include/cf_abi.h, values as float64
 * (integer words carried as exact doubles) */
void cfo_get_state(const cfo_env* e, double* r) {
    memset(r, 0, sizeof(double) * CF_STATE_WORDS);
    memcpy(r + CF_S_POS, e->pos, 24); r[CF_S_EP_RET] = e->ep_ret;
    memcpy(r + CF_S_QUAT, e->quat, 32); memcpy(r + CF_S_VEL, e->vel, 24);
    memcpy(r + CF_S_OMEGA_W, e->omega_w, 24); memcpy(r + CF_S_LPF, e->lpf, 24);
    memcpy(r + CF_S_GYRO_BIAS, e->bias, 24); memcpy(r + CF_S_MOTOR_X, e->x, 32);
    memcpy(r + CF_S_OU, e->ou, 32); memcpy(r + CF_S_RING, e->ring, 64);
    memcpy(r + CF_S_ACT_PREV1, e->act_prev1, 32); memcpy(r + CF_S_ACT_PREV2, e->act_prev2, 32);
    memcpy(r + CF_S_OBS_PREV, e->obs_prev, 8 * 13);
    r[CF_S_EP_LEN] = e->ep_len; r[CF_S_RNG_CTR] = e->rng_ctr;
    r[CF_S_MISC] = (double)(e->level_idx | (e->ring_idx << 8) | ((e->iteration & 1) << 12)) + (e->fresh ? 2147483648.0 : 0.0);
    r[CF_S_DT] = e->dt; r[CF_S_MASS] = e->mass; memcpy(r + CF_S_INERTIA, e->J, 24);
    r[CF_S_FTF1] = e->ftf1; memcpy(r + CF_S_MOTOR_B, e->B, 32); memcpy(r + CF_S_MOTOR_K, e->K, 32);
}
void cfo_set_state(cfo_env* e, const double* r) {
    memcpy(e->pos, r + CF_S_POS, 24); e->ep_ret = r[CF_S_EP_RET];
    memcpy(e->quat, r + CF_S_QUAT, 32); memcpy(e->vel, r + CF_S_VEL, 24);
    memcpy(e->omega_w, r + CF_S_OMEGA_W, 24); memcpy(e->lpf, r + CF_S_LPF, 24);
    memcpy(e->bias, r + CF_S_GYRO_BIAS, 24); memcpy(e->x, r + CF_S_MOTOR_X, 32);
    memcpy(e->ou, r + CF_S_OU, 32); memcpy(e->ring, r + CF_S_RING, 64);
    memcpy(e->act_prev1, r + CF_S_ACT_PREV1, 32); memcpy(e->act_prev2, r + CF_S_ACT_PREV2, 32);
    memcpy(e->obs_prev, r + CF_S_OBS_PREV, 8 * 13);
    memcpy(e->obs_prev + 13, e->act_prev1, 32);     /* noise-off obs17 tail == u(k-1) */
    e->ep_len = (int)r[CF_S_EP_LEN]; e->rng_ctr = (uint32_t)r[CF_S_RNG_CTR];
    double misc = r[CF_S_MISC];
    e->fresh = misc >= 2147483648.0;
    if (e->fresh) misc -= 2147483648.0;
    int m = (int)misc;
    e->level_idx = m & 0xff; e->ring_idx = (m >> 8) & 0xf; e->iteration = (m >> 12) & 1;
    e->dt = r[CF_S_DT]; e->mass = r[CF_S_MASS]; memcpy(e->J, r + CF_S_INERTIA, 24);
    e->ftf1 = r[CF_S_FTF1]; memcpy(e->B, r + CF_S_MOTOR_B, 32); memcpy(e->K, r + CF_S_MOTOR_K, 32);
    for (int i = 0; i < 4; ++i) e->A[i] = 1.0 - e->B[i];
}
void cfo_get_state_batch(const cfo_env* envs, int n, double* recs) {
    for (int i = 0; i < n; ++i) cfo_get_state(&envs[i], recs + (size_t)i * CF_STATE_WORDS);
}
void cfo_set_state_batch(cfo_env* envs, int n, const double* recs) {
    for (int i = 0; i < n; ++i) cfo_set_state(&envs[i], recs + (size_t)i * CF_STATE_WORDS);
}
/* the reference's exact A (= 1 - T_s/T, agents.py:222) when it is known separately */
void cfo_set_motor_A(cfo_env* e, const double A[4]) { memcpy(e->A, A, 32); }

/* ------------------------------------------------------------------ GAE */
/* Buffer.finish_path + calculate_adv_and_value_targets + discount_cumsum,
 * algs/core.py:105-119, 459-480, 498-535; float64 like scipy.signal.lfilter, time-major [T][N]. */
void cfo_gae(const float* rew, const float* val, const uint8_t* done, const float* end_val,
             const float* last_val, double gamma, double lam, double rew_scale, int T, int N,
             float* adv, float* target_v, float* disc_ret) {
    for (int n = 0; n < N; ++n) {
        double next_val = last_val[n], run_adv = 0.0, run_ret = last_val[n];
        for (int t = T - 1; t >= 0; --t) {
            size_t k = (size_t)t * N + n;
            if (done[k]) { next_val = end_val ? end_val[k] : 0.0; run_adv = 0.0; run_ret = next_val; }
            double r = rew[k];
            run_ret = r + gamma * run_ret;
            double rs = r;
            if (rew_scale > 0.0) {            /* ret_oms(rews, subtract_mean=False, clip=True), float32 torch math */
                float q = (float)r * (float)rew_scale;
                q = q < -10.f ? -10.f : (q > 10.f ? 10.f : q);
                rs = q;
            }
            double delta = rs + gamma * next_val - (double)val[k];
            run_adv = delta + gamma * lam * run_adv;
            adv[k] = (float)run_adv;
            target_v[k] = (float)(run_adv + (double)val[k]);
            disc_ret[k] = (float)run_ret;
            next_val = val[k];
        }
    }
}

/* ------------------------------------------------------------------ HJ batch helpers (tests) */
void cfo_hj_index_batch(const double* states, int n, int* idx) {
    for (int i = 0; i < n; ++i) cfo_hj_index(states + 6 * (size_t)i, idx + 6 * (size_t)i);
}
void cfo_hj_signbits_batch(const float* V, const int* idx, int n, int* bits) {
    for (int i = 0; i < n; ++i) bits[i] = cfo_hj_signbits(V, idx + 6 * (size_t)i);
}
/* packed 2-bit sign table of a whole value table: cell c -> bits (c&3)*2 of byte c>>2 */
void cfo_hj_signs_all(const float* V, uint8_t* out) {
    init_static();
    long n_bytes = (CF_HJ_CELLS + 3) / 4;
    memset(out, 0, (size_t)n_bytes);
#pragma omp parallel for schedule(static)
    for (long b = 0; b < n_bytes; ++b) {
        unsigned v = 0;
        for (int j = 0; j < 4; ++j) {
            long c = 4 * b + j;
            if (c >= CF_HJ_CELLS) break;
            int idx[6]; long r = c;
            for (int d = 5; d >= 0; --d) { idx[d] = (int)(r % 15); r /= 15; }
            v |= (unsigned)cfo_hj_signbits(V, idx) << (2 * j);
        }
        out[b] = (uint8_t)v;
    }
}

/* ------------------------------------------------------------------ constructor defaults */
/* The kwargs defaults of the registered env ids (phoenix_drone_simulation/__init__.py:8-109; constructors and
 * compute_done overrides in envs/hover_free.py:18-39,303-326,449-461 and envs/hover.py:18-40,565-590,1290-1317;
 * DroneBaseEnv defaults envs/base.py:26-47), so that bench.py's reference arm needs nothing from the product
 * package.  tests/test_oracle_golden.py checks it field by field against cf_config_default of the CUDA library. */
typedef struct { const char* id; int reward, dstb; double level, rp_deg, rate_deg, init_angle, rate_lim_deg; } cfo_id_defaults;
static const cfo_id_defaults CFO_IDS[] = {
    {"DroneHoverBulletFreeEnvWithoutAdversary-v0", CF_REWARD_FREE, CF_DSTB_NONE, 0.0, 75, 1000, PI / 6, 200},
    {"DroneHoverBulletFreeEnvWithAdversary-v0", CF_REWARD_FREE, CF_DSTB_HJ_FIXED, 1.5, 75, 1000, PI / 6, 200},
    {"DroneHoverBulletFreeEnvWithRandomHJAdversary-v0", CF_REWARD_FREE, CF_DSTB_HJ_BOLTZMANN, 0.0, 75, 1000, PI / 6, 200},
    {"DroneHoverBulletEnv-v0", CF_REWARD_TARGET, CF_DSTB_NONE, 0.0, 60, 300, PI / 6, 200},
    {"DroneHoverBulletEnvWithoutAdversary-v0", CF_REWARD_TARGET, CF_DSTB_NONE, 0.0, 75, 1000, PI / 6, 200},
    {"DroneHoverBulletEnvWithAdversary-v0", CF_REWARD_TARGET, CF_DSTB_HJ_FIXED, 0.0, 75, 1000, PI / 6, 200},
    {"DroneHoverBulletEnvWithRandomHJAdversary-v0", CF_REWARD_TARGET, CF_DSTB_HJ_BOLTZMANN, 0.0, 75, 1000, PI / 6, 200},
    {"DroneHoverBulletEnvWithCurriculumHJAdversary-v0", CF_REWARD_TARGET, CF_DSTB_HJ_BOLTZMANN, 0.0, 75, 1000, PI / 6, 200},
    {"DroneHoverBulletEnvWithRandomAdversary-v0", CF_REWARD_TARGET, CF_DSTB_RANDOM_BOX, 0.0, 75, 1000, PI / 6, 200},
    {"DroneHoverBulletEnvWithAdversaryInitial-v0", CF_REWARD_TARGET, CF_DSTB_HJ_FIXED, 1.5, 75, 1000, PI / 4, 300},
};
int cfo_config_default(const char* env_id, cf_config* c) {
    const cfo_id_defaults* d = NULL;
    for (size_t k = 0; k < sizeof(CFO_IDS) / sizeof(CFO_IDS[0]); ++k)
        if (strcmp(CFO_IDS[k].id, env_id) == 0) d = &CFO_IDS[k];
    if (!d) return -1;
    memset(c, 0, sizeof(*c));
    c->abi_version = CF_ABI_VERSION;
    c->reward_kind = d->reward; c->disturbance_kind = d->dstb;
    c->observation_noise = 1; c->enable_reset_distribution = 1;
    c->aggregate_phy_steps = 2; c->max_episode_steps = 500;
    c->gyroscopic = 1; c->use_ground_effect = 0;
    c->hj_mode = CF_HJ_SIGN_TABLE; c->rng_mode = CF_RNG_PHILOX; c->auto_reset = 1;
    c->domain_randomization = 0.10; c->disturbance_level = d->level;
    if (d->reward == CF_REWARD_FREE) {
        c->penalty_action = 0.0; c->penalty_angle = 1.0; c->penalty_spin = 1.0;
        c->penalty_terminal = 1000.0; c->penalty_velocity = 1.0; c->penalty_z = 0.0;
    } else {
        c->penalty_action = 1e-4; c->penalty_angle = 0.0; c->penalty_spin = 1e-4;
        c->penalty_terminal = 1000.0; c->penalty_velocity = 0.0; c->penalty_z = 0.0;
    }
    c->done_roll_pitch_rad = deg2rad(d->rp_deg); c->done_rate_deg = d->rate_deg;
    c->init_pos_lim = 0.25; c->init_angle = d->init_angle; c->init_yaw = 2.0 * PI;
    c->init_vel_lim = 0.1; c->init_rate_lim = deg2rad(d->rate_lim_deg); c->init_yaw_rate_lim = deg2rad(20.0);
    c->cost_rate_limit = deg2rad(d->rate_lim_deg);
    c->sim_time_step = 1.0 / 200.0; c->motor_time_constant = 0.080;
    c->motor_thrust_noise = 0.05; c->latency = 0.015;
    c->linear_damping = 0.04; c->angular_damping = 0.04;
    return 0;
}
