// cf_abi.cu - host side of the C ABI declared in include/cf_abi.h: handle management, config
// defaults of the registered env ids, HJ table upload, launch wrappers and the host-pointer API.
// No C++ exception crosses the boundary; every entry point returns a cf_status.
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <time.h>

#include <new>
#include <vector>

#include "cf_kernels.h"

using namespace cf;

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, const char* a = "", long long b = 0) {
    snprintf(g_err, sizeof(g_err), fmt, a, b);
    return code;
}
int cuda_fail(cudaError_t e, const char* where) {
    snprintf(g_err, sizeof(g_err), "%s: %s", where, cudaGetErrorString(e));
    return CF_ERR_CUDA;
}
#define CU(call)                                                  \
    do {                                                          \
        cudaError_t e_ = (call);                                  \
        if (e_ != cudaSuccess) return cuda_fail(e_, #call);       \
    } while (0)

}  // namespace
namespace cf {
// error text for entry points implemented in other translation units (cf_hj_solver.cu)
void set_error(const char* where, const char* what) { snprintf(g_err, sizeof(g_err), "%s: %s", where, what); }
}  // namespace cf
namespace {

constexpr double kPiD = 3.141592653589793;
double deg2rad(double x) { return kPiD * x / 180.0; }

struct EventPair { cudaEvent_t a, b; };

}  // namespace

struct cf_env {
    cf_config cfg;
    DevCfg dev;
    DevTables tab;
    int n = 0, n_pad = 0, device = 0, obs_dim = 34;
    float* S = nullptr;                          // [n_pad / CF_TILE][CF_STATE_WORDS][CF_TILE]
    float* value_dev[CF_MAX_LEVELS] = {};        // host copy of the device pointer array tab.value
    bool has_signs[CF_MAX_LEVELS] = {};
    const float** value_ptrs_dev = nullptr;      // device array [CF_MAX_LEVELS]
    uint8_t* signs_slab = nullptr;               // [CF_MAX_LEVELS][CF_SIGN_STRIDE]
    float* thr_dev = nullptr;                    // [6][16]
    const cf_env* table_owner = nullptr;
    float* inj[4] = {};              // step z, step u, reset z, reset u
    float* fin_ret = nullptr;
    int* fin_len = nullptr;
    double* agg = nullptr;
    int* reset_list = nullptr;       // deferred resets of single-step launches: [n] env indices + [2] counters (cf_types.h StepIO)
    int* reset_count = nullptr;
    float* rec_dev = nullptr;        // scratch for get/set_state
    // host-pointer API: private stream + device mirrors of the caller's host buffers
    cudaStream_t hstream = nullptr;
    cudaEvent_t probe_ev[5] = {};                           // tools/ only: CF_PROBE_E2E_TIMELINE (see step_host_enqueue)
    double probe_cpu[3] = {};                               // CPU clock [ms]: enqueue entry, enqueue return, sync entry
    // ordering between the two call families on one handle (both touch S): 1 = device-pointer API on the caller's stream,
    // 2 = host-pointer API on hstream; on a switch the new stream waits for an event recorded on the previous one
    cudaEvent_t order_ev = nullptr;
    int last_family = 0;
    cudaStream_t last_dev_stream = nullptr;
    float *d_act = nullptr, *d_obs = nullptr, *d_rew = nullptr, *d_cost = nullptr, *d_term = nullptr;
    uint8_t *d_done = nullptr, *d_trunc = nullptr, *d_mask = nullptr;
    uint8_t* d_out_slab = nullptr;                          // d_obs | d_rew | d_cost | d_done carved from one allocation (ensure_outputs)
    bool tables_borrowed = false;    // tables belong to another handle (cf_share_hj_tables)
    long long launches = 0;
    unsigned long long* trace = nullptr;   // caller-owned device buffer (cf_set_trace_buffer)
    const cf_env* pf_next = nullptr;       // cf_set_prefetch_next
    const float* pf_act = nullptr;
    bool timing = false;
    std::vector<EventPair> ev_pool;
    size_t ev_used = 0;
    double ev_ms = 0.0;
    long long ev_count = 0;
};

namespace {

struct IdDefaults {
    const char* id;
    int reward, dstb;
    double level, rp_deg, rate_deg, init_angle, cost_rate_deg;
};
// phoenix_drone_simulation/__init__.py:8-109 (registered ids) and the constructors / compute_done
// overrides of envs/hover_free.py and envs/hover.py.  level: the class's hard-coded
// disturbance_level (hover_free.py:319 -> 1.5; hover.py:577 -> 0.0; hover.py:1311 -> 1.5).
const IdDefaults kIds[] = {
    {"DroneHoverBulletFreeEnvWithoutAdversary-v0", CF_REWARD_FREE, CF_DSTB_NONE, 0.0, 75, 1000, kPiD / 6, 200},
    {"DroneHoverBulletFreeEnvWithAdversary-v0", CF_REWARD_FREE, CF_DSTB_HJ_FIXED, 1.5, 75, 1000, kPiD / 6, 200},
    {"DroneHoverBulletFreeEnvWithRandomHJAdversary-v0", CF_REWARD_FREE, CF_DSTB_HJ_BOLTZMANN, 0.0, 75, 1000, kPiD / 6, 200},
    {"DroneHoverBulletEnv-v0", CF_REWARD_TARGET, CF_DSTB_NONE, 0.0, 60, 300, kPiD / 6, 200},
    {"DroneHoverBulletEnvWithoutAdversary-v0", CF_REWARD_TARGET, CF_DSTB_NONE, 0.0, 75, 1000, kPiD / 6, 200},
    {"DroneHoverBulletEnvWithAdversary-v0", CF_REWARD_TARGET, CF_DSTB_HJ_FIXED, 0.0, 75, 1000, kPiD / 6, 200},
    {"DroneHoverBulletEnvWithRandomHJAdversary-v0", CF_REWARD_TARGET, CF_DSTB_HJ_BOLTZMANN, 0.0, 75, 1000, kPiD / 6, 200},
    {"DroneHoverBulletEnvWithCurriculumHJAdversary-v0", CF_REWARD_TARGET, CF_DSTB_HJ_BOLTZMANN, 0.0, 75, 1000, kPiD / 6, 200},
    {"DroneHoverBulletEnvWithRandomAdversary-v0", CF_REWARD_TARGET, CF_DSTB_RANDOM_BOX, 0.0, 75, 1000, kPiD / 6, 200},
    {"DroneHoverBulletEnvWithAdversaryInitial-v0", CF_REWARD_TARGET, CF_DSTB_HJ_FIXED, 1.5, 75, 1000, kPiD / 4, 300},
};

void grid_tables(double grid[6][CF_HJ_PTS], double dx[6], double cdf[21]) {
    // odp/Grid/GridProcessing.py:6-51 with the arguments of distur_gener.py:179; np.linspace:
    // step = (hi-lo)/14, y_i = lo + i*step, last point forced to hi
    for (int d = 0; d < 6; ++d) {
        double lo = d < 3 ? -kPiD / 2.4 : -kPiD, hi = d < 3 ? kPiD / 2.4 : kPiD;
        if (d < 3) hi = lo + (hi - lo) * (1.0 - 1.0 / 15.0);
        dx[d] = (hi - lo) / (15.0 - 1.0);
        double step = (hi - lo) / 14.0;
        for (int i = 0; i < 15; ++i) grid[d][i] = lo + (double)i * step;
        grid[d][14] = hi;
    }
    // Boltzmann() (envs/utils.py:27-39) through legacy np.random.choice(p=): p = w/np.sum(w)
    // (pairwise 8-way unrolled sum for n < 128), cdf = cumsum(p), cdf /= cdf[-1]
    double w[21], s, c = 0.0, r[8];
    for (int i = 0; i < 21; ++i) w[i] = exp(-1.0 * (0.0 + (double)i * 0.1));
    for (int j = 0; j < 8; ++j) r[j] = w[j];
    int i = 8;
    for (; i < 21 - (21 % 8); i += 8)
        for (int j = 0; j < 8; ++j) r[j] += w[i + j];
    s = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
    for (; i < 21; ++i) s += w[i];
    for (int k = 0; k < 21; ++k) { c += w[k] / s; cdf[k] = c; }
    double last = cdf[20];
    for (int k = 0; k < 21; ++k) cdf[k] /= last;
}

// Grid.get_index (odp/Grid/GridProcessing.py:53-72) in float64, as hj_index_dim on the device
int index_f64(const double* g, double s) {
    if (!(s == s)) return CF_HJ_PTS - 1;
    int i = 0;
    while (i < 15 && g[i] < s) ++i;                      /* np.searchsorted(side='left') */
    if (i > 0 && (i == 15 || fabs(s - g[i - 1]) < fabs(s - g[i]))) i -= 1;
    return i;
}
// float32 <-> order-preserving integer
int64_t f2ord(float f) { int32_t b; memcpy(&b, &f, 4); return b >= 0 ? (int64_t)b : -(int64_t)(b & 0x7fffffff); }
float ord2f(int64_t o) { int32_t b = o >= 0 ? (int32_t)o : (int32_t)((uint32_t)(-o) | 0x80000000u); float f; memcpy(&f, &b, 4); return f; }
// thr[d][j] = smallest float32 whose float64-rule index is >= j+1 (the rule is monotone in s), found
// by bisection over the float32 values between grid points j and j+1; slots 14, 15 = NaN (never true)
void index_thresholds(const double grid[6][CF_HJ_PTS], float thr[6][16]) {
    for (int d = 0; d < 6; ++d) {
        for (int j = 0; j < 14; ++j) {
            int64_t lo = f2ord(nextafterf((float)grid[d][j], -INFINITY));       /* index <= j */
            int64_t hi = f2ord(nextafterf((float)grid[d][j + 1], INFINITY));    /* index >= j+1 */
            while (hi - lo > 1) {
                int64_t mid = lo + (hi - lo) / 2;
                if (index_f64(grid[d], (double)ord2f(mid)) >= j + 1) hi = mid; else lo = mid;
            }
            thr[d][j] = ord2f(hi);
        }
        thr[d][14] = thr[d][15] = NAN;
    }
}

int derive(const cf_config& c, uint64_t seed, int64_t gid_base, DevCfg& d) {
    if (c.abi_version != CF_ABI_VERSION) return fail(CF_ERR_INVALID, "cf_config.abi_version mismatch%s (got %lld)", "", c.abi_version);
    if (c.aggregate_phy_steps != 2) return fail(CF_ERR_INVALID, "aggregate_phy_steps must be 2%s (got %lld)", "", c.aggregate_phy_steps);
    if (fabs(c.sim_time_step - 0.005) > 1e-12) return fail(CF_ERR_INVALID, "sim_time_step must be 1/200%s", "");
    // action ring of int(latency // time_step) slots (agents.py:170-175); the kernel holds exactly 2
    {   // Python float floor division
        double mod = fmod(c.latency, c.sim_time_step), div = (c.latency - mod) / c.sim_time_step, fl = floor(div);
        if (div - fl > 0.5) fl += 1.0;
        if (c.latency < c.sim_time_step || (int)fl != 2)
            return fail(CF_ERR_INVALID, "latency must give an action ring of 2 slots (0.010 <= latency < 0.015+)%s", "");
    }
    if (c.reward_kind != CF_REWARD_FREE && c.reward_kind != CF_REWARD_TARGET) return fail(CF_ERR_INVALID, "bad reward_kind%s", "");
    if (c.disturbance_kind < CF_DSTB_NONE || c.disturbance_kind > CF_DSTB_RANDOM_BOX) return fail(CF_ERR_INVALID, "bad disturbance_kind%s", "");
    int li = (int)lround(c.disturbance_level * 10.0);
    if (c.disturbance_level < 0 || fabs(li / 10.0 - c.disturbance_level) > 1e-9)
        return fail(CF_ERR_INVALID, "disturbance_level must be a non-negative multiple of 0.1%s", "");
    if (li > 30) return fail(CF_ERR_INVALID, "disturbance_level must be <= 3.0 (distur_gener.py:154)%s", "");
    memset(&d, 0, sizeof(d));
    d.reward_kind = c.reward_kind; d.dstb_kind = c.disturbance_kind;
    d.obs_noise = c.observation_noise > 0; d.reset_dist = c.enable_reset_distribution != 0;
    d.agg_steps = 2; d.max_ep_steps = c.max_episode_steps;
    d.gyroscopic = c.gyroscopic != 0; d.ground_effect = c.use_ground_effect != 0;
    d.hj_mode = c.hj_mode; d.auto_reset = c.auto_reset != 0;
    d.dr_on = c.domain_randomization > 0; d.injected = c.rng_mode == CF_RNG_INJECTED;
    d.fixed_level_idx = c.disturbance_kind == CF_DSTB_HJ_FIXED ? li : 0;
    d.obs_dim = d.obs_noise ? 34 : 42;
    d.dr = (float)c.domain_randomization;
    d.pen_action = (float)c.penalty_action; d.pen_angle = (float)c.penalty_angle; d.pen_spin = (float)c.penalty_spin;
    d.pen_terminal = (float)c.penalty_terminal; d.pen_velocity = (float)c.penalty_velocity; d.pen_z = (float)c.penalty_z;
    d.done_rp_rad = (float)c.done_roll_pitch_rad; d.done_rate_deg = (float)c.done_rate_deg;
    d.init_pos = (float)c.init_pos_lim; d.init_angle = (float)c.init_angle; d.init_yaw = (float)c.init_yaw;
    d.init_vel = (float)c.init_vel_lim; d.init_rate = (float)c.init_rate_lim; d.init_yaw_rate = (float)c.init_yaw_rate_lim;
    d.cost_rate_limit = (float)c.cost_rate_limit;
    d.sim_dt = (float)c.sim_time_step; d.motor_tc = (float)c.motor_time_constant;
    d.ou_sigma = (float)(0.2 * c.motor_thrust_noise);         // agents.py:206
    d.lin_damp = (float)c.linear_damping; d.ang_damp = (float)c.angular_damping;
    const double M = 0.030, G = 9.81, T2W = 1.8;                // cf21x_bullet.urdf:12,21
    d.nomA = (float)(1.0 - c.sim_time_step / c.motor_time_constant);
    d.nomB = (float)(c.sim_time_step / c.motor_time_constant);
    d.nomK = (float)((G * M) * T2W / 4.0);
    d.hover_x = (float)sqrt(1.0 / T2W); d.hover_a = (float)(2.0 * 1.0 / T2W - 1.0);
    // SensorNoise.add_noise_to_omega constants at dt = 1/SIM_FREQ (sensors.py:121-134, defaults :26-30)
    const double dt = 1.0 / 200.0, dens = 0.000175, corr = 1000.0;
    double sg = dens / pow(dt, 0.5);
    d.gyro_sb = (float)pow(-(sg * sg) * (corr / 2.0) * (exp(-2.0 * dt / corr) - 1.0), 0.5);
    d.gyro_pi = (float)exp(-dt / corr);
    d.gyro_rw = 0.0105f; d.gyro_ton = (float)deg2rad(5.0);
    d.quat_std = (float)deg2rad(0.1); d.quat_unif = (float)deg2rad(0.05);
    {   // calculate_ground_effect, physics.py:27-58
        const double kf = 3.16e-10, gnd = 11.36859, pr = 2.31348e-2;
        double max_thrust = (G * M) * T2W / 4.0, max_rpm = sqrt((T2W * (G * M)) / (4.0 * max_thrust));
        d.gnd_hclip = (float)(0.25 * pr * sqrt((15.0 * max_rpm * max_rpm * kf * gnd) / max_thrust));
    }
    {   // Philox4x32-10 key schedule of the seed (constant-bank operands in the kernels)
        uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
        for (int r = 0; r < 10; ++r) { d.pk[r][0] = k0; d.pk[r][1] = k1; k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
    }
    {
        double grid[6][CF_HJ_PTS], dx[6], cdf[21];
        grid_tables(grid, dx, cdf);
        for (int k = 0; k < 6; ++k) { d.hj_lo[k] = (float)grid[k][0]; d.hj_inv_dx[k] = (float)(1.0 / dx[k]); }
        float thr[6][16];
        index_thresholds(grid, thr);
        memcpy(d.hj_thr, thr, sizeof(thr));
        for (int k = 0; k < 20; ++k) {                      // smallest float32 t with cdf[k] <= (double)t
            float t = (float)cdf[k];
            if ((double)t < cdf[k]) t = nextafterf(t, INFINITY);
            d.boltz_thr[k] = t;
        }
    }
    d.gid_base = gid_base;
    return CF_OK;
}

void begin_timing(cf_env* e, cudaStream_t st, EventPair** out) {
    *out = nullptr;
    if (!e->timing) return;
    if (e->ev_used == e->ev_pool.size()) {
        if (e->ev_pool.size() >= 8192) return;
        EventPair p;
        if (cudaEventCreate(&p.a) != cudaSuccess || cudaEventCreate(&p.b) != cudaSuccess) return;
        e->ev_pool.push_back(p);
    }
    *out = &e->ev_pool[e->ev_used++];
    cudaEventRecord((*out)->a, st);
}
void end_timing(EventPair* p, cudaStream_t st) {
    if (p) cudaEventRecord(p->b, st);
}
void drain_timing(cf_env* e) {
    for (size_t i = 0; i < e->ev_used; ++i) {
        float ms = 0.f;
        cudaEventSynchronize(e->ev_pool[i].b);
        if (cudaEventElapsedTime(&ms, e->ev_pool[i].a, e->ev_pool[i].b) == cudaSuccess) { e->ev_ms += ms; e->ev_count += 1; }
    }
    e->ev_used = 0;
}

int check_tables(cf_env* e) {
    const cf_config& c = e->cfg;
    const cf_env* o = e->tables_borrowed ? e->table_owner : e;
    auto have = [&](int li) { return c.hj_mode == CF_HJ_SIGN_TABLE ? o->has_signs[li] : o->value_dev[li] != nullptr; };
    if (c.disturbance_kind == CF_DSTB_HJ_FIXED) {
        int li = e->dev.fixed_level_idx;
        if (li > 0 && !have(li)) return fail(CF_ERR_NO_TABLE, "no HJ value table loaded for level index %s%lld", "", li);
    } else if (c.disturbance_kind == CF_DSTB_HJ_BOLTZMANN) {
        for (int li = 1; li <= 20; ++li)
            if (!have(li)) return fail(CF_ERR_NO_TABLE, "no HJ value table loaded for level index %s%lld", "", li);
    }
    return CF_OK;
}

void order_families(cf_env* e, cudaStream_t st) {
    const int family = st == e->hstream ? 2 : 1;
    if (e->last_family && e->last_family != family && e->order_ev) {
        cudaStream_t prev = e->last_family == 2 ? e->hstream : e->last_dev_stream;
        if (cudaEventRecord(e->order_ev, prev) == cudaSuccess) cudaStreamWaitEvent(st, e->order_ev, 0);
        else cudaGetLastError();                            // the caller's previous stream is gone: nothing left to order against
    }
    e->last_family = family;
    if (family == 1) e->last_dev_stream = st;
}

int step_common(cf_env* e, int k, const float* act, float* obs, float* rew, uint8_t* done, float* cost,
                float* term_obs, uint8_t* truncated, cudaStream_t st) {
    if (!e || !act || !obs || !rew || !done || k < 1) return fail(CF_ERR_INVALID, "cf_step: null argument%s", "");
    int rc = check_tables(e);
    if (rc != CF_OK) return rc;
    if (e->dev.injected && (!e->inj[0] || !e->inj[1] || (e->dev.auto_reset && (!e->inj[2] || !e->inj[3]))))
        return fail(CF_ERR_INVALID, "rng_mode INJECTED: call cf_set_noise_override first%s", "");
    CU(cudaSetDevice(e->device));
    StepIO io;
    io.act = act; io.obs = obs; io.rew = rew; io.done = done; io.cost = cost;
    io.term_obs = term_obs; io.truncated = truncated;
    io.inj_step_z = e->inj[0]; io.inj_step_u = e->inj[1]; io.inj_reset_z = e->inj[2]; io.inj_reset_u = e->inj[3];
    io.fin_ret = e->fin_ret; io.fin_len = e->fin_len; io.agg = e->agg;
    io.reset_list = e->reset_list; io.reset_count = e->reset_count;
    // a full tile's observation rows leave with one bulk copy when every tile's first row is 16-byte aligned
    io.trace = e->trace;
    io.pf_S = e->pf_next ? e->pf_next->S : nullptr; io.pf_act = e->pf_next ? e->pf_act : nullptr;
    io.pf_tiles = e->pf_next ? e->pf_next->n_pad / CF_TILE : 0; io.pf_n = e->pf_next ? e->pf_next->n : 0;
    io.bulk_obs = ((uintptr_t)obs % 16 == 0) && (k == 1 || ((size_t)e->n * e->obs_dim * 4) % 16 == 0);
    if ((uintptr_t)act % 16 != 0) return fail(CF_ERR_INVALID, "cf_step: act must be 16-byte aligned%s", "");
    if ((uintptr_t)obs % 8 != 0 || (term_obs && (uintptr_t)term_obs % 8 != 0))
        return fail(CF_ERR_INVALID, "cf_step: obs rows must be 8-byte aligned%s", "");
    order_families(e, st);
    EventPair* ep;
    begin_timing(e, st, &ep);
    cudaError_t lerr = launch_step(e->S, e->n, k, e->dev, e->tab, io, st);
    end_timing(ep, st);
    if (lerr != cudaSuccess) return cuda_fail(lerr, "launch_step");
    e->launches += (k == 1 && e->dev.auto_reset) ? 2 : 1;    // the step kernel + the list-driven reset kernel of single-step launches
    CU(cudaGetLastError());
    return CF_OK;
}

template <typename T>
int ensure(T** p, size_t count) {
    if (*p) return CF_OK;
    cudaError_t err = cudaMalloc((void**)p, count * sizeof(T));
    if (err != cudaSuccess) { cuda_fail(err, "cudaMalloc"); return CF_ERR_ALLOC; }
    return CF_OK;
}

// Device-side outputs of the host-pointer API, in ONE allocation laid out [obs n x od floats | rew n floats | cost n floats | done n
// bytes] (n % 4 == 0: every section starts 16-byte aligned): a caller whose host arrays are adjacent in the same order gets them
// in one download (step_host_enqueue).  Other n: four allocations as before.
int ensure_outputs(cf_env* e) {
    if (e->d_obs && e->d_rew && e->d_cost && e->d_done) return CF_OK;
    const size_t n = e->n, od = e->obs_dim;
    if (n % 4 == 0 && !e->d_obs && !e->d_rew && !e->d_cost && !e->d_done) {
        cudaError_t err = cudaMalloc((void**)&e->d_out_slab, n * (od * sizeof(float) + 2 * sizeof(float) + 1));
        if (err != cudaSuccess) { cuda_fail(err, "cudaMalloc"); return CF_ERR_ALLOC; }
        e->d_obs = (float*)e->d_out_slab;
        e->d_rew = e->d_obs + n * od;
        e->d_cost = e->d_rew + n;
        e->d_done = (uint8_t*)(e->d_cost + n);
        return CF_OK;
    }
    int rc = ensure(&e->d_obs, n * od);
    if (rc == CF_OK) rc = ensure(&e->d_rew, n);
    if (rc == CF_OK) rc = ensure(&e->d_cost, n);
    if (rc == CF_OK) rc = ensure(&e->d_done, n);
    return rc;
}

}  // namespace

extern "C" {

const char* cf_last_error(void) { return g_err; }
int cf_abi_version(void) { return CF_ABI_VERSION; }

int cf_config_default(const char* env_id, cf_config* c) {
    if (!env_id || !c) return fail(CF_ERR_INVALID, "cf_config_default: null argument%s", "");
    const IdDefaults* d = nullptr;
    for (const IdDefaults& k : kIds)
        if (strcmp(k.id, env_id) == 0) d = &k;
    if (!d) return fail(CF_ERR_INVALID, "unknown env id '%s'", env_id);
    memset(c, 0, sizeof(*c));
    c->abi_version = CF_ABI_VERSION;
    c->reward_kind = d->reward; c->disturbance_kind = d->dstb;
    c->observation_noise = 1; c->enable_reset_distribution = 1;          // base.py:26-47
    c->aggregate_phy_steps = 2; c->max_episode_steps = 500;              // __init__.py max_episode_steps
    c->gyroscopic = 1; c->use_ground_effect = 0;                         // physics.py:18,25
    c->hj_mode = CF_HJ_SIGN_TABLE; c->rng_mode = CF_RNG_PHILOX; c->auto_reset = 1;
    c->domain_randomization = 0.10; c->disturbance_level = d->level;
    if (d->reward == CF_REWARD_FREE) {                                   // hover_free.py:29-34
        c->penalty_action = 0.0; c->penalty_angle = 1.0; c->penalty_spin = 1.0;
        c->penalty_terminal = 1000.0; c->penalty_velocity = 1.0; c->penalty_z = 0.0;
    } else {                                                             // hover.py:28-32
        c->penalty_action = 1e-4; c->penalty_angle = 0.0; c->penalty_spin = 1e-4;
        c->penalty_terminal = 1000.0; c->penalty_velocity = 0.0; c->penalty_z = 0.0;
    }
    c->done_roll_pitch_rad = deg2rad(d->rp_deg); c->done_rate_deg = d->rate_deg;
    c->init_pos_lim = 0.25; c->init_angle = d->init_angle; c->init_yaw = 2.0 * kPiD;   // hover_free.py:249-274
    c->init_vel_lim = 0.1; c->init_rate_lim = deg2rad(d->cost_rate_deg); c->init_yaw_rate_lim = deg2rad(20.0);
    c->cost_rate_limit = deg2rad(d->cost_rate_deg);                      // rpy_dot_limit, hover_free.py:75 / hover.py:1317
    c->sim_time_step = 1.0 / 200.0; c->motor_time_constant = 0.080;
    c->motor_thrust_noise = 0.05; c->latency = 0.015;
    c->linear_damping = 0.04; c->angular_damping = 0.04;
    return CF_OK;
}

int cf_create(const cf_config* cfg, int32_t n_envs, int32_t device, uint64_t seed, int64_t env_id_base, cf_env** out) {
    if (!cfg || !out || n_envs < 1) return fail(CF_ERR_INVALID, "cf_create: bad argument%s", "");
    *out = nullptr;
    DevCfg dev;
    int rc = derive(*cfg, seed, env_id_base, dev);
    if (rc != CF_OK) return rc;
    int count = 0;
    CU(cudaGetDeviceCount(&count));
    if (device < 0 || device >= count) return fail(CF_ERR_INVALID, "no CUDA device %s%lld", "", device);
    CU(cudaSetDevice(device));
    cf_env* e = new (std::nothrow) cf_env();
    if (!e) return fail(CF_ERR_ALLOC, "out of host memory%s", "");
    e->cfg = *cfg; e->dev = dev; e->n = n_envs; e->device = device; e->obs_dim = dev.obs_dim;
    e->n_pad = (n_envs + CF_TILE - 1) / CF_TILE * CF_TILE;
    memset(&e->tab, 0, sizeof(e->tab));
    double grid[6][CF_HJ_PTS], dx[6], cdf[21];
    float thr[6][16];
    grid_tables(grid, dx, cdf);
    index_thresholds(grid, thr);
    const bool hj = cfg->disturbance_kind == CF_DSTB_HJ_FIXED || cfg->disturbance_kind == CF_DSTB_HJ_BOLTZMANN;
    cudaError_t err = upload_constants(grid, dx, cdf);
    if (err == cudaSuccess) err = cudaMalloc((void**)&e->S, sizeof(float) * CF_STATE_WORDS * (size_t)e->n_pad);
    if (err == cudaSuccess) err = cudaMalloc((void**)&e->fin_ret, sizeof(float) * (size_t)n_envs);
    if (err == cudaSuccess) err = cudaMalloc((void**)&e->fin_len, sizeof(int) * (size_t)n_envs);
    if (err == cudaSuccess) err = cudaMalloc((void**)&e->agg, sizeof(double) * 4 * CF_AGG_SLOTS);
    if (err == cudaSuccess) err = cudaMalloc((void**)&e->reset_list, sizeof(int) * (size_t)n_envs);
    if (err == cudaSuccess) err = cudaMalloc((void**)&e->reset_count, sizeof(int) * 2);
    if (err == cudaSuccess) err = cudaMemset(e->reset_count, 0, sizeof(int) * 2);
    if (err == cudaSuccess) err = cudaMalloc((void**)&e->thr_dev, sizeof(thr));
    if (err == cudaSuccess) err = cudaMalloc((void**)&e->value_ptrs_dev, sizeof(float*) * CF_MAX_LEVELS);
    if (err == cudaSuccess && hj) err = cudaMalloc((void**)&e->signs_slab, (size_t)CF_SIGN_STRIDE * CF_MAX_LEVELS);
    if (err == cudaSuccess) err = cudaMemset(e->fin_ret, 0, sizeof(float) * (size_t)n_envs);
    if (err == cudaSuccess) err = cudaMemset(e->fin_len, 0, sizeof(int) * (size_t)n_envs);
    if (err == cudaSuccess) err = cudaMemset(e->agg, 0, sizeof(double) * 4 * CF_AGG_SLOTS);
    if (err == cudaSuccess) err = cudaMemset(e->value_ptrs_dev, 0, sizeof(float*) * CF_MAX_LEVELS);
    if (err == cudaSuccess) err = cudaMemcpy(e->thr_dev, thr, sizeof(thr), cudaMemcpyHostToDevice);
    if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&e->hstream, cudaStreamNonBlocking);
    if (err == cudaSuccess) err = cudaEventCreateWithFlags(&e->order_ev, cudaEventDisableTiming);
    if (err == cudaSuccess) { launch_init(e->S, e->n_pad, e->dev, 0); e->launches += 1; err = cudaDeviceSynchronize(); }
    if (err != cudaSuccess) { cuda_fail(err, "cf_create"); cf_destroy(e); return err == cudaErrorMemoryAllocation ? CF_ERR_ALLOC : CF_ERR_CUDA; }
    e->tab.value = e->value_ptrs_dev; e->tab.signs = e->signs_slab; e->tab.thr = e->thr_dev;
    *out = e;
    return CF_OK;
}

int cf_destroy(cf_env* e) {
    if (!e) return CF_OK;
    cudaSetDevice(e->device);
    cudaDeviceSynchronize();
    for (int i = 0; i < CF_MAX_LEVELS; ++i) cudaFree(e->value_dev[i]);
    cudaFree(e->signs_slab); cudaFree((void*)e->value_ptrs_dev); cudaFree(e->thr_dev);
    for (int i = 0; i < 4; ++i) cudaFree(e->inj[i]);
    cudaFree(e->reset_list); cudaFree(e->reset_count); cudaFree(e->S); cudaFree(e->fin_ret); cudaFree(e->fin_len); cudaFree(e->agg); cudaFree(e->rec_dev);
    if (e->d_out_slab) cudaFree(e->d_out_slab);
    else { cudaFree(e->d_obs); cudaFree(e->d_rew); cudaFree(e->d_cost); cudaFree(e->d_done); }
    cudaFree(e->d_act); cudaFree(e->d_term); cudaFree(e->d_trunc); cudaFree(e->d_mask);
    for (EventPair& p : e->ev_pool) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    if (e->order_ev) cudaEventDestroy(e->order_ev);
    if (e->hstream) cudaStreamDestroy(e->hstream);
    delete e;
    return CF_OK;
}

int cf_obs_dim(const cf_env* e) { return e ? e->obs_dim : CF_ERR_INVALID; }
int cf_num_envs(const cf_env* e) { return e ? e->n : CF_ERR_INVALID; }
int64_t cf_launch_count(const cf_env* e) { return e ? e->launches : 0; }

int cf_load_hj_table(cf_env* e, int32_t li, const float* V_host) {
    if (!e || !V_host || li < 0 || li >= CF_MAX_LEVELS) return fail(CF_ERR_INVALID, "cf_load_hj_table: bad argument%s", "");
    if (e->tables_borrowed) return fail(CF_ERR_INVALID, "cf_load_hj_table: this handle borrows its tables (cf_share_hj_tables)%s", "");
    CU(cudaSetDevice(e->device));
    const size_t vbytes = sizeof(float) * (size_t)CF_HJ_CELLS;
    int rc = ensure(&e->signs_slab, (size_t)CF_SIGN_STRIDE * CF_MAX_LEVELS);
    if (rc != CF_OK) return rc;
    e->tab.signs = e->signs_slab;
    float* V = e->value_dev[li];
    if (!V) { cudaError_t err = cudaMalloc((void**)&V, vbytes); if (err != cudaSuccess) { cuda_fail(err, "cudaMalloc(value table)"); return CF_ERR_ALLOC; } }
    CU(cudaMemcpy(V, V_host, vbytes, cudaMemcpyHostToDevice));
    launch_hj_signs(V, e->signs_slab + (size_t)li * CF_SIGN_STRIDE, 0);
    e->launches += 1;
    CU(cudaDeviceSynchronize());
    if (e->cfg.hj_mode == CF_HJ_SIGN_TABLE) { cudaFree(V); V = nullptr; }   // 2.85 MB/level stays, 45.6 MB/level goes
    e->value_dev[li] = V;
    e->has_signs[li] = true;
    CU(cudaMemcpy((void*)(e->value_ptrs_dev + li), &V, sizeof(float*), cudaMemcpyHostToDevice));
    return CF_OK;
}

int cf_share_hj_tables(cf_env* e, const cf_env* owner) {
    if (!e || !owner || e == owner) return fail(CF_ERR_INVALID, "cf_share_hj_tables: bad argument%s", "");
    if (e->device != owner->device || e->cfg.hj_mode != owner->cfg.hj_mode)
        return fail(CF_ERR_INVALID, "cf_share_hj_tables: handles must be on the same device with the same hj_mode%s", "");
    if (owner->tables_borrowed) return fail(CF_ERR_INVALID, "cf_share_hj_tables: owner itself borrows its tables%s", "");
    CU(cudaSetDevice(e->device));
    // this handle's own (empty or loaded) tables are released; it reads the owner's from now on
    for (int i = 0; i < CF_MAX_LEVELS; ++i) { cudaFree(e->value_dev[i]); e->value_dev[i] = nullptr; e->has_signs[i] = false; }
    cudaFree(e->signs_slab); e->signs_slab = nullptr;
    e->tab.value = owner->value_ptrs_dev; e->tab.signs = owner->signs_slab;
    e->table_owner = owner;
    e->tables_borrowed = true;
    return CF_OK;
}

int cf_get_hj_signs(cf_env* e, int32_t li, uint8_t* out_host) {
    if (!e || !out_host || li < 0 || li >= CF_MAX_LEVELS) return fail(CF_ERR_INVALID, "cf_get_hj_signs: bad argument%s", "");
    const cf_env* o = e->tables_borrowed ? e->table_owner : e;
    if (!o->has_signs[li]) return fail(CF_ERR_NO_TABLE, "no table for level index %s%lld", "", li);
    CU(cudaSetDevice(e->device));
    CU(cudaMemcpy(out_host, o->signs_slab + (size_t)li * CF_SIGN_STRIDE, (CF_HJ_CELLS + 3) / 4, cudaMemcpyDeviceToHost));
    return CF_OK;
}

int cf_reset_dev(cf_env* e, const uint8_t* mask, float* obs, void* stream) {
    if (!e || !obs) return fail(CF_ERR_INVALID, "cf_reset: null argument%s", "");
    if (e->dev.injected && (!e->inj[2] || !e->inj[3])) return fail(CF_ERR_INVALID, "rng_mode INJECTED: call cf_set_noise_override first%s", "");
    CU(cudaSetDevice(e->device));
    order_families(e, (cudaStream_t)stream);
    launch_reset(e->S, e->n, e->dev, mask, obs, e->inj[2], e->inj[3], (cudaStream_t)stream);
    e->launches += 1;
    CU(cudaGetLastError());
    return CF_OK;
}

int cf_step_dev(cf_env* e, const float* act, float* obs, float* rew, uint8_t* done, float* cost, float* term_obs,
                uint8_t* truncated, void* stream) {
    return step_common(e, 1, act, obs, rew, done, cost, term_obs, truncated, (cudaStream_t)stream);
}

int cf_step_k_dev(cf_env* e, int32_t k, const float* act, float* obs, float* rew, uint8_t* done, float* cost, void* stream) {
    if (e && e->dev.injected && k != 1) return fail(CF_ERR_INVALID, "cf_step_k: injected noise supports k = 1 only%s", "");
    return step_common(e, k, act, obs, rew, done, cost, nullptr, nullptr, (cudaStream_t)stream);
}

int cf_reset_host(cf_env* e, const uint8_t* mask, float* obs) {
    if (!e || !obs) return fail(CF_ERR_INVALID, "cf_reset: null argument%s", "");
    CU(cudaSetDevice(e->device));
    const size_t n = e->n, od = e->obs_dim;
    int rc = ensure_outputs(e);
    if (rc == CF_OK && mask) rc = ensure(&e->d_mask, n);
    if (rc != CF_OK) return rc;
    cudaStream_t st = e->hstream;
    if (mask) {
        CU(cudaMemcpyAsync(e->d_mask, mask, n, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(e->d_obs, obs, n * od * sizeof(float), cudaMemcpyHostToDevice, st));   // keep unmasked rows
    }
    rc = cf_reset_dev(e, mask ? e->d_mask : nullptr, e->d_obs, st);
    if (rc != CF_OK) return rc;
    CU(cudaMemcpyAsync(obs, e->d_obs, n * od * sizeof(float), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return CF_OK;
}

// tools/ only: CF_PROBE_E2E_TIMELINE=<csv path> records CUDA events around the upload, the kernels, the observation download and
// the last small download of every host-pointer step against one process-wide origin and writes them at exit (no nsys here).
struct HostTimeline {
    const char* path = getenv("CF_PROBE_E2E_TIMELINE");
    cudaEvent_t origin = nullptr;
    std::vector<float> rows;                                   // 10 floats per step: handle tag, t0..t4 [ms since origin], 4 CPU times
    double cpu0 = 0.0;
    static double now_ms() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return 1e3 * ts.tv_sec + 1e-6 * ts.tv_nsec; }
    ~HostTimeline() {
        if (!path || rows.empty()) return;
        if (FILE* f = fopen(path, "w")) {
            fprintf(f, "handle,before_h2d,after_h2d,after_kernels,after_obs_d2h,after_small_d2h,cpu_enq_in,cpu_enq_out,cpu_sync_in,cpu_sync_out\n");
            for (size_t k = 0; k + 10 <= rows.size(); k += 10)
                fprintf(f, "%d,%.4f,%.4f,%.4f,%.4f,%.4f,%.4f,%.4f,%.4f,%.4f\n", (int)rows[k], rows[k + 1], rows[k + 2], rows[k + 3], rows[k + 4],
                        rows[k + 5], rows[k + 6], rows[k + 7], rows[k + 8], rows[k + 9]);
            fclose(f);
        }
    }
};
static HostTimeline g_timeline;

static int step_host_enqueue(cf_env* e, const float* act, float* obs, float* rew, uint8_t* done, float* cost, float* term_obs,
                             uint8_t* truncated, bool wait) {
    if (!e || !act || !obs || !rew || !done) return fail(CF_ERR_INVALID, "cf_step: null argument%s", "");
    CU(cudaSetDevice(e->device));
    const size_t n = e->n, od = e->obs_dim;
    int rc = ensure(&e->d_act, n * 4);
    if (rc == CF_OK) rc = ensure_outputs(e);
    if (rc == CF_OK && term_obs) rc = ensure(&e->d_term, n * od);
    if (rc == CF_OK && truncated) rc = ensure(&e->d_trunc, n);
    if (rc != CF_OK) return rc;
    cudaStream_t st = e->hstream;
    const bool tl = g_timeline.path != nullptr;
    if (tl) {
        if (!g_timeline.origin) { CU(cudaEventCreate(&g_timeline.origin)); CU(cudaEventRecord(g_timeline.origin, st)); g_timeline.cpu0 = HostTimeline::now_ms(); }
        e->probe_cpu[0] = HostTimeline::now_ms() - g_timeline.cpu0;
        for (int k = 0; k < 5; ++k) if (!e->probe_ev[k]) CU(cudaEventCreate(&e->probe_ev[k]));
        CU(cudaEventRecord(e->probe_ev[0], st));
    }
    // pinned caller buffers make these true async DMA transfers; pageable ones are staged by the driver.  (An upload by kernel
    // from the pinned buffer was measured against this copy and is never faster: profiles/r2_e2e_timeline.md.)
    CU(cudaMemcpyAsync(e->d_act, act, n * 4 * sizeof(float), cudaMemcpyHostToDevice, st));
    if (tl) CU(cudaEventRecord(e->probe_ev[1], st));
    if (term_obs) CU(cudaMemcpyAsync(e->d_term, term_obs, n * od * sizeof(float), cudaMemcpyHostToDevice, st));
    rc = step_common(e, 1, e->d_act, e->d_obs, e->d_rew, e->d_done, cost ? e->d_cost : nullptr,
                     term_obs ? e->d_term : nullptr, truncated ? e->d_trunc : nullptr, st);
    if (rc != CF_OK) return rc;
    if (tl) CU(cudaEventRecord(e->probe_ev[2], st));
    // obs | rew | cost | done adjacent in host memory, in the order of the device slab: one download instead of four (each small copy
    // costs ~4 us of fixed time on the copy engine that paces this path: profiles/r2_e2e_timeline.md)
    static const bool no_pack = getenv("CF_HOST_NO_PACK") != nullptr;     // tools/ only: A/B
    const uintptr_t ho = (uintptr_t)obs, hr = (uintptr_t)rew, hc = (uintptr_t)cost, hd = (uintptr_t)done;   // addresses: the arrays may be unrelated objects
    const bool packed = !no_pack && e->d_out_slab && cost && hr == ho + n * od * sizeof(float) && hc == hr + n * sizeof(float) &&
                        hd == hc + n * sizeof(float);
    if (packed) {
        CU(cudaMemcpyAsync(obs, e->d_obs, n * (od * sizeof(float) + 2 * sizeof(float) + 1), cudaMemcpyDeviceToHost, st));
        if (tl) CU(cudaEventRecord(e->probe_ev[3], st));
    } else {
        CU(cudaMemcpyAsync(obs, e->d_obs, n * od * sizeof(float), cudaMemcpyDeviceToHost, st));
        if (tl) CU(cudaEventRecord(e->probe_ev[3], st));
        CU(cudaMemcpyAsync(rew, e->d_rew, n * sizeof(float), cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(done, e->d_done, n, cudaMemcpyDeviceToHost, st));
        if (cost) CU(cudaMemcpyAsync(cost, e->d_cost, n * sizeof(float), cudaMemcpyDeviceToHost, st));
    }
    if (term_obs) CU(cudaMemcpyAsync(term_obs, e->d_term, n * od * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (truncated) CU(cudaMemcpyAsync(truncated, e->d_trunc, n, cudaMemcpyDeviceToHost, st));
    if (tl) { CU(cudaEventRecord(e->probe_ev[4], st)); e->probe_cpu[1] = HostTimeline::now_ms() - g_timeline.cpu0; }
    if (wait) CU(cudaStreamSynchronize(st));
    return CF_OK;
}

int cf_step_host(cf_env* e, const float* act, float* obs, float* rew, uint8_t* done, float* cost, float* term_obs,
                 uint8_t* truncated) {
    return step_host_enqueue(e, act, obs, rew, done, cost, term_obs, truncated, true);
}
int cf_step_host_async(cf_env* e, const float* act, float* obs, float* rew, uint8_t* done, float* cost, float* term_obs,
                       uint8_t* truncated) {
    return step_host_enqueue(e, act, obs, rew, done, cost, term_obs, truncated, false);
}
int cf_host_sync(cf_env* e) {
    if (!e) return fail(CF_ERR_INVALID, "cf_host_sync: null handle%s", "");
    CU(cudaSetDevice(e->device));
    if (g_timeline.path) e->probe_cpu[2] = HostTimeline::now_ms() - g_timeline.cpu0;
    CU(cudaStreamSynchronize(e->hstream));
    if (g_timeline.path && e->probe_ev[4] && g_timeline.rows.size() < 10 * 4096) {
        const double out = HostTimeline::now_ms() - g_timeline.cpu0;
        float t[5];
        for (int k = 0; k < 5; ++k) CU(cudaEventElapsedTime(&t[k], g_timeline.origin, e->probe_ev[k]));
        g_timeline.rows.push_back((float)((uintptr_t)e >> 4 & 0xffff));
        for (int k = 0; k < 5; ++k) g_timeline.rows.push_back(t[k]);
        for (int k = 0; k < 3; ++k) g_timeline.rows.push_back((float)e->probe_cpu[k]);
        g_timeline.rows.push_back((float)out);
    }
    return CF_OK;
}

int cf_get_state(cf_env* e, float* state_host) {
    if (!e || !state_host) return fail(CF_ERR_INVALID, "cf_get_state: null argument%s", "");
    CU(cudaSetDevice(e->device));
    int rc = ensure(&e->rec_dev, (size_t)CF_STATE_WORDS * e->n);
    if (rc != CF_OK) return rc;
    CU(cudaDeviceSynchronize());
    launch_unpack_state(e->S, e->n, e->rec_dev, 0);
    e->launches += 1;
    CU(cudaMemcpy(state_host, e->rec_dev, sizeof(float) * CF_STATE_WORDS * (size_t)e->n, cudaMemcpyDeviceToHost));
    return CF_OK;
}

int cf_set_state(cf_env* e, const float* state_host) {
    if (!e || !state_host) return fail(CF_ERR_INVALID, "cf_set_state: null argument%s", "");
    CU(cudaSetDevice(e->device));
    int rc = ensure(&e->rec_dev, (size_t)CF_STATE_WORDS * e->n);
    if (rc != CF_OK) return rc;
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(e->rec_dev, state_host, sizeof(float) * CF_STATE_WORDS * (size_t)e->n, cudaMemcpyHostToDevice));
    launch_pack_state(e->S, e->n, e->rec_dev, 0);
    e->launches += 1;
    CU(cudaDeviceSynchronize());
    return CF_OK;
}

int cf_debug_draws(cf_env* e, int32_t first_env, int32_t count, uint32_t ctr, uint32_t tag, int32_t nzb, int32_t nub,
                   uint32_t* words_host, float* draws_host) {
    if (!e || !words_host || !draws_host || count < 1 || first_env < 0 || nzb < 0 || nub < 0 || nzb + nub < 1 || nzb > 16)
        return fail(CF_ERR_INVALID, "cf_debug_draws: bad argument%s", "");
    CU(cudaSetDevice(e->device));
    const size_t words = (size_t)count * (nzb + nub) * 4;
    uint32_t* dw = nullptr; float* dd = nullptr;
    cudaError_t err = cudaMalloc((void**)&dw, words * 4);
    if (err == cudaSuccess) err = cudaMalloc((void**)&dd, words * 4);
    if (err == cudaSuccess) {
        launch_debug_draws(e->dev, first_env, count, ctr, tag, nzb, nub, dw, dd, 0);
        e->launches += 1;
        err = cudaMemcpy(words_host, dw, words * 4, cudaMemcpyDeviceToHost);
    }
    if (err == cudaSuccess) err = cudaMemcpy(draws_host, dd, words * 4, cudaMemcpyDeviceToHost);
    cudaFree(dw); cudaFree(dd);
    if (err != cudaSuccess) return cuda_fail(err, "cf_debug_draws");
    return CF_OK;
}

int cf_debug_box_muller(int32_t device, const uint32_t* words_host, int64_t n, float* normals_host, float* u01_host) {
    if (!words_host || !normals_host || !u01_host || n < 1) return fail(CF_ERR_INVALID, "cf_debug_box_muller: bad argument%s", "");
    CU(cudaSetDevice(device));
    const size_t bytes = (size_t)n * 16;
    uint32_t* dw = nullptr; float *dz = nullptr, *du = nullptr;
    cudaError_t err = cudaMalloc((void**)&dw, bytes);
    if (err == cudaSuccess) err = cudaMalloc((void**)&dz, bytes);
    if (err == cudaSuccess) err = cudaMalloc((void**)&du, bytes);
    if (err == cudaSuccess) err = cudaMemcpy(dw, words_host, bytes, cudaMemcpyHostToDevice);
    if (err == cudaSuccess) {
        launch_debug_box_muller(dw, n, dz, du, 0);
        err = cudaMemcpy(normals_host, dz, bytes, cudaMemcpyDeviceToHost);
    }
    if (err == cudaSuccess) err = cudaMemcpy(u01_host, du, bytes, cudaMemcpyDeviceToHost);
    cudaFree(dw); cudaFree(dz); cudaFree(du);
    if (err != cudaSuccess) return cuda_fail(err, "cf_debug_box_muller");
    return CF_OK;
}

int cf_debug_euler(int32_t device, const float* quat_host, int64_t n, float* rpy_bullet_host, float* rpy_hj_host) {
    if (!quat_host || !rpy_bullet_host || !rpy_hj_host || n < 1) return fail(CF_ERR_INVALID, "cf_debug_euler: bad argument%s", "");
    CU(cudaSetDevice(device));
    float *dq = nullptr, *da = nullptr, *db = nullptr;
    cudaError_t err = cudaMalloc((void**)&dq, (size_t)n * 16);
    if (err == cudaSuccess) err = cudaMalloc((void**)&da, (size_t)n * 12);
    if (err == cudaSuccess) err = cudaMalloc((void**)&db, (size_t)n * 12);
    if (err == cudaSuccess) err = cudaMemcpy(dq, quat_host, (size_t)n * 16, cudaMemcpyHostToDevice);
    if (err == cudaSuccess) {
        launch_debug_euler(dq, n, da, db, 0);
        err = cudaMemcpy(rpy_bullet_host, da, (size_t)n * 12, cudaMemcpyDeviceToHost);
    }
    if (err == cudaSuccess) err = cudaMemcpy(rpy_hj_host, db, (size_t)n * 12, cudaMemcpyDeviceToHost);
    cudaFree(dq); cudaFree(da); cudaFree(db);
    if (err != cudaSuccess) return cuda_fail(err, "cf_debug_euler");
    return CF_OK;
}

int cf_get_poses_dev(cf_env* e, const int32_t* env_idx_dev, int32_t count, float* poses_dev, void* stream) {
    if (!e || !poses_dev || count < 0) return fail(CF_ERR_INVALID, "cf_get_poses_dev: bad argument%s", "");
    if (count == 0) return CF_OK;
    if (!env_idx_dev && count > e->n) return fail(CF_ERR_INVALID, "cf_get_poses_dev: count exceeds n_envs%s", "");
    CU(cudaSetDevice(e->device));
    launch_gather_poses(e->S, e->n, env_idx_dev, count, poses_dev, (cudaStream_t)stream);
    e->launches += 1;
    CU(cudaGetLastError());
    return CF_OK;
}

int cf_get_levels(cf_env* e, float* levels_host) {
    if (!e || !levels_host) return fail(CF_ERR_INVALID, "cf_get_levels: null argument%s", "");
    CU(cudaSetDevice(e->device));
    int rc = ensure_outputs(e);
    if (rc != CF_OK) return rc;
    CU(cudaDeviceSynchronize());
    launch_levels(e->S, e->n, e->d_rew, 0);
    e->launches += 1;
    CU(cudaMemcpy(levels_host, e->d_rew, sizeof(float) * (size_t)e->n, cudaMemcpyDeviceToHost));
    return CF_OK;
}

int cf_set_noise_override(cf_env* e, const float* sz, const float* su, const float* rz, const float* ru) {
    if (!e) return fail(CF_ERR_INVALID, "cf_set_noise_override: null handle%s", "");
    CU(cudaSetDevice(e->device));
    const float* src[4] = {sz, su, rz, ru};
    const size_t slots[4] = {CF_STEP_NORMALS, CF_STEP_UNIFORMS, CF_RESET_NORMALS, CF_RESET_UNIFORMS};
    CU(cudaDeviceSynchronize());
    for (int i = 0; i < 4; ++i) {
        if (!src[i]) continue;
        int rc = ensure(&e->inj[i], slots[i] * (size_t)e->n);
        if (rc != CF_OK) return rc;
        CU(cudaMemcpy(e->inj[i], src[i], sizeof(float) * slots[i] * (size_t)e->n, cudaMemcpyHostToDevice));
    }
    return CF_OK;
}

static int hj_lookup_impl(cf_env* e, int32_t li, const void* states, bool f32, int32_t n, int32_t* idx, float* dstb_sign) {
    if (!e || !states || !idx || !dstb_sign || n < 1 || li < 0 || li >= CF_MAX_LEVELS)
        return fail(CF_ERR_INVALID, "cf_hj_lookup: bad argument%s", "");
    const cf_env* o = e->tables_borrowed ? e->table_owner : e;
    const bool use_value = e->cfg.hj_mode == CF_HJ_VALUE_TABLE;
    if (use_value ? !o->value_dev[li] : !o->has_signs[li]) return fail(CF_ERR_NO_TABLE, "no table for level index %s%lld", "", li);
    CU(cudaSetDevice(e->device));
    const size_t sbytes = (f32 ? sizeof(float) : sizeof(double)) * 6 * (size_t)n;
    void* ds = nullptr; int* di = nullptr; float* dd = nullptr;
    cudaError_t err = cudaMalloc(&ds, sbytes);
    if (err == cudaSuccess) err = cudaMalloc((void**)&di, sizeof(int) * 6 * (size_t)n);
    if (err == cudaSuccess) err = cudaMalloc((void**)&dd, sizeof(float) * 2 * (size_t)n);
    if (err == cudaSuccess) err = cudaMemcpy(ds, states, sbytes, cudaMemcpyHostToDevice);
    if (err == cudaSuccess) {
        const float* V = use_value ? o->value_dev[li] : nullptr;
        const uint8_t* P = use_value ? nullptr : o->signs_slab + (size_t)li * CF_SIGN_STRIDE;
        if (f32) launch_hj_lookup_f32((const float*)ds, n, e->dev, V, P, e->thr_dev, di, dd, 0);
        else launch_hj_lookup((const double*)ds, n, V, P, di, dd, 0);
        e->launches += 1;
        err = cudaMemcpy(idx, di, sizeof(int) * 6 * (size_t)n, cudaMemcpyDeviceToHost);
    }
    if (err == cudaSuccess) err = cudaMemcpy(dstb_sign, dd, sizeof(float) * 2 * (size_t)n, cudaMemcpyDeviceToHost);
    cudaFree(ds); cudaFree(di); cudaFree(dd);
    if (err != cudaSuccess) return cuda_fail(err, "cf_hj_lookup");
    return CF_OK;
}

int cf_hj_lookup_host(cf_env* e, int32_t li, const double* states, int32_t n, int32_t* idx, float* dstb_sign) {
    return hj_lookup_impl(e, li, states, false, n, idx, dstb_sign);
}
int cf_hj_lookup_f32_host(cf_env* e, int32_t li, const float* states, int32_t n, int32_t* idx, float* dstb_sign) {
    return hj_lookup_impl(e, li, states, true, n, idx, dstb_sign);
}

int cf_episode_stats(cf_env* e, int32_t reset, double out[4], float* fin_ret_host, int32_t* fin_len_host) {
    if (!e || !out) return fail(CF_ERR_INVALID, "cf_episode_stats: null argument%s", "");
    CU(cudaSetDevice(e->device));
    CU(cudaDeviceSynchronize());
    double slots[4 * CF_AGG_SLOTS];
    CU(cudaMemcpy(slots, e->agg, sizeof(slots), cudaMemcpyDeviceToHost));
    for (int k = 0; k < 4; ++k) { out[k] = 0.0; for (int j = 0; j < CF_AGG_SLOTS; ++j) out[k] += slots[4 * j + k]; }
    if (fin_ret_host) CU(cudaMemcpy(fin_ret_host, e->fin_ret, sizeof(float) * (size_t)e->n, cudaMemcpyDeviceToHost));
    if (fin_len_host) {
        CU(cudaMemcpy(fin_len_host, e->fin_len, sizeof(int) * (size_t)e->n, cudaMemcpyDeviceToHost));
        for (int i = 0; i < e->n; ++i) fin_len_host[i] &= 0xffff;         // the device word also carries the episode's level index
    }
    if (reset) CU(cudaMemset(e->agg, 0, sizeof(double) * 4 * CF_AGG_SLOTS));
    return CF_OK;
}

int cf_last_episodes(cf_env* e, float* ret_host, int32_t* len_host, float* level_host) {
    if (!e || !ret_host || !len_host || !level_host) return fail(CF_ERR_INVALID, "cf_last_episodes: null argument%s", "");
    CU(cudaSetDevice(e->device));
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(ret_host, e->fin_ret, sizeof(float) * (size_t)e->n, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(len_host, e->fin_len, sizeof(int) * (size_t)e->n, cudaMemcpyDeviceToHost));
    for (int i = 0; i < e->n; ++i) { level_host[i] = (float)(len_host[i] >> 16) / 10.0f; len_host[i] &= 0xffff; }
    return CF_OK;
}

int cf_gae_dev(const float* rew, const float* val, const uint8_t* done, const float* end_val, const float* last_val,
               float gamma, float lam, float rew_scale, int32_t T, int32_t N, float* adv, float* target_v,
               float* disc_ret, void* stream) {
    if (!rew || !val || !done || !last_val || !adv || !target_v || !disc_ret || T < 1 || N < 1)
        return fail(CF_ERR_INVALID, "cf_gae: bad argument%s", "");
    launch_gae(rew, val, done, end_val, last_val, gamma, lam, rew_scale, T, N, adv, target_v, disc_ret, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return CF_OK;
}

int cf_policy_step_dev(cf_env* e, const cf_policy_weights* w, const float* obs, int32_t deterministic, float* act, float* val,
                       float* logp, void* stream) {
    if (!e || !w || !obs || !act || !val || !logp) return fail(CF_ERR_INVALID, "cf_policy_step: null argument%s", "");
    if (w->pi_hidden < 1 || w->pi_hidden > 64 || w->v_hidden < 1 || w->v_hidden > 64)
        return fail(CF_ERR_INVALID, "cf_policy_step: hidden widths must be 1..64%s", "");
    for (int l = 0; l < 3; ++l)
        if (!w->pi_w[l] || !w->pi_b[l] || !w->v_w[l] || !w->v_b[l]) return fail(CF_ERR_INVALID, "cf_policy_step: null weight pointer%s", "");
    if (!deterministic && !w->log_std) return fail(CF_ERR_INVALID, "cf_policy_step: log_std needed for sampling%s", "");
    if (((uintptr_t)obs & 7) || ((uintptr_t)act & 15)) return fail(CF_ERR_INVALID, "cf_policy_step: obs must be 8-byte, act 16-byte aligned%s", "");
    CU(cudaSetDevice(e->device));
    order_families(e, (cudaStream_t)stream);                 // reads the event counters the step kernels write
    PolicyHostArgs h;
    h.obs = obs; h.obs_mean = w->obs_mean; h.obs_std = w->obs_std;
    for (int l = 0; l < 3; ++l) { h.pi_w[l] = w->pi_w[l]; h.pi_b[l] = w->pi_b[l]; h.v_w[l] = w->v_w[l]; h.v_b[l] = w->v_b[l]; }
    h.log_std = w->log_std; h.S = e->S; h.act = act; h.val = val; h.logp = logp;
    h.n = e->n; h.obs_dim = e->obs_dim; h.pi_h = w->pi_hidden; h.v_h = w->v_hidden; h.deterministic = deterministic ? 1 : 0;
    h.eps = w->obs_eps;
    cudaError_t err = launch_policy(h, e->dev, (cudaStream_t)stream);
    if (err != cudaSuccess) return cuda_fail(err, "launch_policy");
    e->launches += 1;
    CU(cudaGetLastError());
    return CF_OK;
}

static const int kPpoCtas = 296;                             // persistent CTAs of the fused update kernel (2 per SM on 148 SMs)
int cf_ppo_workspace_floats(void) { return kPpoCtas * ppo_partial_floats(); }

int cf_ppo_pi_grad_dev(const cf_policy_weights* w, int32_t obs_dim, const float* obs, const float* act, const float* adv,
                       const float* logp_old, int64_t B, float clip_ratio, float* const grad_w[3], float* const grad_b[3],
                       float* loss_ratio, float* workspace, void* stream) {
    if (!w || !obs || !act || !adv || !logp_old || !grad_w || !grad_b || !loss_ratio || !workspace || B < 1 || B > 0x7fffffff)
        return fail(CF_ERR_INVALID, "cf_ppo_pi_grad: bad argument%s", "");
    if ((obs_dim != 34 && obs_dim != 42) || w->pi_hidden < 1 || w->pi_hidden > 64 || !w->log_std)
        return fail(CF_ERR_INVALID, "cf_ppo_pi_grad: obs_dim must be 34 or 42, hidden width 1..64%s", "");
    if (((uintptr_t)obs & 7) || ((uintptr_t)act & 15)) return fail(CF_ERR_INVALID, "cf_ppo_pi_grad: obs must be 8-byte, act 16-byte aligned%s", "");
    PpoHostArgs a{};
    a.net = 0; a.obs = obs; a.act = act; a.adv = adv; a.logp_old = logp_old; a.log_std = w->log_std;
    for (int l = 0; l < 3; ++l) {
        if (!w->pi_w[l] || !w->pi_b[l] || !grad_w[l] || !grad_b[l]) return fail(CF_ERR_INVALID, "cf_ppo_pi_grad: null tensor%s", "");
        a.w[l] = w->pi_w[l]; a.b[l] = w->pi_b[l]; a.gw[l] = grad_w[l]; a.gb[l] = grad_b[l];
    }
    a.loss_ratio = loss_ratio; a.partial = workspace; a.B = (int)B; a.obs_dim = obs_dim; a.h = w->pi_hidden; a.n_cta = kPpoCtas;
    a.clip = clip_ratio;
    cudaError_t err = launch_ppo_grad(a, (cudaStream_t)stream);
    if (err != cudaSuccess) return cuda_fail(err, "launch_ppo_grad");
    CU(cudaGetLastError());
    return CF_OK;
}

// The allocation-free variant of cf_ppo_pi_grad_dev: no workspace argument, the per-CTA partial sums go to a file-scope
// __device__ array of cf_ppo.cu (address resolved there with cudaGetSymbolAddress, once per device).  Calls nothing that
// allocates: no cudaMalloc / cudaFree, no library routine; the error text goes to a thread-local buffer.
int64_t cf_ppo_static_max_samples(void) { return (int64_t)ppo_static_max_samples(); }

int cf_ppo_pi_grad_static_dev(const cf_policy_weights* w, int32_t obs_dim, const float* obs, const float* act, const float* adv,
                              const float* logp_old, int64_t B, float clip_ratio, float* const grad_w[3], float* const grad_b[3],
                              float* loss_ratio, void* stream) {
    if (!w || !obs || !act || !adv || !logp_old || !grad_w || !grad_b || !loss_ratio || B < 1)
        return fail(CF_ERR_INVALID, "cf_ppo_pi_grad_static: bad argument%s", "");
    if (B > ppo_static_max_samples())
        return fail(CF_ERR_INVALID, "cf_ppo_pi_grad_static: %s%lld samples exceed the static scratch's limit of 4194304 "
                                    "(cf_ppo_static_max_samples); use cf_ppo_pi_grad_dev with a caller workspace", "", (long long)B);
    if ((obs_dim != 34 && obs_dim != 42) || w->pi_hidden < 1 || w->pi_hidden > 64 || !w->log_std)
        return fail(CF_ERR_INVALID, "cf_ppo_pi_grad_static: obs_dim must be 34 or 42, hidden width 1..64%s", "");
    if (((uintptr_t)obs & 7) || ((uintptr_t)act & 15))
        return fail(CF_ERR_INVALID, "cf_ppo_pi_grad_static: obs must be 8-byte, act 16-byte aligned%s", "");
    PpoHostArgs a{};
    a.net = 0; a.obs = obs; a.act = act; a.adv = adv; a.logp_old = logp_old; a.log_std = w->log_std;
    for (int l = 0; l < 3; ++l) {
        if (!w->pi_w[l] || !w->pi_b[l] || !grad_w[l] || !grad_b[l]) return fail(CF_ERR_INVALID, "cf_ppo_pi_grad_static: null tensor%s", "");
        a.w[l] = w->pi_w[l]; a.b[l] = w->pi_b[l]; a.gw[l] = grad_w[l]; a.gb[l] = grad_b[l];
    }
    a.loss_ratio = loss_ratio; a.B = (int)B; a.obs_dim = obs_dim; a.h = w->pi_hidden; a.clip = clip_ratio;
    cudaError_t err = launch_ppo_pi_grad_static(a, (cudaStream_t)stream);
    if (err != cudaSuccess) return cuda_fail(err, "launch_ppo_pi_grad_static (cudaGetSymbolAddress / launch)");
    CU(cudaGetLastError());
    return CF_OK;
}

int cf_ppo_v_grad_dev(const cf_policy_weights* w, int32_t obs_dim, const float* obs, const float* target_v, const int64_t* idx,
                      int64_t B, float* const grad_w[3], float* const grad_b[3], float* loss, float* workspace, void* stream) {
    if (!w || !obs || !target_v || !grad_w || !grad_b || !loss || !workspace || B < 1 || B > 0x7fffffff)
        return fail(CF_ERR_INVALID, "cf_ppo_v_grad: bad argument%s", "");
    if ((obs_dim != 34 && obs_dim != 42) || w->v_hidden < 1 || w->v_hidden > 64)
        return fail(CF_ERR_INVALID, "cf_ppo_v_grad: obs_dim must be 34 or 42, hidden width 1..64%s", "");
    if ((uintptr_t)obs & 7) return fail(CF_ERR_INVALID, "cf_ppo_v_grad: obs must be 8-byte aligned%s", "");
    PpoHostArgs a{};
    a.net = 1; a.obs = obs; a.tgt = target_v; a.idx = (const long long*)idx;
    for (int l = 0; l < 3; ++l) {
        if (!w->v_w[l] || !w->v_b[l] || !grad_w[l] || !grad_b[l]) return fail(CF_ERR_INVALID, "cf_ppo_v_grad: null tensor%s", "");
        a.w[l] = w->v_w[l]; a.b[l] = w->v_b[l]; a.gw[l] = grad_w[l]; a.gb[l] = grad_b[l];
    }
    a.loss_ratio = loss; a.partial = workspace; a.B = (int)B; a.obs_dim = obs_dim; a.h = w->v_hidden; a.n_cta = kPpoCtas;
    cudaError_t err = launch_ppo_grad(a, (cudaStream_t)stream);
    if (err != cudaSuccess) return cuda_fail(err, "launch_ppo_grad");
    CU(cudaGetLastError());
    return CF_OK;
}

int cf_pd_actions_dev(const float* obs, int32_t obs_dim, int32_t n, float* act, const float gains[6], void* stream) {
    if (!obs || !act || !gains || n < 1 || (obs_dim != 34 && obs_dim != 42) || ((uintptr_t)act & 15))
        return fail(CF_ERR_INVALID, "cf_pd_actions: bad argument%s", "");
    const PdGains g{gains[0], gains[1], gains[2], gains[3], gains[4], gains[5]};
    launch_pd_actions(obs, obs_dim, obs_dim == 34 ? 17 : 21, n, act, g, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return CF_OK;
}

int cf_gae_host(const float* rew, const float* val, const uint8_t* done, const float* end_val, const float* last_val,
                float gamma, float lam, float rew_scale, int32_t T, int32_t N, float* adv, float* target_v,
                float* disc_ret) {
    if (!rew || !val || !done || !last_val || !adv || !target_v || !disc_ret || T < 1 || N < 1)
        return fail(CF_ERR_INVALID, "cf_gae: bad argument%s", "");
    const size_t tn = (size_t)T * N;
    float *d_f = nullptr; uint8_t* d_d = nullptr;
    // one slab: rew, val, end_val, adv, tgt, ret [T*N each] + last_val [N]
    cudaError_t err = cudaMalloc((void**)&d_f, sizeof(float) * (6 * tn + N));
    if (err == cudaSuccess) err = cudaMalloc((void**)&d_d, tn);
    float *d_rew = d_f, *d_val = d_f + tn, *d_end = d_f + 2 * tn, *d_adv = d_f + 3 * tn, *d_tgt = d_f + 4 * tn,
          *d_ret = d_f + 5 * tn, *d_last = d_f + 6 * tn;
    if (err == cudaSuccess) err = cudaMemcpy(d_rew, rew, sizeof(float) * tn, cudaMemcpyHostToDevice);
    if (err == cudaSuccess) err = cudaMemcpy(d_val, val, sizeof(float) * tn, cudaMemcpyHostToDevice);
    if (err == cudaSuccess && end_val) err = cudaMemcpy(d_end, end_val, sizeof(float) * tn, cudaMemcpyHostToDevice);
    if (err == cudaSuccess) err = cudaMemcpy(d_last, last_val, sizeof(float) * N, cudaMemcpyHostToDevice);
    if (err == cudaSuccess) err = cudaMemcpy(d_d, done, tn, cudaMemcpyHostToDevice);
    if (err == cudaSuccess) {
        launch_gae(d_rew, d_val, d_d, end_val ? d_end : nullptr, d_last, gamma, lam, rew_scale, T, N, d_adv, d_tgt, d_ret, 0);
        err = cudaMemcpy(adv, d_adv, sizeof(float) * tn, cudaMemcpyDeviceToHost);
    }
    if (err == cudaSuccess) err = cudaMemcpy(target_v, d_tgt, sizeof(float) * tn, cudaMemcpyDeviceToHost);
    if (err == cudaSuccess) err = cudaMemcpy(disc_ret, d_ret, sizeof(float) * tn, cudaMemcpyDeviceToHost);
    cudaFree(d_f); cudaFree(d_d);
    if (err != cudaSuccess) return cuda_fail(err, "cf_gae_host");
    return CF_OK;
}

int cf_set_prefetch_next(cf_env* e, const cf_env* next, const float* next_act) {
    if (!e) return fail(CF_ERR_INVALID, "cf_set_prefetch_next: null handle%s", "");
    if (next && next->device != e->device) return fail(CF_ERR_INVALID, "cf_set_prefetch_next: handles must be on the same device%s", "");
    if (next_act && (uintptr_t)next_act % 16 != 0) return fail(CF_ERR_INVALID, "cf_set_prefetch_next: next_act must be 16-byte aligned%s", "");
    e->pf_next = next;
    e->pf_act = next ? next_act : nullptr;
    return CF_OK;
}

int cf_set_trace_buffer(cf_env* e, uint64_t* trace_dev) {
    if (!e) return fail(CF_ERR_INVALID, "cf_set_trace_buffer: null handle%s", "");
    e->trace = (unsigned long long*)trace_dev;
    return CF_OK;
}
int cf_num_tiles(const cf_env* e) { return e ? e->n_pad / CF_TILE : CF_ERR_INVALID; }

int cf_kernel_timing(cf_env* e, int32_t enable, int32_t reset, double* avg_ms, int64_t* launches) {
    if (!e) return fail(CF_ERR_INVALID, "cf_kernel_timing: null handle%s", "");
    CU(cudaSetDevice(e->device));
    drain_timing(e);
    if (avg_ms) *avg_ms = e->ev_count ? e->ev_ms / (double)e->ev_count : 0.0;
    if (launches) *launches = e->ev_count;
    if (reset) { e->ev_ms = 0.0; e->ev_count = 0; }
    e->timing = enable != 0;
    return CF_OK;
}

void* cf_host_alloc(uint64_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) { cuda_fail(cudaGetLastError(), "cudaHostAlloc"); return nullptr; }
    return p;
}
void cf_host_free(void* p) { if (p) cudaFreeHost(p); }

}  // extern "C"
