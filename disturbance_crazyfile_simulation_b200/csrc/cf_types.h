// cf_types.h - plain structs shared by the host side (cf_abi.cu) and the kernels (cf_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/cf_abi.h"

namespace cf {

constexpr int kGroups = CF_STATE_WORDS / 4;   // 19 float4 groups per env

// Everything the kernels need from cf_config, pre-derived in float64 on the host (cf_abi.cu).
struct DevCfg {
    int reward_kind, dstb_kind, obs_noise, reset_dist, agg_steps, max_ep_steps;
    int gyroscopic, ground_effect, hj_mode, auto_reset, dr_on, injected, fixed_level_idx, obs_dim;
    float dr;
    float pen_action, pen_angle, pen_spin, pen_terminal, pen_velocity, pen_z;
    float done_rp_rad, done_rate_deg;       // compared as rad2deg(|w|) > done_rate_deg
    float init_pos, init_angle, init_yaw, init_vel, init_rate, init_yaw_rate;
    float cost_rate_limit;
    float sim_dt, motor_tc, ou_sigma, lin_damp, ang_damp;
    float nomA, nomB, nomK;                 // agents.py:200-204
    float hover_x, hover_a;                 // agents.py:152-153
    float gyro_sb, gyro_pi, gyro_rw, gyro_ton;   // sensors.py:121-134 constants at dt = 1/200
    float quat_std, quat_unif;
    float gnd_hclip;
    uint32_t seed_lo, seed_hi;
    long long gid_base;
};

// Device-side tables owned by the handle.
struct DevTables {
    const float* value[CF_MAX_LEVELS];      // fp32 (15,)*6 value tables (CF_HJ_VALUE_TABLE) or null
    const uint8_t* signs[CF_MAX_LEVELS];    // packed 2-bit sign tables, 4 cells / byte, or null
};

// Caller-owned I/O of one step launch plus the handle-owned side buffers.
struct StepIO {
    const float* act;      // [k][n][4]
    float* obs;            // [k][n][obs_dim]
    float* rew;            // [k][n]
    uint8_t* done;         // [k][n]
    float* cost;           // [k][n] or null
    float* term_obs;       // [n][obs_dim] or null (single-step only)
    uint8_t* truncated;    // [n] or null (single-step only)
    const float* inj_step_z;   // parity mode: [44][n], [8][n], [48][n], [42][n]
    const float* inj_step_u;
    const float* inj_reset_z;
    const float* inj_reset_u;
    float* fin_ret;        // [n] return of the last finished episode
    int* fin_len;          // [n]
    double* agg;           // [4]: episodes, sum return, sum length, (spare)
};

}  // namespace cf
