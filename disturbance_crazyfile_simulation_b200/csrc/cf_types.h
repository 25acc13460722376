// cf_types.h - plain structs shared by the host side (cf_abi.cu) and the kernels (cf_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/cf_abi.h"

// envs per tile = threads per CTA of the step kernel.  State in HBM: float S[n_tiles][CF_STATE_WORDS][CF_TILE]
#ifndef CF_TILE
#define CF_TILE 128
#endif
// words 0..CF_DYN_WORDS-1 change every step (staged through shared memory by the step kernel);
// the rest are the domain-randomised parameters, rewritten by reset only
#define CF_DYN_WORDS CF_S_DT
// Device word order inside a tile (the host-visible record keeps the CF_S_* order of cf_abi.h): the three integer words
// move in front of the 21 history words (u(k-1), u(k-2), last emitted obs13), so that words 0..CF_BULK_WORDS-1 are what
// the step kernel stages through shared memory with one bulk copy, and the history words - which a step only copies
// through into the observation row and overwrites at its very end - stay in global memory (coalesced LDG / STG).
#define CF_BULK_WORDS 39
#define CF_HIST_WORDS 21
#if defined(__CUDACC__)
#define CF_HD __host__ __device__
#else
#define CF_HD
#endif
CF_HD constexpr int cf_dev_word(int w) {
    return w < CF_S_ACT_PREV1 ? w : (w < CF_S_EP_LEN ? w + 3 : (w < CF_S_DT ? w - CF_HIST_WORDS : w));
}
// bytes between the packed 2-bit sign tables of consecutive levels (>= ceil(15^6 / 4), 128-B multiple)
#define CF_SIGN_STRIDE 2847744
// episode-statistics accumulators are spread over slots (by tile) to avoid same-address atomics
#define CF_AGG_SLOTS 64
// uint64 words per tile of the debug timeline (cf_set_trace_buffer)
#define CF_TRACE_WORDS 16

namespace cf {

static_assert(CF_DYN_WORDS == 60 && CF_STATE_WORDS == 76, "state record layout");
static_assert((CF_BULK_WORDS * CF_TILE * 4) % 16 == 0 && (CF_HIST_WORDS * CF_TILE * 4) % 16 == 0, "bulk copies need 16-byte multiples");
static_assert(cf_dev_word(CF_S_ACT_PREV1) == CF_BULK_WORDS && cf_dev_word(CF_S_OBS_PREV + 12) == CF_DYN_WORDS - 1 &&
              cf_dev_word(CF_S_EP_LEN) == 36 && cf_dev_word(CF_S_MISC) == 38 && CF_BULK_WORDS + CF_HIST_WORDS == CF_DYN_WORDS,
              "device word order");

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
    float hj_lo[6], hj_inv_dx[6];           // grid origin and 1/dx per dim (index guess, float32)
    float hj_thr[96];                       // [6][16] float32 index thresholds (cf_device.cuh hj_index_f32), same as DevTables::thr
    float boltz_thr[20];                    // Boltzmann level draw: smallest float32 >= cdf[k] (envs/utils.py:27-39)
    uint32_t pk[10][2];                     // Philox4x32-10 round keys of the seed
    long long gid_base;
};

// Device-side tables owned by the handle (or borrowed from another handle on the same GPU).
struct DevTables {
    const float* const* value;              // device array [CF_MAX_LEVELS] of fp32 (15,)*6 value tables (or null entries)
    const uint8_t* signs;                   // slab [CF_MAX_LEVELS][CF_SIGN_STRIDE]: packed 2-bit sign tables, 4 cells / byte
    const float* thr;                       // [6][16] float32 index thresholds (cf_device.cuh hj_index_f32)
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
    double* agg;           // [CF_AGG_SLOTS][4]: episodes, sum return, sum length, (spare)
    int bulk_obs;          // obs rows of a full tile may leave with one bulk copy (16-byte aligned)
    const float* pf_S;     // L2 prefetch hint (cf_set_prefetch_next): tiled state of the handle stepped next, or null
    const float* pf_act;   // its action rows [pf_n][4] or null
    int pf_tiles, pf_n;
    unsigned long long* trace;   // debug: [n_tiles][CF_TRACE_WORDS] per-CTA timeline (cf_set_trace_buffer) or null
    int* reset_list;       // deferred resets (single-step launches): env indices that finished an episode in this launch ...
    int* reset_count;      // ... [0] their number, [1] ticket of the list kernel's CTAs; both zeroed again by k_reset_list
};

}  // namespace cf
