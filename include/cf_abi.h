/* cf_abi.h - C ABI of the B200 batched CrazyFlie hover environment and fused GAE.
 *
 * The reference (Hu-Hanyang/disturbance-CrazyFile-simulation) is pure Python and has
 * no FFI; the interface this library sits behind is the gym `Env` made by
 * `gym.make(id)` and its `reset()` / `step(action)` (envs/base.py:420-507 and the
 * step overrides envs/hover_free.py:391-444, 608-665, 778-833, envs/hover.py:649-705,
 * 866-926, 1036-1094, 1388-1444), plus `Buffer.finish_path` for GAE
 * (algs/core.py:498-535).  Each entry point below cites what it replaces.  The Python
 * binding a maintainer would add (ctypes) is shown in INTEGRATION.md.
 *
 * Conventions: every call returns 0 on success or a negative cf_status; no C++
 * exception crosses the boundary; `cf_last_error()` gives the message of the last
 * failure on the calling thread.  All env-major buffers (`act`, `obs`, `rew`, ...)
 * are caller-owned and contiguous.  The `_dev` entry points take DEVICE pointers and
 * are asynchronous on `stream` (a cudaStream_t passed as void*; NULL = legacy default
 * stream).  The `_host` entry points take HOST pointers, do the host<->device copies
 * themselves and return when the results are in the host buffers.
 * One handle <-> one device <-> one stream at a time; distinct handles are independent.
 */
#ifndef CF_ABI_H
#define CF_ABI_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CF_ABI_VERSION 1

typedef enum cf_status {
    CF_OK = 0,
    CF_ERR_INVALID = -1,   /* bad argument / unsupported configuration */
    CF_ERR_CUDA = -2,      /* CUDA runtime error (message in cf_last_error) */
    CF_ERR_NO_TABLE = -3,  /* an HJ disturbance level has no value table loaded */
    CF_ERR_ALLOC = -4
} cf_status;

/* disturbance source of the env variant (the classes differ only in this) */
enum {
    CF_DSTB_NONE = 0,          /* *WithoutAdversary: dstb = (0,0,0)   hover_free.py:814 */
    CF_DSTB_HJ_FIXED = 1,      /* *WithAdversary[Initial]: fixed level  hover_free.py:319,425 */
    CF_DSTB_HJ_BOLTZMANN = 2,  /* *WithRandomHJAdversary: Boltzmann() at the end of reset, hover_free.py:536 */
    CF_DSTB_RANDOM_BOX = 3     /* *WithRandomAdversary: dstb_space.sample(), hover_free.py:878,986 */
};

enum { CF_REWARD_FREE = 0,    /* envs/hover_free.py:206-235 */
       CF_REWARD_TARGET = 1   /* envs/hover.py:184-202 (-|xyz - target|) */ };

/* noise source */
enum { CF_RNG_PHILOX = 0,     /* counter-based Philox4x32-10 keyed by (seed, global env id) */
       CF_RNG_INJECTED = 1    /* parity mode: standard draws supplied by cf_set_noise_override */ };

enum { CF_HJ_VALUE_TABLE = 0, /* gather 5 fp32 values of V per env-step (distur_gener.py:40-99) */
       CF_HJ_SIGN_TABLE = 1   /* 2-bit/cell sign table precomputed on load with the same fp32 rule */ };

#define CF_ACT_DIM 4
#define CF_OBS13 13
#define CF_HJ_PTS 15
#define CF_HJ_CELLS 11390625      /* 15^6 */
#define CF_MAX_LEVELS 31          /* level index = round(level*10), levels <= 3.0 (distur_gener.py:154) */

/* live standard draws consumed per env-step / per reset */
#define CF_STEP_NORMALS 44
#define CF_STEP_UNIFORMS 8        /* 6 sensor-noise + 2 random-box disturbance */
#define CF_RESET_NORMALS 48
#define CF_RESET_UNIFORMS 42

/* Constructor kwargs of the reference env classes, flattened.  Defaults are filled by
 * cf_config_default(); cites in cf_abi.cu. */
typedef struct cf_config {
    int32_t abi_version;             /* CF_ABI_VERSION */
    int32_t reward_kind;             /* CF_REWARD_* */
    int32_t disturbance_kind;        /* CF_DSTB_* */
    int32_t observation_noise;       /* >0: SensorNoise pipeline, obs 34-d; <=0: exact state, obs 42-d */
    int32_t enable_reset_distribution;
    int32_t aggregate_phy_steps;     /* physics substeps per control step (2) */
    int32_t max_episode_steps;       /* gym TimeLimit (500); <=0 disables */
    int32_t gyroscopic;              /* Bullet model switch B2 */
    int32_t use_ground_effect;       /* physics.py:18,25 (reference never enables it) */
    int32_t hj_mode;                 /* CF_HJ_* */
    int32_t rng_mode;                /* CF_RNG_* */
    int32_t auto_reset;              /* 1: done envs are reset inside cf_step */
    double domain_randomization;     /* <=0 off; 0.1 = +-10 % */
    double disturbance_level;        /* CF_DSTB_HJ_FIXED */
    double penalty_action, penalty_angle, penalty_spin, penalty_terminal, penalty_velocity, penalty_z;
    double done_roll_pitch_rad;      /* deg2rad(75) (60 for the base classes) */
    double done_rate_deg;            /* 1000 (300) deg/s, compared as rad2deg(|w|) > limit */
    double init_pos_lim, init_angle, init_yaw, init_vel_lim, init_rate_lim, init_yaw_rate_lim;
    double cost_rate_limit;          /* rpy_dot_limit, deg2rad(200) (Initial variant: deg2rad(300)) */
    double sim_time_step;            /* 1/200 */
    double motor_time_constant;      /* 0.080 */
    double motor_thrust_noise;       /* 0.05 -> OU sigma 0.2*0.05 */
    double latency;                  /* 0.015 -> ring of int(latency // time_step) = 2 */
    double linear_damping, angular_damping;   /* Bullet model switch B3 (0.04) */
} cf_config;

typedef struct cf_env cf_env;

/* Host-visible state record: CF_STATE_WORDS 32-bit words per env, env-major
 * ([n_envs][CF_STATE_WORDS]); words are float unless marked (i) = int32/uint32 bits. */
enum {
    CF_S_POS = 0,        /* 3 */
    CF_S_EP_RET = 3,     /* 1  running episode return */
    CF_S_QUAT = 4,       /* 4  [x,y,z,w] */
    CF_S_VEL = 8,        /* 3  world */
    CF_S_OMEGA_W = 11,   /* 3  world-frame angular velocity (as Bullet stores it) */
    CF_S_LPF = 14,       /* 3  gyro low-pass state */
    CF_S_GYRO_BIAS = 17, /* 3 */
    CF_S_MOTOR_X = 20,   /* 4  PT1 motor state */
    CF_S_OU = 24,        /* 4  thrust-noise OU state */
    CF_S_RING = 28,      /* 8  action latency ring [2][4] */
    CF_S_ACT_PREV1 = 36, /* 4  u(k-1) */
    CF_S_ACT_PREV2 = 40, /* 4  u(k-2) */
    CF_S_OBS_PREV = 44,  /* 13 last emitted obs13 */
    CF_S_EP_LEN = 57,    /* (i) steps since reset */
    CF_S_RNG_CTR = 58,   /* (i) per-env RNG event counter */
    CF_S_MISC = 59,      /* (i) level_idx | ring_idx<<8 | iteration<<12 | fresh<<31 */
    CF_S_DT = 60,        /* domain-randomised parameters: 14 words */
    CF_S_MASS = 61,
    CF_S_INERTIA = 62,   /* 3 */
    CF_S_FTF1 = 65,
    CF_S_MOTOR_B = 66,   /* 4  T_s/T  (A = 1 - B) */
    CF_S_MOTOR_K = 70,   /* 4 */
    CF_STATE_WORDS = 76
};

/* ---- lifecycle -------------------------------------------------------------------- */
/* Fill `cfg` with the constructor defaults of a registered env id, e.g.
 * "DroneHoverBulletFreeEnvWithRandomHJAdversary-v0" (phoenix_drone_simulation/__init__.py:8-109). */
int cf_config_default(const char* env_id, cf_config* cfg);

/* gym.make(id, **kwargs) for n_envs envs whose global ids are env_id_base..env_id_base+n_envs-1
 * (RNG is keyed by the global id, so results do not depend on how envs are sharded over GPUs). */
int cf_create(const cf_config* cfg, int32_t n_envs, int32_t device, uint64_t seed,
              int64_t env_id_base, cf_env** out);
int cf_destroy(cf_env* env);
int cf_obs_dim(const cf_env* env);       /* 34 (noise on) or 42 (noise off), envs/base.py:139-141 */
int cf_num_envs(const cf_env* env);

/* np.load(fastrack_{level}_15x15.npy) of distur_gener.py:155, hoisted: upload one float32
 * (15,)*6 C-order value table (HOST pointer) for level index round(level*10). */
int cf_load_hj_table(cf_env* env, int32_t level_idx, const float* V_host);
/* Let `env` use the device tables already uploaded to `owner` (same device, same hj_mode) instead of
 * its own copy: one table set per GPU shared by all env shards.  `owner` must outlive `env`. */
int cf_share_hj_tables(cf_env* env, const cf_env* owner);
/* read back the packed 2-bit sign table of a level (HOST pointer, CF_HJ_CELLS/4+1 bytes) */
int cf_get_hj_signs(cf_env* env, int32_t level_idx, uint8_t* out_host);

/* ---- device-pointer API (asynchronous on `stream`) ---------------------------------- */
/* env.reset() (envs/base.py:420-464; hover_free.py:492-538) for envs with mask[i]!=0 (mask NULL = all).
 * obs: [n_envs][obs_dim] float32 (rows of unmasked envs untouched). */
int cf_reset_dev(cf_env* env, const uint8_t* mask, float* obs, void* stream);

/* env.step(action) under TimeLimit for all envs.  act [n][4] (16-byte aligned); obs [n][obs_dim]
 * (8-byte aligned; 16-byte aligned rows leave the SM as bulk copies); rew [n];
 * done [n] (1 = terminal or time-limit); cost [n] (info['cost'], may be NULL);
 * term_obs [n][obs_dim] (observation before the auto-reset, written for done envs only; may be NULL);
 * truncated [n] (info['TimeLimit.truncated'], may be NULL). */
int cf_step_dev(cf_env* env, const float* act, float* obs, float* rew, uint8_t* done,
                float* cost, float* term_obs, uint8_t* truncated, void* stream);

/* k env.step calls in one launch, the tile's state kept in shared memory: act [k][n][4], obs [k][n][obs_dim],
 * rew/done/cost [k][n] (cost may be NULL).  Open-loop / throughput entry: the k action batches must exist up front, and it
 * has NO truncated / terminal-observation outputs - with auto_reset on, a caller cannot tell time-limit cuts from true
 * terminals in its [k][n] done flags, so PPO-style bootstrapping from V(s_T) needs cf_step_dev (one step per launch,
 * `truncated` and `term_obs` outputs).  The last finished episode's return / length per env stay available through
 * cf_last_episodes. */
int cf_step_k_dev(cf_env* env, int32_t k, const float* act, float* obs, float* rew,
                  uint8_t* done, float* cost, void* stream);

/* Shard pipelining hint for callers that step several handles of one GPU in turn (their combined state does not
 * fit the L2): while `env` is stepped, its step kernel asks the L2 to prefetch the state of `next` (and the
 * `next_act` rows, [n_next][4], nullable), so that the next handle's step finds its inputs in L2 instead of
 * waiting for HBM at its start.  A hint only: results never depend on it.  next = NULL switches it off.
 * `next` must be on the same device and outlive the setting. */
int cf_set_prefetch_next(cf_env* env, const cf_env* next, const float* next_act);

/* ---- host-pointer API (what a gym-style caller with NumPy buffers uses) -------------- */
int cf_reset_host(cf_env* env, const uint8_t* mask, float* obs);
int cf_step_host(cf_env* env, const float* act, float* obs, float* rew, uint8_t* done,
                 float* cost, float* term_obs, uint8_t* truncated);

/* The same step, not waited for: the copies and the kernel are queued on the handle's private stream and the call
 * returns; the host buffers (pinned: cf_host_alloc) must stay untouched until cf_host_sync(env) returns.  Lets a caller
 * with several handles overlap one handle's download with the next one's upload and kernel.
 * ONE DOWNLOAD PER STEP (both calls): when num_envs is a multiple of 4 and the output arrays are ADJACENT in host memory in the
 * order obs [n][obs_dim] | rew [n] | cost [n] | done [n] - i.e. rew == obs + n * obs_dim, cost == rew + n, done == (uint8_t*)(cost + n),
 * one block of n * (4 * obs_dim + 9) bytes - they arrive as a single copy instead of four (each separate copy costs ~4 us of fixed
 * time on the copy engine that paces this path: 3.4e8 -> 3.6e8 env-steps/s on one B200).  Any other arrangement works as before. */
int cf_step_host_async(cf_env* env, const float* act, float* obs, float* rew, uint8_t* done,
                       float* cost, float* term_obs, uint8_t* truncated);
int cf_host_sync(cf_env* env);

/* ---- state / parity hooks ----------------------------------------------------------- */
int cf_get_state(cf_env* env, float* state_host /* [n][CF_STATE_WORDS] */);
int cf_set_state(cf_env* env, const float* state_host);
/* disturbance_level of each env (info['disturbance_level']) */
int cf_get_levels(cf_env* env, float* levels_host);
/* parity mode (rng_mode = CF_RNG_INJECTED): standard normal / U[0,1) draws for the NEXT call.
 * step:  normals [CF_STEP_NORMALS][n],  uniforms [CF_STEP_UNIFORMS][n]
 * reset: normals [CF_RESET_NORMALS][n], uniforms [CF_RESET_UNIFORMS][n]  (slot-major, HOST pointers;
 * either may be NULL to leave that set unchanged).  Layout of the slots: DESIGN.md "noise slots". */
int cf_set_noise_override(cf_env* env, const float* step_normals, const float* step_uniforms,
                          const float* reset_normals, const float* reset_uniforms);
/* grid indices and disturbance of the HJ lookup for given (roll,pitch,yaw,p,q,r) states
 * (Grid.get_index + spa_deriv + opt_dstb_non_hcl, distur_gener.py:40-125); states [n][6] float64,
 * idx [n][6] int32, dstb [n][2] float32 sign codes (+1/-1); HOST pointers. */
int cf_hj_lookup_host(cf_env* env, int32_t level_idx, const double* states, int32_t n,
                      int32_t* idx, float* dstb_sign);
/* the same for float32 states [n][6] through the step kernel's own float32 rule (per-boundary float32
 * thresholds derived on the host from the float64 rule): must equal cf_hj_lookup_host on (double)state */
int cf_hj_lookup_f32_host(cf_env* env, int32_t level_idx, const float* states, int32_t n,
                          int32_t* idx, float* dstb_sign);

/* ---- PPO GAE (Buffer.finish_path, algs/core.py:498-535; discount_cumsum :105-119) ----- */
/* Time-major rollout [T][N].  A path ends after step t of env n when done[t][n] != 0; its
 * bootstrap value is end_val[t][n] (0 for a true terminal, V(s_T) for a time-limit cut,
 * iwpg.py:392-402); paths still open at t = T-1 bootstrap from last_val[n].
 * rew_scale: reward divisor 1/(ret_std+1e-5) applied with clip +-10 when > 0
 * (use_scaled_rewards, online_mean_std.py:32-48); <= 0 disables scaling.
 * Outputs adv, target_v, disc_ret [T][N].  DEVICE pointers, asynchronous on `stream`. */
int cf_gae_dev(const float* rew, const float* val, const uint8_t* done, const float* end_val,
               const float* last_val, float gamma, float lam, float rew_scale,
               int32_t T, int32_t N, float* adv, float* target_v, float* disc_ret, void* stream);
int cf_gae_host(const float* rew, const float* val, const uint8_t* done, const float* end_val,
                const float* last_val, float gamma, float lam, float rew_scale,
                int32_t T, int32_t N, float* adv, float* target_v, float* disc_ret);

/* ---- episode statistics (what the reference logs per finished episode, loggers.py:612-633) -- */
/* out[0..3] = episodes finished, sum of their returns, sum of their lengths, spare - accumulated
 * on the device since creation or the last call with reset != 0.  fin_ret_host / fin_len_host
 * (nullable, [n]) receive the return and length of each env's LAST finished episode. */
int cf_episode_stats(cf_env* env, int32_t reset, double out[4], float* fin_ret_host, int32_t* fin_len_host);
/* One row of the reference's episodic_data.csv per env (utils/loggers.py:612-618: episode, disturbance_level, steps, return):
 * return, length and disturbance level of each env's LAST finished episode (length 0 = none yet); host arrays [n].  The level is
 * the one the episode was flown with (hover_free.py:536 draws the next one at the end of reset). */
int cf_last_episodes(cf_env* env, float* ret_host, int32_t* len_host, float* level_host);

/* ---- the rollout's networks, fused -------------------------------------------------------
 * ActorCritic.step of the reference (algs/core.py:371-394) as IWPGAlgorithm.roll_out calls it once per control step
 * (algs/iwpg/iwpg.py:372-410), for all envs of a handle in ONE launch: obs_oms(obs) = (obs - mean) / (std + eps)
 * (utils/online_mean_std.py:32-48), value MLP obs_dim -> v_hidden -> v_hidden -> 1 with tanh (core.py:298-312), policy MLP
 * obs_dim -> pi_hidden -> pi_hidden -> 4 with relu (core.py:227-260), then a = mu + exp(log_std) * N(0,1) with
 * logp = sum_j Normal(mu_j, std_j).log_prob(a_j) (core.py:257-266), or a = mu, logp = 1 when `deterministic` (eval mode,
 * core.py:284-290).  float32 FFMA arithmetic.  All pointers are DEVICE pointers; weight matrices in torch's nn.Linear
 * layout [out][in] row-major (the tensors of the reference's state_dict: pi.net.{0,2,4}.{weight,bias}, v.net.{0,2,4}.*,
 * pi.log_std, obs_oms.{mean,std}); obs_mean / obs_std may be NULL (no standardisation).  Hidden widths <= 64.
 * The exploration noise of env i is Philox4x32-10 keyed by the handle's seed with counter (global env id, the env's event
 * counter CF_S_RNG_CTR, block 0, tag 4): call it BEFORE the cf_step_* that consumes the actions; results do not depend on
 * how envs are sharded.  obs [n][obs_dim] (8-byte aligned), act [n][4] (16-byte aligned), val [n], logp [n]; asynchronous
 * on `stream`. */
typedef struct cf_policy_weights {
    const float* obs_mean; const float* obs_std;
    const float* pi_w[3]; const float* pi_b[3];
    const float* v_w[3]; const float* v_b[3];
    const float* log_std;
    int32_t pi_hidden, v_hidden;
    float obs_eps;              /* 1e-5, online_mean_std.py:12 */
    int32_t reserved;
} cf_policy_weights;
int cf_policy_step_dev(cf_env* env, const cf_policy_weights* w, const float* obs, int32_t deterministic, float* act, float* val,
                       float* logp, void* stream);

/* One policy-update iteration of PPO up to the optimiser step, fused (IWPGAlgorithm.update_policy_net, algs/iwpg/iwpg.py:448-459:
 * compute_loss_pi of algs/ppo/ppo.py:22-40 + loss.backward()): forward of the policy MLP on the standardised observations,
 * Gaussian log-prob of the stored actions, ratio, clipped surrogate -mean(min(ratio adv, clip(ratio, 1 - c, 1 + c) adv)) and the
 * gradients of the six pi.net tensors, in ONE kernel over tiles of 128 samples + a deterministic reduction.  All DEVICE pointers;
 * weights and gradients in torch's nn.Linear layout [out][in] (w->pi_w / pi_b / log_std / pi_hidden are read, the rest of the
 * struct is ignored); obs [B][obs_dim] ALREADY standardised (pre_process_data, iwpg.py:350-370), act [B][4] 16-byte aligned,
 * adv [B], logp_old [B]; grad_w[l] / grad_b[l] receive d loss / d pi.net.{0,2,4}.{weight,bias}; loss_ratio [2] = mean loss (without
 * the entropy bonus, which has no gradient: log_std is not trained), mean ratio; workspace: cf_ppo_workspace_floats() floats. */
int cf_ppo_workspace_floats(void);
int cf_ppo_pi_grad_dev(const cf_policy_weights* w, int32_t obs_dim, const float* obs, const float* act, const float* adv,
                       const float* logp_old, int64_t B, float clip_ratio, float* const grad_w[3], float* const grad_b[3],
                       float* loss_ratio, float* workspace, void* stream);
/* cf_ppo_pi_grad_dev without a workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and arithmetic as cf_ppo_pi_grad_dev; the per-CTA partial sums (the kernel's scratch)
 * live in a file-scope `static __device__ __align__(256)` array of 148 x 7560 floats = 4.27 MiB, which every GPU that loads the
 * library holds until the process exits.  LIMIT: B <= cf_ppo_static_max_samples() = 4 194 304 samples; more is rejected with
 * CF_ERR_INVALID and a message naming the limit.  The call performs no cudaMalloc / cudaFree and calls no allocating library
 * routine; on the first call per device it resolves the array's address (cudaGetSymbolAddress).  It runs 148 persistent CTAs
 * where cf_ppo_pi_grad_dev runs 296, so the two sum in a different order: results agree to float32 rounding (2e-5 of each
 * tensor's largest entry in tests/test_zz_no_alloc.py), not bit for bit.  ALL CALLS SHARE THE ONE ARRAY: it keeps the last
 * call's contents (every launch overwrites all it reads, nothing needs clearing), and calls that overlap on the GPU - other
 * streams, other host threads - race on it; order them on one stream or with events. */
int64_t cf_ppo_static_max_samples(void);
int cf_ppo_pi_grad_static_dev(const cf_policy_weights* w, int32_t obs_dim, const float* obs, const float* act, const float* adv,
                              const float* logp_old, int64_t B, float clip_ratio, float* const grad_w[3], float* const grad_b[3],
                              float* loss_ratio, void* stream);

/* One mini-batch step of the value update up to the optimiser step, the same way (IWPGAlgorithm.update_value_net,
 * algs/iwpg/iwpg.py:478-510; loss = mean((v(obs[idx]) - target_v[idx])^2), iwpg.py:250-255): v.net forward (tanh), loss and the six
 * gradients.  obs [all][obs_dim] standardised, target_v [all]; idx [B] int64 rows of this mini-batch (torch.randperm slice), or
 * NULL for rows 0..B-1; w->v_w / v_b / v_hidden are read; loss [2]: loss[0] = mean squared error (loss[1] unused). */
int cf_ppo_v_grad_dev(const cf_policy_weights* w, int32_t obs_dim, const float* obs, const float* target_v, const int64_t* idx,
                      int64_t B, float* const grad_w[3], float* const grad_b[3], float* loss, float* workspace, void* stream);

/* ---- synthetic "PID actions" source of the throughput runs ------------------------------------
 * Attitude -> rate -> motor-mix PD law on the NOISY observation rows (device pointers, asynchronous): structure of
 * the firmware cascade the reference cites (envs/control.py:13-26 gains, :33-50 rpy_control_factors_to_PWM,
 * :94-100 PWM.act), in action units.  obs [n][obs_dim] (34 or 42), act [n][4] 16-byte aligned,
 * gains = {hover action, kz, kvz, k_attitude [1/s], k_rate [action per rad/s], k_yaw_rate}. */
int cf_pd_actions_dev(const float* obs, int32_t obs_dim, int32_t n, float* act, const float gains[6], void* stream);

/* ---- HJ value-table generator ---------------------------------------------------------
 * The FasTrack solve that produces fastrack_{level}_15x15.npy in the reference
 * (adversarial_generation/FasTrack_data/Fastrack.py:281-320 -> odp/solver.py:83-233 -> odp/computeGraphs/graph_6D.py
 * with odp/spatialDerivatives/first_orderENO6D.py, accuracy "low", TargetSetMode maxVWithV0), float32 like the
 * reference's HeteroCL build.  All scalars arrive rounded to float32 by the caller; sin / cos of the grid values of
 * dims 0, 1 are tabulated by the caller.  Value arrays are C-order [pts]^6 floats. */
#define CF_HJ_MAX_PTS 16
typedef struct cf_hj_problem {
    int32_t pts;                        /* grid points per dimension, 3..CF_HJ_MAX_PTS */
    int32_t reserved;
    float x[6][CF_HJ_MAX_PTS];          /* grid values (odp Grid.vs, GridProcessing.py:6-51) */
    float sin_phi[CF_HJ_MAX_PTS], cos_phi[CF_HJ_MAX_PTS], tan_theta[CF_HJ_MAX_PTS], cos_theta[CF_HJ_MAX_PTS];
    double target_sq[6][CF_HJ_MAX_PTS]; /* float64 (x_d - center_d)^2 per dim, 0 for ignored dims: l0 = (float)sum (Fastrack.py:296-306) */
    float dx[6];
    float u_min[3], u_max[3], d_min[3], d_max[3];   /* Fastrack.py:22-23, 42-44 */
    float c[3], inertia[3];                         /* (Iyy-Izz)/Ixx, (Izz-Ixx)/Iyy, (Ixx-Iyy)/Izz; Ixx, Iyy, Izz */
} cf_hj_problem;
typedef struct cf_hj_solver cf_hj_solver;
/* V starts at l0 (solver.py:108) */
int cf_hj_solver_create(const cf_hj_problem* prob, int32_t device, cf_hj_solver** out);
int cf_hj_solver_destroy(cf_hj_solver* s);
int cf_hj_solver_set_value(cf_hj_solver* s, const float* V_host);
int cf_hj_solver_get_value(cf_hj_solver* s, float* V_host);
/* one call of the compiled graph_6D kernel with t = (t0, t1): V <- max(V + rate * dt, l0); returns the advanced
 * t0 and, when info != NULL, {min_deriv[6], max_deriv[6], max_alpha[6], dt} (graph_6D.py:64-330) */
int cf_hj_solver_update(cf_hj_solver* s, float t0, float t1, float* t0_new, float info[19]);
/* the time loop of HJSolver (solver.py:152-206) over tau = arange(0, lookback + 1e-5, t_step); max_updates <= 0: no cap */
int cf_hj_solver_run(cf_hj_solver* s, double lookback, double t_step, int64_t max_updates, int64_t* updates_done,
                     double* t_final);
/* average device time of one update (all its kernels) over the updates run so far, in ms */
int cf_hj_solver_timing(cf_hj_solver* s, double* avg_ms_per_update, int64_t* updates);

/* ---- f4: headless rasteriser ---------------------------------------------------------------------
 * Replaces BulletClient.getCameraImage(..., renderer=ER_BULLET_HARDWARE_OPENGL) behind the reference's
 * `capture_image()` (envs/base.py:520-544) and `render('rgb_array')` (envs/base.py:379-416): 1920x1080x3 uint8
 * frames of the drone over the ground plane, written into a device buffer - no graphics API involved (graphics
 * is unavailable on this box).  The camera is the one those calls build: computeViewMatrixFromYawPitchRoll
 * (target = drone position, distance 1.8, yaw 10, pitch -20, roll 0, up axis z) and computeProjectionMatrixFOV
 * (fov 60, aspect w/h, near 0.1); the host derives the basis below from those numbers (raster.py).
 * The drone is a triangle soup in the body frame (the URDF's visual meshes loaded by the host, or the built-in
 * proxy made of the URDF's collision shapes); triangles are flat-shaded, two-sided, z-tested per pixel; a triangle
 * with a vertex in front of the near plane is dropped, not clipped.  The ground is the z = 0 plane out to
 * +-ground_half metres with a procedural checker.  All arithmetic is float32 without fused multiply-adds so that
 * oracle/raster_oracle.py reproduces every byte. */
typedef struct cf_camera {
    int32_t width, height;       /* width % 4 == 0 */
    int32_t follow;              /* 1: target = the frame's drone position (capture_image); 0: target[] (render) */
    float target[3];
    float eye_offset[3];         /* eye = target + eye_offset */
    float right[3], up[3], fwd[3];   /* orthonormal camera basis in the world frame */
    float tan_x, tan_y;          /* tan(fov/2) * aspect, tan(fov/2) */
    float near_plane;
    float light[3];              /* unit vector towards the light */
    float ambient, diffuse;
    float background[3], ground_a[3], ground_b[3];   /* colours in 0..1 */
    float inv_checker;           /* 1 / edge of a ground square, 1/m */
    float ground_half;           /* half extent of the ground plane, m */
} cf_camera;

typedef struct cf_raster cf_raster;
/* tris_host [n_tris][3][3] body-frame vertices, rgb_host [n_tris][3] colours in 0..1 (both copied);
 * bound_radius >= the largest vertex norm (used only to skip pixel blocks the drone cannot touch) */
int cf_raster_create(cf_raster** out, int32_t device, const float* tris_host, const float* rgb_host, int32_t n_tris,
                     float bound_radius);
void cf_raster_destroy(cf_raster* r);
/* poses_dev [n_frames][7] = position xyz, quaternion xyzw (DEVICE); rgb_dev [n_frames][height][width][3] uint8 (DEVICE);
 * asynchronous on `stream` (a cudaStream_t, NULL = default stream) */
int cf_raster_render_dev(cf_raster* r, const cf_camera* cam, const float* poses_dev, int32_t n_frames, uint8_t* rgb_dev,
                         void* stream);
/* the same with HOST buffers; synchronous */
int cf_raster_render_host(cf_raster* r, const cf_camera* cam, const float* poses_host, int32_t n_frames, uint8_t* rgb_host);
/* pose rows of envs env_idx_dev[0..count) (DEVICE int32; NULL = envs 0..count-1) into poses_dev [count][7] (DEVICE) */
int cf_get_poses_dev(cf_env* env, const int32_t* env_idx_dev, int32_t count, float* poses_dev, void* stream);

/* ---- pinned host memory for the host-pointer API (so its copies are asynchronous DMA) ------- */
void* cf_host_alloc(uint64_t bytes);     /* NULL on failure (cf_last_error) */
void cf_host_free(void* p);

/* number of kernels this library has launched on this handle since creation */
int64_t cf_launch_count(const cf_env* env);
/* average device time (ms) of the step kernel launches since the last call with reset!=0,
 * measured with CUDA events on the launching stream when profiling was enabled */
int cf_kernel_timing(cf_env* env, int32_t enable, int32_t reset, double* avg_ms, int64_t* launches);

/* Parity hooks for the production RNG path (rng_mode = CF_RNG_PHILOX), host pointers, synchronous.
 * cf_debug_draws: the raw Philox4x32-10 words and the standard draws of ONE event of envs first_env .. first_env+count-1,
 * produced by the device functions the step / reset kernels call.  An event is (global env id, event counter, tag):
 * tag 0 = step, 1 = in-place reset inside a step launch, 2 = cf_reset_*, 3 = constructor-time gyro-bias draw; the
 * counter is the env's CF_S_RNG_CTR word when the event runs.  Blocks 0..n_norm_blocks-1 are normal blocks (Philox
 * counter word 2 = b; four normals each by Box-Muller), the following n_unif_blocks are uniform blocks (counter word 2 =
 * 16 + b).  words_host [count][n_norm_blocks+n_unif_blocks][4] uint32, draws_host the same shape in float32.
 * cf_debug_box_muller: the normals [n][4] and the u01 values [n][4] the kernels derive from given raw words [n][4]
 * (no handle needed): lets a test sweep all 2^23 mantissa patterns. */
int cf_debug_draws(cf_env* env, int32_t first_env, int32_t count, uint32_t event_ctr, uint32_t tag, int32_t n_norm_blocks,
                   int32_t n_unif_blocks, uint32_t* words_host, float* draws_host);
int cf_debug_box_muller(int32_t device, const uint32_t* words_host, int64_t n, float* normals_host, float* u01_host);
/* The two quaternion -> Euler conversions of a step exactly as the kernels compute them (pybullet.getEulerFromQuaternion
 * of agents.py:446 and quat2euler of distur_gener.py:186-207, both on a polynomial atan2): quat_host [n][4] (x, y, z, w),
 * outputs [n][3] (roll, pitch, yaw). */
int cf_debug_euler(int32_t device, const float* quat_host, int64_t n, float* rpy_bullet_host, float* rpy_hj_host);

/* Tracing hook: when `trace_dev` (DEVICE pointer, [cf_num_tiles][16] uint64, caller-owned, zeroed by the caller)
 * is set, every step launch records per tile (times in %globaltimer ns): [0] SM id, [1] CTA start, [2] state tile
 * landed in shared memory, [3..6] warp 0..3 finished the step phase, [7] tile barrier passed, [8] in-place resets
 * done, [9] CTA end, [10] envs of the tile that finished an episode, [11] first half of the step's draws generated,
 * [12] event counter arrived, [13] table gather issued + second half of the draws generated, [14] / [15] reset
 * draws ready / reset tasks done (first group of finished envs).  NULL switches tracing off (the default; no cost
 * beyond a predicate). */
int cf_set_trace_buffer(cf_env* env, uint64_t* trace_dev);
int cf_num_tiles(const cf_env* env);     /* tiles of envs the step kernel launches (one CTA each) */

const char* cf_last_error(void);
int cf_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* CF_ABI_H */
