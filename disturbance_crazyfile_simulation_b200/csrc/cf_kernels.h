// cf_kernels.h - launchers implemented in cf_kernels.cu (all asynchronous on the given stream).
#pragma once
#include "cf_types.h"

namespace cf {

cudaError_t upload_constants(const double grid[6][CF_HJ_PTS], const double dx[6], const double cdf[21]);
void launch_init(float* S, int n_pad, const DevCfg& cfg, cudaStream_t st);
void launch_reset(float* S, int n, const DevCfg& cfg, const uint8_t* mask, float* obs, const float* inj_z,
                  const float* inj_u, cudaStream_t st);
cudaError_t launch_step(float* S, int n, int k, const DevCfg& cfg, const DevTables& tab, const StepIO& io, cudaStream_t st);
size_t step_smem_bytes(int obs_dim);
void launch_hj_signs(const float* V, uint8_t* P, cudaStream_t st);
void launch_hj_lookup(const double* states, int n, const float* V, const uint8_t* P, int* idx, float* sign,
                      cudaStream_t st);
void launch_hj_lookup_f32(const float* states, int n, const DevCfg& cfg, const float* V, const uint8_t* P, const float* thr,
                          int* idx, float* sign, cudaStream_t st);
void launch_unpack_state(const float* S, int n, float* rec, cudaStream_t st);
void launch_pack_state(float* S, int n, const float* rec, cudaStream_t st);
void launch_levels(const float* S, int n, float* out, cudaStream_t st);
void launch_gather_poses(const float* S, int n, const int* idx, int count, float* poses, cudaStream_t st);
void launch_gae(const float* rew, const float* val, const uint8_t* done, const float* end_val, const float* last_val,
                float gamma, float lam, float rew_scale, int T, int N, float* adv, float* tgt, float* disc_ret,
                cudaStream_t st);
void launch_debug_draws(const DevCfg& cfg, int first_env, int count, uint32_t ctr, uint32_t tag, int nzb, int nub, uint32_t* words,
                        float* draws, cudaStream_t st);
void launch_debug_euler(const float* quat, long long n, float* rpy_bullet, float* rpy_hj, cudaStream_t st);
void launch_debug_box_muller(const uint32_t* words, long long n, float* z, float* u, cudaStream_t st);
// fused rollout networks (cf_policy.cu): device pointers, torch Linear layout [out][in]
struct PolicyHostArgs {
    const float* obs; const float* obs_mean; const float* obs_std;
    const float* pi_w[3]; const float* pi_b[3]; const float* v_w[3]; const float* v_b[3];
    const float* log_std; const float* S;
    float* act; float* val; float* logp;
    int n, obs_dim, pi_h, v_h, deterministic;
    float eps;
};
cudaError_t launch_policy(const PolicyHostArgs& h, const DevCfg& cfg, cudaStream_t st);
// fused PPO policy update step (cf_ppo.cu): forward + clipped-surrogate loss + backward for pi.net; device pointers
struct PpoHostArgs {
    const float *obs, *act, *adv, *logp_old;      // [B][obs_dim] standardised, [B][4], [B], [B]   (policy net)
    const float* tgt; const long long* idx;       // value net: targets [all], mini-batch sample indices [B] (or null)
    int net;                                      // 0 policy (pi.net), 1 value (v.net)
    const float *w[3], *b[3], *log_std;           // net.{0,2,4} in torch layout, pi.log_std
    float *gw[3], *gb[3];                         // gradient tensors, same layouts
    float* loss_ratio;                            // [2]: mean loss, mean ratio (policy)
    float* partial;                               // workspace [n_cta][ppo_partial_floats()]
    int B, obs_dim, h, n_cta;
    float clip;
};
cudaError_t launch_ppo_grad(const PpoHostArgs& a, cudaStream_t st);
int ppo_partial_floats();
cudaError_t launch_ppo_pi_grad_static(const PpoHostArgs& a, cudaStream_t st);    // scratch in a file-scope __device__ array of cf_ppo.cu
int ppo_static_max_ctas();
long long ppo_static_max_samples();
void set_error(const char* where, const char* what);     // cf_abi.cu: text behind cf_last_error()
struct PdGains { float hover, kz, kvz, katt, krate, kyaw; };
void launch_pd_actions(const float* obs, int obs_dim, int off, int n, float* act, const PdGains& g, cudaStream_t st);

}  // namespace cf
