/* oracle/ppo_oracle.h — TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C (gcc, -ffp-contract=off) CPU restatement of the PPO-LibTorch hot path, one function
 * per SURVEY.md §8(a) row.  It is the checker the CUDA kernels are compared against on the
 * GPU box (where /root/reference does not exist).  It is pinned against the compiled
 * reference itself (oracle/_ref/libppo_ref.so, built from /root/reference by oracle/Makefile)
 * in tests/test_oracle_vs_ref.py and against the committed fixtures tests/golden/ *.npz
 * generated from that reference build (tests/golden/make_golden.py).
 *
 * Parity status: the reference ships NO tests or golden vectors of its own, so
 * the pins are outputs of the reference run in the build container -- "pinned by reference
 * execution", not by reference-authored known answers.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library.
 *
 * All citations are file:line under /root/reference.
 */
#ifndef PPO_ORACLE_H
#define PPO_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_ENV_CARTPOLE = 0, ORC_ENV_MOUNTAINCAR = 1, ORC_ENV_MULTIWALK = 2 };
enum { ORC_ENTROPY_REFERENCE = 0, ORC_ENTROPY_EXACT = 1 };

/* ---- a1..a4: single-env dynamics ------------------------------------------------ */
/* Environments/CartPole.cpp:47-94 */
void orc_cartpole_step(float* state4, int64_t action, float* reward, int* terminated);
/* Environments/CartPole.cpp:34-45,97-101 with injected canonical uniforms u in [0,1):
 * libstdc++ uniform_real_distribution<float>(a,b) returns u*(b-a)+a. */
void orc_cartpole_reset(float* state4, const float* u4);
/* libstdc++ mt19937 -> generate_canonical<float,24> stream, to reproduce CartPole(seed).reset(). */
void orc_mt19937_canonical_f32(uint32_t seed, int n, float* u_out);
/* Environments/MountainCar.cpp:29-57 */
void orc_mountaincar_step(float* state2, int64_t action, float* reward, int* terminated);
/* Environments/MountainCar.cpp:60-67,79-88 with injected uniform */
void orc_mountaincar_reset(float* state2, float u);

/* MultiWalk (synthetic MultiDiscrete env with a state-dependent action mask; csrc/env.cuh: MultiWalkEnv).  No reference
 * counterpart: parity for this env is against this restatement only ("parity unpinned"). */
void orc_multiwalk_reset(float* state4, const float* u4);
void orc_multiwalk_step(float* state4, const int32_t* acts8, float* reward, int* terminated);
void orc_multiwalk_mask(const float* state4, uint8_t* mask21);

/* ---- a5: vectorised step (PPO/PPO_Discrete.cpp:413-483) ------------------------------ */
/* state AoS [n,S] (S=4 cartpole, 2 mountaincar); reset_u [n,4] (row i used iff env i resets;
 * mountaincar uses column 0).  done_out is float 0/1 as stored into m_dones (:530,546).
 * fin_len/fin_ret: episode length/return of envs that finished this step, else 0. */
void orc_env_step(int kind, int n, float* state, int32_t* ep_len, float* ep_ret,
                  const int32_t* actions, int max_episode_steps, const float* reset_u,
                  float* obs_out, float* reward_out, float* done_out,
                  int32_t* fin_len, float* fin_ret);

/* ---- a6: actor-critic MLP (PPO/Agent.cpp:25-59,107-128) ------------------------------ */
/* params: flat, reference parameter order (critic W1,b1,W2,b2,W3,b3 then actor ...),
 * weights [out,in] row-major.  K = total logits. */
int64_t orc_num_params(int O, int H, int K);
void orc_mlp_forward(const float* params, int O, int H, int K, int n, const float* obs,
                     float* logits /*[n,K]*/, float* value /*[n]*/);

/* ---- a7/a8: Categorical / CategoricalMasked over `heads` heads --------------------------- */
/* Distributions/Categorical.cpp:28-39,73-79,92-101,112-119; CategoricalMasked.cpp:31-46,87-144;
 * multi-head sum PPO/Agent.cpp:137-170.
 * mask: u8 [n,sumK] or NULL (unmasked).  actions_in: i32 [n,heads] or NULL -> sample by
 * inverse CDF on uniforms [n,heads]: action = first j with u < cumsum(p)_j (fallback: last
 * j with p_j>0).  probs_out / normlogits_out optional [n,sumK]. */
void orc_categorical(int n, int heads, const int32_t* nvec, const float* logits,
                     const uint8_t* mask, const int32_t* actions_in, const float* uniforms,
                     int entropy_mode, int32_t* actions_out, float* logprob, float* entropy,
                     float* probs_out, float* normlogits_out);

/* ---- a9: GAE / returns (PPO/PPO_Discrete.cpp:274-331) ------------------------------- */
void orc_gae(int T, int N, const float* rewards, const float* values, const float* dones,
             const float* next_value, const float* next_done, float gamma, float lam,
             int use_gae, float* adv_out, float* ret_out);

/* ---- a10/a11: minibatch loss + closed-form backward (PPO/PPO_Discrete.cpp:579-638) ---- */
typedef struct {
    float clip_coef;
    int norm_adv;
    int clip_vloss;
    float ent_coef;
    float vf_coef;
    int entropy_mode;
} orc_hyper;
/* Inputs are the already-gathered minibatch rows.  grads: flat parameter order, PRE-clip.
 * diag[8] = {pg_loss, v_loss, entropy_mean, loss, approx_kl, clipfrac, adv_mean, adv_std}. */
void orc_minibatch_grads(const float* params, int O, int H, int heads, const int32_t* nvec,
                         int n, const float* obs, const uint8_t* mask, const int32_t* actions,
                         const float* old_logprob, const float* adv, const float* ret,
                         const float* old_val, const orc_hyper* hp, float* grads, float* diag);

/* ---- a12: clip_grad_norm_ + AdamW (PPO/PPO_Discrete.cpp:640-641; LibTorch adamw.cpp) ---- */
/* step = 1-based step count AFTER increment.  Returns pre-clip total L2 norm. */
float orc_clip_adamw(float* params, float* grads, float* m, float* v, int64_t n_params,
                     float lr, float beta1, float beta2, float eps, float weight_decay,
                     int64_t step, float max_grad_norm);

/* explained variance (PPO/PPO_Discrete.cpp:647-648), unbiased variances */
float orc_explained_variance(int64_t n, const float* returns, const float* values);

#ifdef __cplusplus
}
#endif
#endif
