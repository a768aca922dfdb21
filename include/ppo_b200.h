/* include/ppo_b200.h — C ABI of the B200-native PPO hot path (libppo_b200.so).
 *
 * This is the drop-in boundary underneath the reference's C++ class surface
 * (Agent / PPO_Discrete / PPO_MultiDiscrete / Categorical / CategoricalMasked and the
 * duck-typed env interface).  The reference has no FFI of its own; each entry point below
 * names the reference code (file:line under AidanShipperley/PPO-LibTorch) it replaces.
 *
 * Conventions
 *  - Plain C: pointers and sizes only, no torch types.
 *  - Every pointer is a DEVICE pointer unless its name starts with `h_` (host).
 *  - Every launch function takes the stream to launch on (a cudaStream_t passed as void*),
 *    returns 0 on success / non-zero on error, never throws, and never synchronises.
 *    ppo_last_error() returns a thread-local description of the last failure.
 *  - The caller owns all buffers.  The ppo_trainer_* family owns its own device arena.
 *  - There is NO CPU fallback: without a CUDA device every launch function fails with
 *    PPO_ERR_CUDA.
 *
 * Layouts (all fp32 unless noted)
 *  - params / grads / Adam moments: one flat slab in the reference's parameter order
 *    (PPO/Agent.cpp:41-66): critic {W1[H,O], b1[H], W2[H,H], b2[H], W3[1,H], b3[1]} then
 *    actor {W1[H,O], b1[H], W2[H,H], b2[H], W3[K,H], b3[K]}, weights [out,in] row-major.
 *  - env state: SoA, state[s*n + i] for state component s of env i (CartPole S=4:
 *    x, x_dot, theta, theta_dot; MountainCar S=2: position, velocity).
 *  - rollout buffers: time-major like the reference (PPO/PPO_Discrete.cpp:90-95):
 *    obs [T,N,O], actions int32 [T,N,heads], logprobs/rewards/dones/values/adv/ret [T,N].
 *  - masks: uint8 [rows, sumK] (non-zero = valid), NULL = unmasked Categorical.
 */
#ifndef PPO_B200_H
#define PPO_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PPO_B200_ABI_VERSION 4
#define PPO_MAX_HEADS 8
#define PPO_MAX_LOGITS 64

enum { PPO_OK = 0, PPO_ERR_ARG = 1, PPO_ERR_CUDA = 2, PPO_ERR_UNSUPPORTED = 3, PPO_ERR_COMM = 4 };
enum { PPO_ENV_CARTPOLE = 0, PPO_ENV_MOUNTAINCAR = 1, PPO_ENV_MULTIWALK = 2,
       PPO_ENV_HOST = 1000 /* environments stepped by the caller on the host (ppo_trainer_update) */ };
/* PPO_ENTROPY_REFERENCE reproduces Distributions/Categorical.cpp:114-117 bit-for-bit in effect
 * (clamp min = FLT_MIN => entropy == -FLT_MIN*sum(p), zero gradient) for UNMASKED heads;
 * masked heads always use the correct formula (CategoricalMasked.cpp:140-142). */
enum { PPO_ENTROPY_REFERENCE = 0, PPO_ENTROPY_EXACT = 1 };

typedef void* ppo_stream_t; /* cudaStream_t */

typedef struct {
    int32_t obs_dim;             /* O  (m_obs_size)                                  */
    int32_t hidden;              /* H  (64 in PPO/Agent.cpp:25-37; 256 for config C) */
    int32_t heads;               /* number of categorical heads (1 for PPO_Discrete) */
    int32_t nvec[PPO_MAX_HEADS]; /* logits per head; sumK <= PPO_MAX_LOGITS          */
} ppo_dims;

typedef struct {
    float clip_coef;      /* m_clip_coef      */
    int32_t norm_adv;     /* m_norm_adv       */
    int32_t clip_vloss;   /* m_clip_vloss     */
    float ent_coef;       /* m_ent_coef       */
    float vf_coef;        /* m_vf_coef        */
    int32_t entropy_mode; /* PPO_ENTROPY_*    */
} ppo_hyper;

/* Arithmetic of the MLP GEMMs in the fused update kernel (K5).  The reference's switch for this is
 * at::globalContext().setAllowTF32CuBLAS(true) (PPO/PPO_Discrete.cpp:56): reduced-precision tensor-core operands with
 * fp32 accumulation on CUDA, plain fp32 on CPU.  PPO_MATH_FP16 (default; 11 significant bits like TF32, conversions saturate) / PPO_MATH_BF16: tcgen05 tensor cores,
 * 16-bit operands, fp32 accumulate, master weights / activations' non-GEMM math / Adam in fp32; PPO_MATH_FP32: fp32 FMA
 * kernels (the parity mode: matches the oracle to 1e-4).  Shapes outside the tensor-core kernel's envelope
 * (hidden != 64, obs_dim > 7, sum of logits > 32) use the FMA kernels in every mode.  Process-wide, like the
 * reference's; a captured CUDA graph keeps the mode it was captured with. */
enum { PPO_MATH_FP32 = 0, PPO_MATH_BF16 = 1, PPO_MATH_FP16 = 2 };
int ppo_set_math_mode(int mode);
int ppo_get_math_mode(void);
/* Which kernels serve ppo_update_minibatch / the trainer's update phase for this shape in the current math mode:
 * 0 = fp32 FMA kernels, 1 = tcgen05 fused kernel (hidden = 64), 2 = tcgen05 per-layer pipeline (hidden = 256). */
int ppo_update_path(const ppo_dims* dims);

/* Upper bound (milliseconds of SM clock, 0 = default ~20 s) of every in-kernel wait on another CTA or another rank.  A
 * wait that runs out sets the trainer's sticky error flag and SKIPS the optimiser step on that rank; the next
 * synchronised ppo_trainer_iterate returns PPO_ERR_COMM. */
int ppo_set_spin_timeout_ms(int64_t ms);

const char* ppo_last_error(void);
int ppo_abi_version(void);
/* number of visible CUDA devices (0 => nothing below can run) */
int ppo_device_count(void);
/* sm count / compute capability of the current device */
int ppo_device_info(int* sm_count, int* cc_major, int* cc_minor);
/* PPO/Agent.cpp:25-59: parameter count of critic + actor */
int64_t ppo_num_params(const ppo_dims* dims);
int ppo_env_state_dim(int kind); /* 4 / 2 / 4 */
int ppo_env_num_actions(int kind); /* 2 / 3 / 3 (logits of the first head) */
int ppo_env_num_heads(int kind);   /* sub-actions an env consumes per step: 1 / 1 / 8 */
/* The environment registry (csrc/env.cuh: PPO_ENV_LIST): the reference swaps environments by changing the element type of
 * m_envs (PPO/PPO_Discrete.h:106, README "Custom Environment Framework"); here an environment is a device functor with a
 * registered name, and `[b200] env = "<name>"` in PPOConfig.toml picks it.  -1 / NULL for an unknown name / kind. */
int ppo_env_kind_from_name(const char* name);
const char* ppo_env_name(int kind);

/* ---- K3: vectorised envs -------------------------------------------------------------- */
/* Environments/CartPole.cpp:34-45, MountainCar.cpp:60-67 + PPO_Discrete::initEnvs (:365-402).
 * uniforms: optional [n,4] canonical uniforms in [0,1) (parity mode; MountainCar uses col 0);
 * NULL => Philox4x32-10 keyed by (seed, env index, counter).  obs_out [n,S] optional. */
int ppo_env_reset(int kind, int n, float* state_soa, int32_t* ep_len, float* ep_ret,
                  const float* uniforms, uint64_t seed, uint64_t counter, float* obs_out,
                  ppo_stream_t stream);
/* CartPole::step (CartPole.cpp:47-94) / MountainCar::step (MountainCar.cpp:29-57) fused with
 * PPO_Discrete::stepEnvs (PPO_Discrete.cpp:413-483): truncation at max_episode_steps (:443),
 * actions int32 [n, ppo_env_num_heads(kind)] (one sub-action per env for the reference's two environments),
 * auto-reset on done (:453-458), episode statistics.  done_out is 0/1 fp32 as stored in
 * m_dones; fin_len/fin_ret (optional) hold the finished episode's length/return, else 0. */
int ppo_env_step(int kind, int n, float* state_soa, int32_t* ep_len, float* ep_ret,
                 const int32_t* actions, int max_episode_steps, const float* reset_uniforms,
                 uint64_t seed, uint64_t counter, float* obs_out, float* reward_out,
                 float* done_out, int32_t* fin_len, float* fin_ret, ppo_stream_t stream);

/* ---- K2: Categorical / CategoricalMasked ---------------------------------------------- */
/* Distributions/Categorical.cpp:28-39,73-79,92-101,112-119; CategoricalMasked.cpp:31-46,
 * 87-94,107-116,127-144; multi-head sum PPO/Agent.cpp:137-170.
 * logits [n,sumK]; mask u8 [n,sumK] or NULL; actions_in int32 [n,heads] or NULL (=> sample).
 * Sampling is inverse-CDF on one uniform per (row, head): from `uniforms` [n,heads] when
 * given (parity), else Philox(seed, row, counter).  probs_out [n,sumK] optional. */
int ppo_categorical(int n, int heads, const int32_t* h_nvec, const float* logits,
                    const uint8_t* mask, const int32_t* actions_in, const float* uniforms,
                    uint64_t seed, uint64_t counter, int entropy_mode, int32_t* actions_out,
                    float* logprob_out, float* entropy_out, float* probs_out,
                    ppo_stream_t stream);

/* The distribution classes' public members (Distributions/Categorical.cpp:33-35, CategoricalMasked.cpp:35-40):
 * logits_norm_out [n,sumK] = x' - logsumexp(x') per head, probs_out [n,sumK] = softmax(x') per head, x' = where(mask, x,
 * -1e8).  Either output may be NULL. */
int ppo_categorical_normalize(int n, int heads, const int32_t* h_nvec, const float* logits, const uint8_t* mask,
                              float* logits_norm_out, float* probs_out, ppo_stream_t stream);

/* ---- K1: fused actor-critic forward ------------------------------------------------------ */
/* Agent::getActionAndValueDiscrete / getActionAndValueMasked (PPO/Agent.cpp:117-128,137-170)
 * = actor MLP + critic MLP + K2 in one launch.  obs [n,O].  actions_in NULL => sample.
 * Optional outputs may be NULL (entropy_out, logits_out [n,sumK]). */
int ppo_policy_forward(const float* params, const ppo_dims* dims, int n, const float* obs,
                       const uint8_t* mask, const int32_t* actions_in, const float* uniforms,
                       uint64_t seed, uint64_t counter, int entropy_mode, int32_t* actions_out,
                       float* logprob_out, float* entropy_out, float* value_out,
                       float* logits_out, ppo_stream_t stream);
/* Agent::getValue (PPO/Agent.cpp:107-109) */
int ppo_value_forward(const float* params, const ppo_dims* dims, int n, const float* obs,
                      float* value_out, ppo_stream_t stream);

/* ---- K4: GAE / returns ------------------------------------------------------------------- */
/* PPO_Discrete::calcAdvantage (PPO/PPO_Discrete.cpp:289-305 GAE, :309-329 plain returns).
 * rewards/values/dones [T,n] time-major; next_value/next_done [n]. */
int ppo_gae(int T, int n, const float* rewards, const float* values, const float* dones,
            const float* next_value, const float* next_done, float gamma, float gae_lambda,
            int use_gae, float* adv_out, float* ret_out, ppo_stream_t stream);

/* ---- fused rollout: T x (K1 + K2 + K3 + rollout store) + bootstrap in one launch, then K4 ---- */
/* PPO_Discrete::train rollout phase (PPO/PPO_Discrete.cpp:524-554) + calcAdvantage (:274-331).
 * Every env runs its own T steps (the policy is frozen during a rollout, so no grid-wide step
 * barrier is needed).
 * In/out: env state, ep_len, ep_ret, next_done [n] (done flag entering / leaving the rollout).
 * Out: next_value [n] (critic(next_obs), :280), obs [T,n,O], actions [T,n,heads],
 * logprobs/rewards/dones/values [T,n]; adv/ret [T,n] when both non-NULL (the GAE launch);
 * mask_buf u8 [T,n,sumK] optional; episode statistics accumulated into ep_stats[3] =
 * {finished episodes, sum of lengths, sum of returns} (fp64, caller zeroes). */
typedef struct {
    int32_t env_kind;
    int32_t n_envs;
    int32_t n_steps;
    int32_t max_episode_steps;
    float gamma;
    float gae_lambda;
    int32_t use_gae;
    int32_t entropy_mode;
    int32_t masked;       /* 1 => CategoricalMasked with the env's action mask (PPO_MultiDiscrete) */
    uint32_t env_id_base; /* global index of env 0 (rank * n_envs under data parallelism): RNG key */
    uint64_t seed;
    uint64_t counter; /* advances by n_steps per rollout */
    const uint64_t* d_counter; /* optional DEVICE scalar added to `counter` (CUDA-graph replay) */
    /* optional episode log [T,n] (both or neither): length / return of the episode that FINISHED at (t, env) - by
     * termination or truncation, before the auto-reset - and 0 / 0 where none did: the values stepEnvs pushes into its
     * CircularBuffer(100) (PPO/PPO_Discrete.cpp:453-458, 474-480), in the same (step, env) order. */
    int32_t* fin_len;
    float* fin_ret;
} ppo_rollout_cfg;
int ppo_rollout(const float* params, const ppo_dims* dims, const ppo_rollout_cfg* cfg,
                float* state_soa, int32_t* ep_len, float* ep_ret, float* next_done,
                float* next_value, float* obs_buf, int32_t* actions_buf, float* logprobs_buf, float* rewards_buf,
                float* dones_buf, float* values_buf, float* adv_buf, float* ret_buf,
                uint8_t* mask_buf, double* ep_stats, ppo_stream_t stream);

/* The newest (at most max_n <= 1024) finished episodes of an episode log, oldest first: what the reference's
 * CircularBuffer(max_n) holds after pushing the whole log (PPO/PPO_Discrete.cpp:474-480; Utils.h:30-79).  One CTA walks
 * the log backwards from (T-1, n-1) until max_n episodes are found.  out_count: device int32; out_len / out_ret:
 * device [max_n] (entries past *out_count are untouched). */
int ppo_episode_tail(const int32_t* fin_len, const float* fin_ret, int64_t rows, int max_n, int32_t* out_count,
                     int32_t* out_len, float* out_ret, ppo_stream_t stream);

/* ---- K5: fused minibatch forward + PPO loss + backward ----------------------------------- */
/* PPO/PPO_Discrete.cpp:591-594: per-minibatch advantage mean / unbiased std for the n_mb
 * consecutive minibatches of `mb_size` rows of one epoch.  perm (int32 [batch]) NULL =>
 * identity.  stats_out [n_mb,2] = {mean, std}.  partial_out != NULL => also (or only, if
 * stats_out == NULL) write raw double partials [n_mb,3] = {count, sum, sum of squares} for a
 * data-parallel allreduce; ppo_adv_stats_finalize turns (reduced) partials into stats.
 * workspace: optional scratch of ppo_adv_stats_workspace_bytes(n_mb) bytes, zeroed once by the caller (left zeroed);
 * with it, large minibatches are reduced by 32 CTAs each (raw moments in double) instead of one. */
int64_t ppo_adv_stats_workspace_bytes(int n_mb);
int ppo_adv_stats(const float* adv, const int32_t* perm, int64_t batch, int mb_size, int n_mb,
                  float* stats_out, double* partial_out, void* workspace, ppo_stream_t stream);
int ppo_adv_stats_finalize(const double* partial, int n_mb, float* stats_out, ppo_stream_t stream);
/* Packed rollout rows for the minibatch gather (no reference counterpart: the reference indexes six tensors with
 * mbInds, PPO/PPO_Discrete.cpp:577-594).  rec_out [batch][16] floats, 64-byte aligned:
 *   {obs[0..4) zero padded} {old_logprob, adv, ret, old_val} {action bits, 0, 0, 0} {0, 0, 0, 0}
 * A permuted row then costs ONE aligned 64-byte DRAM burst instead of six scattered 32-byte sectors (measured at
 * 16.8 M rows: 635 B of DRAM traffic per row before, profiles/r01j).  Single action head, obs_dim <= 4; old_val may
 * be NULL (zeros). */
int ppo_pack_rows(int64_t batch, int obs_dim, const float* obs, const int32_t* actions, const float* old_logprob,
                  const float* adv, const float* ret, const float* old_val, float* rec_out, ppo_stream_t stream);
/* The same statistics for all n_epochs x n_mb minibatches of one PPO update in ONE launch (the advantages are fixed
 * during the update, PPO/PPO_Discrete.cpp:567-594, and the permutation of epoch e is ppo_make_permutation(batch, seed,
 * epoch0 + e [+ *d_epoch]) evaluated inline, no index buffer).  stats_out [n_epochs*n_mb, 2] and / or partial_out
 * [n_epochs*n_mb, 3]; workspace: ppo_adv_stats_workspace_bytes(n_epochs * n_mb) bytes, zeroed once. */
int ppo_adv_stats_all_epochs(const float* adv, int64_t batch, int mb_size, int n_mb, int n_epochs, uint64_t seed,
                             uint64_t epoch0, const uint64_t* d_epoch, float* stats_out, double* partial_out,
                             void* workspace, ppo_stream_t stream);
/* Scratch the update kernel needs: returns bytes for `workspace`. */
int64_t ppo_update_workspace_bytes(const ppo_dims* dims);
/* PPO/PPO_Discrete.cpp:579-638 for the minibatch rows perm[mb_start .. mb_start+mb_size):
 * forward with the stored actions, ratio, clipped surrogate, (clipped) value loss, entropy
 * bonus, closed-form backward (SURVEY.md §8 a-bis).  `inv_n` scales sample sums into the
 * minibatch mean: pass 1/mb_size (single GPU) or 1/global_mb_size (data parallel, followed by
 * an allreduce(sum) of grads_out and diag_out).  adv_stats {mean,std} for this minibatch
 * (ignored unless norm_adv).  grads_out: flat parameter order, PRE-clip, overwritten.
 * diag_out[8] = {pg_loss, v_loss, entropy_mean, loss, approx_kl, clipfrac, 0, 0} (overwritten). */
int ppo_update_minibatch(const float* params, const ppo_dims* dims, const float* obs,
                         const uint8_t* mask, const int32_t* actions, const float* old_logprob,
                         const float* adv, const float* ret, const float* old_val,
                         const int32_t* perm, int64_t mb_start, int mb_size, float inv_n,
                         const float* adv_stats, const ppo_hyper* h_hyper, void* workspace,
                         float* grads_out, float* diag_out, ppo_stream_t stream);

/* ---- K6: clip_grad_norm_ + AdamW ----------------------------------------------------------- */
/* PPO/PPO_Discrete.cpp:640-641; torch/nn/utils/clip_grad.h; torch optim adamw.cpp.
 * grads are multiplied by grad_scale first (1 for single GPU).  step = 1-based step count;
 * when d_step != NULL the DEVICE scalar is added to it.  lr is read from the device scalar
 * d_lr when non-NULL (both let a captured CUDA graph replay with fresh values).
 * norm_out (optional, 1 float) receives the pre-clip total L2 norm. */
int ppo_clip_adamw(float* params, float* grads, float* exp_avg, float* exp_avg_sq,
                   int64_t n_params, float lr, const float* d_lr, float beta1, float beta2,
                   float eps, float weight_decay, int64_t step, const int64_t* d_step,
                   float max_grad_norm, float grad_scale, float* norm_out, ppo_stream_t stream);

/* PPO/PPO_Discrete.cpp:647-648: 1 - var(R - V)/var(R), unbiased.  out: 1 float.
 * workspace: ppo_explained_variance_workspace_bytes() bytes, zeroed once by the caller (the
 * kernel leaves it zeroed). */
int64_t ppo_explained_variance_workspace_bytes(void);
int ppo_explained_variance(int64_t n, const float* returns, const float* values, void* workspace,
                           float* out, ppo_stream_t stream);
/* No counterpart in the reference: added for callers that cannot allocate
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The same kernel and the
 * same result as ppo_explained_variance, with the scratch (per-CTA partial moments + arrival ticket, 32 784 bytes)
 * in a file-scope static __device__ array of the library: every GPU that loads the library holds it until exit, and
 * the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine; one 16-byte cudaMemsetAsync of
 * the ticket + one kernel on `stream`).  Limit: n <= PPO_EV_STATIC_MAX_N rows, larger n => PPO_ERR_ARG with a message
 * (the workspace variant above has no such limit).  ALL calls on a device share the one array: calls that overlap
 * on the GPU (different streams, no dependency between them) race on it — serialise them. */
#define PPO_EV_STATIC_MAX_N ((int64_t)1 << 28)
int64_t ppo_explained_variance_static_max_n(void);
int ppo_explained_variance_static(int64_t n, const float* returns, const float* values, float* out,
                                  ppo_stream_t stream);

/* ---- minibatch permutation (replaces torch::randperm, PPO/PPO_Discrete.cpp:569) ----------- */
/* Counter-based bijection on [0,batch): perm_out[i] = pi_{seed,epoch}(i).  d_epoch: optional
 * DEVICE scalar added to `epoch` (CUDA-graph replay). */
int ppo_make_permutation(int64_t batch, uint64_t seed, uint64_t epoch, const uint64_t* d_epoch,
                         int32_t* perm_out, ppo_stream_t stream);

/* ---- host-side trainer: PPO_Discrete / PPO_MultiDiscrete::train() loop body ---------------- */
/* Owns one device arena (params, grads, moments, env state, rollout slabs) and runs one PPO
 * update iteration (PPO/PPO_Discrete.cpp:511-673 minus printing/checkpoint) without host
 * round trips.  All pointers in this family are HOST pointers. */
typedef struct {
    ppo_dims dims;
    int32_t env_kind;
    int32_t num_envs;           /* m_num_envs (per GPU)          */
    int32_t num_steps;          /* m_num_steps                   */
    int32_t num_minibatches;    /* m_num_minibatches             */
    int32_t update_epochs;      /* m_update_epochs               */
    int32_t max_episode_steps;  /* m_max_episode_steps           */
    int32_t use_gae;
    float gamma;
    float gae_lambda;
    ppo_hyper hyper;
    float max_grad_norm;
    float beta1, beta2, adam_eps, weight_decay; /* AdamW(lr).eps(1e-5), wd 1e-2 (:76-78) */
    uint64_t seed;
    int32_t world_size;         /* data-parallel ranks (1 = single GPU) */
    int32_t rank;
    int32_t use_cuda_graph;     /* capture one whole iteration into a CUDA graph and replay it */
    int32_t masked;             /* PPO_MultiDiscrete: CategoricalMasked heads + [T,N,sumK] mask buffer */
} ppo_trainer_cfg;

typedef struct {
    float pg_loss, v_loss, entropy, loss, approx_kl, clipfrac; /* last minibatch, clipfrac = mean over update */
    float explained_variance;
    float grad_norm;
    double episodes, ep_len_sum, ep_ret_sum; /* finished during this rollout */
    float lr;
    int32_t pad;
} ppo_iter_result;

typedef struct ppo_trainer ppo_trainer;
ppo_trainer* ppo_trainer_create(const ppo_trainer_cfg* h_cfg);
void ppo_trainer_destroy(ppo_trainer* t);
int ppo_trainer_set_params(ppo_trainer* t, const float* h_params);
int ppo_trainer_get_params(ppo_trainer* t, float* h_params);
/* optimizer state for checkpointing: exp_avg, exp_avg_sq, step */
int ppo_trainer_get_opt_state(ppo_trainer* t, float* h_exp_avg, float* h_exp_avg_sq, int64_t* h_step);
/* After a synchronised ppo_trainer_iterate(want_result) / ppo_trainer_finish: the newest finished episodes of that
 * iteration, oldest first, at most min(max_n, PPO_EPISODE_TAIL) of them -> host arrays; returns how many.  Pushing them
 * into a CircularBuffer(100) reproduces the reference's ring (PPO/PPO_Discrete.cpp:474-480) exactly: the ring only ever
 * holds the last 100 episodes, so older ones of the same iteration could not have survived. */
#define PPO_EPISODE_TAIL 100
int ppo_trainer_last_episodes(ppo_trainer* t, int max_n, int32_t* h_len, float* h_ret);
int ppo_trainer_set_opt_state(ppo_trainer* t, const float* h_exp_avg, const float* h_exp_avg_sq, int64_t step);
/* initEnvs (PPO_Discrete.cpp:365-402) */
int ppo_trainer_reset_envs(ppo_trainer* t);
/* device pointers of the arena for zero-copy views / collectives / tests:
 * name in {"params","grads","exp_avg","exp_avg_sq","obs","actions","logprobs","rewards","dones",
 * "values","adv","ret","masks","state","ep_len","ep_ret","next_done","next_value","diag","perm",
 * "adv_stats","ep_stats"} */
void* ppo_trainer_buffer(ppo_trainer* t, const char* name, int64_t* bytes);
/* One PPO iteration on `stream`.  h_lr: learning rate for this update (the anneal of
 * :514-518 is host arithmetic).  h_result is written after the iteration's diagnostics
 * have been copied back (this call synchronises the stream once, at the end) unless
 * h_result is NULL (fully asynchronous). */
int ppo_trainer_iterate(ppo_trainer* t, float h_lr, ppo_iter_result* h_result, ppo_stream_t stream);
/* The part of an iteration behind the rollout, on rollout rows the CALLER filled through ppo_trainer_buffer ("obs",
 * "actions", "logprobs", "rewards", "dones", "values", "next_value", "next_done"; optionally "fin_len" / "fin_ret" /
 * "ep_stats"): calcAdvantage + E x MB optimiser steps + diagnostics (PPO/PPO_Discrete.cpp:556-673).  This is what a
 * host-environment bridge calls (env_kind = PPO_ENV_HOST: PPO_Discrete.cpp:413-483 runs on the host, README "embed in game
 * engines"); eager launches, h_result as in ppo_trainer_iterate. */
int ppo_trainer_update(ppo_trainer* t, float h_lr, ppo_iter_result* h_result, ppo_stream_t stream);
/* Split phases of one iteration, in the order iterate() issues them (for host loops that
 * interleave their own work, e.g. tests that inspect the rollout before the update). */
int ppo_trainer_rollout(ppo_trainer* t, float h_lr, ppo_stream_t stream);
int ppo_trainer_begin_epoch(ppo_trainer* t, int epoch, ppo_stream_t stream);
int ppo_trainer_minibatch_grads(ppo_trainer* t, int mb, ppo_stream_t stream);
int ppo_trainer_apply(ppo_trainer* t, ppo_stream_t stream); /* [allreduce] + clip + AdamW */
int ppo_trainer_finish(ppo_trainer* t, float h_lr, ppo_iter_result* h_result, ppo_stream_t stream);

/* Per-phase device timing for bench.py: profile(1) brackets every phase of later iterations with a
 * cudaEvent pair on its launch stream (CUDA-graph replay is bypassed meanwhile); phase_ms
 * synchronises, writes accumulated milliseconds / launch counts for the 6 phases {rollout, gae,
 * epoch setup (permutation + advantage statistics), update (K5), apply (allreduce + K6), tail} and
 * resets the accumulators. */
int ppo_trainer_profile(ppo_trainer* t, int enable);
int ppo_trainer_phase_ms(ppo_trainer* t, float* h_ms6, int32_t* h_count6);

/* ---- data parallelism (new: the reference is single-device, PPO/PPO_Discrete.cpp:65) ------- */
/* One process per GPU.  Envs shard by rank; parameters and Adam moments are replicated; per
 * minibatch the gradient slab (+8 diagnostics) is allreduced, per epoch the advantage moments.
 * NCCL is resolved with dlopen at first use.  h_id128: 128-byte ncclUniqueId made on rank 0 and
 * distributed by the launcher (bench.py uses torch.distributed for that). */
typedef struct ppo_comm ppo_comm;
int ppo_comm_unique_id(char* h_id128);
ppo_comm* ppo_comm_init(const char* h_id128, int rank, int world_size);
void ppo_comm_destroy(ppo_comm* c);
int ppo_comm_world_size(const ppo_comm* c);
int ppo_comm_rank(const ppo_comm* c);
int ppo_comm_allreduce_f32(ppo_comm* c, float* buf, int64_t count, ppo_stream_t stream);
int ppo_comm_allreduce_f64(ppo_comm* c, double* buf, int64_t count, ppo_stream_t stream);
int ppo_trainer_set_comm(ppo_trainer* t, ppo_comm* comm);
/* Peer-memory gradient exchange over NVLink (replaces reduce -> ncclAllReduce -> clip/AdamW by ONE kernel per rank that
 * publishes its reduced gradient blocks, reads the peers' blocks with P2P loads, sums them in rank order and applies
 * clip + AdamW: csrc/update.cu reduce_peer_apply_kernel).  Each rank exports a 64-byte CUDA IPC handle of its exchange
 * buffer; the launcher gathers the world_size handles (rank order) and hands all of them to every rank.  After
 * ppo_trainer_set_peers, ppo_trainer_iterate uses the peer path (the per-epoch advantage moments still go through the
 * communicator set with ppo_trainer_set_comm).  PPO_ERR_UNSUPPORTED: parameter count too large for the single-wave
 * kernel (the NCCL path stays in use). */
int ppo_trainer_peer_handle(ppo_trainer* t, char* h_handle64);
int ppo_trainer_set_peers(ppo_trainer* t, const char* h_handles, int world_size);

/* Diagnostics, A/B switches and in-kernel timelines (ppo_debug_*) live in ppo_b200_debug.h: they are not part of the
 * drop-in surface. */

#ifdef __cplusplus
}
#endif
#endif /* PPO_B200_H */
