// rollout.cu — fused rollout: T x (actor-critic forward + Categorical sample + env step + rollout
// store) + bootstrap value in ONE launch, followed by the GAE scan (gae.cu).
// Replaces the rollout phase of PPO_Discrete::train (PPO/PPO_Discrete.cpp:524-554): per step
// ~25 LibTorch op dispatches, a D2H of the actions, a ThreadPool barrier and 3*N tiny H2D copies.
//
// The policy is frozen during a rollout and envs are independent, so a CTA owns a tile of TM envs
// for all T steps: env state, episode counters and the done flag stay in registers; the tile's
// observations feed the shared-memory MLP directly; nothing but the [T,N,...] rollout rows touches HBM.
#include "env.cuh"
#include "policy.cuh"

namespace ppo {

struct RollArgs {
    const float* params;
    Dims d;
    ppo_rollout_cfg cfg;
    float* state;
    int32_t* ep_len;
    float* ep_ret;
    float* next_done;
    float* next_value;
    float* obs_buf;
    int32_t* actions_buf;
    float* logprobs_buf;
    float* rewards_buf;
    float* dones_buf;
    float* values_buf;
    uint8_t* mask_buf;
    double* ep_stats;
};

template <class Env, int TM, int H>
__global__ void __launch_bounds__(MLP_THREADS) rollout_kernel(const RollArgs a) {
    extern __shared__ __align__(16) float sm[];
    constexpr int S = Env::S, O = Env::OBS;
    static_assert(S == O, "observation = state for the built-in envs");
    const int K = a.d.K, heads = a.d.heads;
    const TileLayout L = tile_layout(TM, H, O, K);
    const NetOff co = critic_off(O, H), ao = actor_off(O, H, K);
    const int n = a.cfg.n_envs, T = a.cfg.n_steps;
    const bool masked = a.cfg.masked != 0;
    const int t = threadIdx.x;
    double fin_cnt = 0.0, fin_len = 0.0, fin_ret = 0.0;
    const uint64_t ctr0 = a.cfg.counter + (a.cfg.d_counter ? *a.cfg.d_counter : 0ull);

    const int ntiles = (n + TM - 1) / TM;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int e0 = tile * TM;
        const int rows = min(TM, n - e0);
        const int e = e0 + t;
        const bool on = t < rows;
        float s[S];
        int32_t len = 0;
        float ret = 0.0f, nd = 0.0f;
        if (on) {
#pragma unroll
            for (int k = 0; k < S; k++) s[k] = a.state[(size_t)k * n + e];
            len = a.ep_len[e];
            ret = a.ep_ret[e];
            nd = a.next_done[e];
        } else {
#pragma unroll
            for (int k = 0; k < S; k++) s[k] = 0.0f;
        }
        __syncthreads();
        for (int step = 0; step <= T; step++) {
            // m_obs[step] = next_obs; m_dones[step] = next_done   (:529-530)
            if (t < TM) {
#pragma unroll
                for (int k = 0; k < O; k++) sm[L.xs + t * O + k] = s[k];
                if (on && step < T) {
                    const size_t row = (size_t)step * n + e;
                    if constexpr (O == 4)
                        *reinterpret_cast<float4*>(a.obs_buf + row * 4) = make_float4(s[0], s[1], s[2], s[3]);
                    else if constexpr (O == 2)
                        *reinterpret_cast<float2*>(a.obs_buf + row * 2) = make_float2(s[0], s[1]);
                    else
                        for (int k = 0; k < O; k++) a.obs_buf[row * O + k] = s[k];
                    a.dones_buf[row] = nd;
                }
            }
            __syncthreads();
            if (step == T) {  // bootstrap: next_value = critic(next_obs)  (calcAdvantage :280)
                net_forward<TM, H>(sm, L, O, a.params, co, 1, sm + L.vs, 1);
                if (on) {
                    a.next_value[e] = sm[L.vs + t];
                    a.next_done[e] = nd;
                }
                break;
            }
            net_forward<TM, H>(sm, L, O, a.params, ao, K, sm + L.ls, L.ldo);  // computeActionLogic (:534)
            net_forward<TM, H>(sm, L, O, a.params, co, 1, sm + L.vs, 1);
            if (on) {
                const size_t row = (size_t)step * n + e;
                const uint32_t env_id = a.cfg.env_id_base + (uint32_t)e;
                const uint64_t ctr = ctr0 + (uint64_t)step;
                const uint64_t mbits = masked ? (uint64_t)Env::action_mask(s) : ~0ull;
                int32_t acts[PPO_MAX_HEADS];
                const RowOut o = row_eval(sm + L.ls + t * L.ldo, heads, a.d.nvec, mbits, masked, nullptr, nullptr,
                                          a.cfg.seed, env_id, ctr, a.cfg.entropy_mode, acts);
                for (int h = 0; h < heads; h++) a.actions_buf[row * heads + h] = acts[h];
                a.logprobs_buf[row] = o.logp;
                a.values_buf[row] = sm[L.vs + t];
                if (a.mask_buf)
                    for (int j = 0; j < K; j++) a.mask_buf[row * K + j] = (uint8_t)((mbits >> j) & 1ull);
                // stepEnvs (:413-483): step, truncate, auto-reset, episode statistics
                float u[4] = {0.f, 0.f, 0.f, 0.f};
                float reward, fr;
                int32_t fl;
                StepOut so = Env::step(s, acts[0]);
                len += 1;
                ret = __fadd_rn(ret, so.reward);
                const bool done = so.terminated || (len == a.cfg.max_episode_steps);
                reward = so.reward;
                if (done) {
                    fl = len; fr = ret;
                    fin_cnt += 1.0; fin_len += (double)fl; fin_ret += (double)fr;
                    reset_uniforms(a.cfg.seed, env_id, ctr, u);
                    Env::reset(s, u);
                    len = 0; ret = 0.0f;
                }
                a.rewards_buf[row] = reward;
                nd = done ? 1.0f : 0.0f;
            }
            // the next iteration's write to xs/ls/vs happens after the barriers inside net_forward
        }
        if (on) {
#pragma unroll
            for (int k = 0; k < S; k++) a.state[(size_t)k * n + e] = s[k];
            a.ep_len[e] = len;
            a.ep_ret[e] = ret;
        }
        __syncthreads();
    }
    // finished-episode statistics of this CTA -> global accumulators (one atomic triple per CTA)
    if (a.ep_stats) {
        __syncthreads();
        double* red = reinterpret_cast<double*>(sm);
        if (t < TM) { red[t] = fin_cnt; red[TM + t] = fin_len; red[2 * TM + t] = fin_ret; }
        __syncthreads();
        if (t < 3) {
            double acc = 0.0;
            for (int m = 0; m < TM; m++) acc += red[t * TM + m];
            if (acc != 0.0) atomicAdd(a.ep_stats + t, acc);
        }
    }
}

template <class Env, int TM, int H>
static int launch_rollout(const RollArgs& a, cudaStream_t st) {
    const TileLayout L = tile_layout(TM, H, Env::OBS, a.d.K);
    const size_t smem = (size_t)L.end * sizeof(float);
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(rollout_kernel<Env, TM, H>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      200 * 1024));
        attr_done = true;
    }
    PPO_REQUIRE(smem <= 200 * 1024, "rollout: tile needs %zu B of shared memory", smem);
    const int ntiles = (a.cfg.n_envs + TM - 1) / TM;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int grid = ntiles < 4 * sms ? ntiles : 4 * sms;
    rollout_kernel<Env, TM, H><<<grid, MLP_THREADS, smem, st>>>(a);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

template <class Env>
static int dispatch_rollout(const RollArgs& a, cudaStream_t st) {
    PPO_REQUIRE(a.d.O == Env::OBS, "rollout: obs_dim=%d but the env observes %d floats", a.d.O, Env::OBS);
    switch (a.d.H) {
        case 64: return launch_rollout<Env, 32, 64>(a, st);
        case 128: return launch_rollout<Env, 32, 128>(a, st);
        case 256: return launch_rollout<Env, 32, 256>(a, st);
    }
    set_error("hidden=%d unsupported", a.d.H);
    return PPO_ERR_UNSUPPORTED;
}

}  // namespace ppo

using namespace ppo;

extern "C" int ppo_rollout(const float* params, const ppo_dims* dims, const ppo_rollout_cfg* cfg, float* state_soa,
                           int32_t* ep_len, float* ep_ret, float* next_done, float* next_value, float* obs_buf,
                           int32_t* actions_buf, float* logprobs_buf, float* rewards_buf, float* dones_buf,
                           float* values_buf, float* adv_buf, float* ret_buf, uint8_t* mask_buf, double* ep_stats,
                           ppo_stream_t stream) {
    RollArgs a{};
    if (int rc = make_dims(dims, &a.d)) return rc;
    PPO_REQUIRE(cfg && params && state_soa && ep_len && ep_ret && next_done && next_value && obs_buf && actions_buf &&
                    logprobs_buf && rewards_buf && dones_buf && values_buf,
                "ppo_rollout: null buffer");
    PPO_REQUIRE(cfg->n_envs >= 0 && cfg->n_steps >= 0, "ppo_rollout: negative shape");
    if (cfg->n_envs == 0) return PPO_OK;
    a.params = params; a.cfg = *cfg; a.state = state_soa; a.ep_len = ep_len; a.ep_ret = ep_ret;
    a.next_done = next_done; a.next_value = next_value; a.obs_buf = obs_buf; a.actions_buf = actions_buf;
    a.logprobs_buf = logprobs_buf; a.rewards_buf = rewards_buf; a.dones_buf = dones_buf; a.values_buf = values_buf;
    a.mask_buf = mask_buf; a.ep_stats = ep_stats;
    cudaStream_t st = as_stream(stream);
    int rc;
    if (cfg->env_kind == PPO_ENV_CARTPOLE) rc = dispatch_rollout<CartPoleEnv>(a, st);
    else if (cfg->env_kind == PPO_ENV_MOUNTAINCAR) rc = dispatch_rollout<MountainCarEnv>(a, st);
    else { set_error("ppo_rollout: unknown env kind %d", cfg->env_kind); return PPO_ERR_ARG; }
    if (rc) return rc;
    if (adv_buf && ret_buf && cfg->n_steps > 0)
        return ppo_gae(cfg->n_steps, cfg->n_envs, rewards_buf, values_buf, dones_buf, next_value, next_done,
                       cfg->gamma, cfg->gae_lambda, cfg->use_gae, adv_buf, ret_buf, stream);
    return PPO_OK;
}
