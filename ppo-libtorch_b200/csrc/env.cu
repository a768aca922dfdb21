// env.cu — K3: batched environment reset / step kernels (one thread per VEC envs, SoA state,
// in-kernel truncation + auto-reset + episode statistics).  HBM-bound: 60 B per CartPole
// env-step, 44 B per MountainCar env-step (+ O*4 B for the observation row).
// Replaces PPO_Discrete::initEnvs / stepEnvs + ThreadPool (PPO/PPO_Discrete.cpp:365-483).
#include <string.h>

#include "env.cuh"

namespace ppo {

template <int V> struct VecF;
template <> struct VecF<1> { using type = float; };
template <> struct VecF<4> { using type = float4; };
template <int V> struct VecI;
template <> struct VecI<1> { using type = int; };
template <> struct VecI<4> { using type = int4; };

template <int V, class T> __device__ inline void ldv(const T* p, T (&r)[V]) {
    if constexpr (V == 4) {
        const uint4 v = *reinterpret_cast<const uint4*>(p);
        r[0] = reinterpret_cast<const T&>(v.x); r[1] = reinterpret_cast<const T&>(v.y);
        r[2] = reinterpret_cast<const T&>(v.z); r[3] = reinterpret_cast<const T&>(v.w);
    } else {
        r[0] = *p;
    }
}
template <int V, class T> __device__ inline void stv(T* p, const T (&r)[V]) {
    if constexpr (V == 4) {
        uint4 v;
        v.x = reinterpret_cast<const uint32_t&>(r[0]); v.y = reinterpret_cast<const uint32_t&>(r[1]);
        v.z = reinterpret_cast<const uint32_t&>(r[2]); v.w = reinterpret_cast<const uint32_t&>(r[3]);
        *reinterpret_cast<uint4*>(p) = v;
    } else {
        *p = r[0];
    }
}

template <class Env, int V>
__global__ void __launch_bounds__(256) env_step_kernel(
    int n, int first, float* __restrict__ state, int32_t* __restrict__ ep_len, float* __restrict__ ep_ret,
    const int32_t* __restrict__ actions, int max_episode_steps, const float* __restrict__ reset_u,
    uint64_t seed, uint64_t counter, float* __restrict__ obs_out, float* __restrict__ reward_out,
    float* __restrict__ done_out, int32_t* __restrict__ fin_len, float* __restrict__ fin_ret) {
    constexpr int S = Env::S;
    const int64_t i0 = (int64_t)first + ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * V;
    if (i0 >= n) return;
    float st[S][V];
    constexpr int AH = Env::HEADS;   // sub-actions per env: actions is int32 [n, AH]
    int32_t len[V], act[V][AH];
    float ret[V];
#pragma unroll
    for (int s = 0; s < S; s++) ldv<V>(state + (size_t)s * n + i0, st[s]);
    ldv<V>(ep_len + i0, len);
    ldv<V>(ep_ret + i0, ret);
    if constexpr (AH == 1) {
        int32_t a1[V];
        ldv<V>(actions + i0, a1);
#pragma unroll
        for (int v = 0; v < V; v++) act[v][0] = a1[v];
    } else {
#pragma unroll
        for (int v = 0; v < V; v++)
#pragma unroll
            for (int h = 0; h < AH; h++) act[v][h] = __ldg(actions + (size_t)(i0 + v) * AH + h);
    }
    float rew[V], dn[V], fr[V];
    int32_t fl[V];
#pragma unroll
    for (int v = 0; v < V; v++) {
        float s_[S];
#pragma unroll
        for (int s = 0; s < S; s++) s_[s] = st[s][v];
        // Peek whether this transition ends the episode is only known after the step, so the
        // reset draws are produced lazily inside the branch (Philox) or read (parity mode).
        float u[4] = {0.f, 0.f, 0.f, 0.f};
        StepOut o = Env::step_multi(s_, act[v]);
        len[v] += 1;
        ret[v] = __fadd_rn(ret[v], o.reward);
        const bool done = o.terminated || (len[v] == max_episode_steps);
        rew[v] = o.reward;
        fl[v] = 0; fr[v] = 0.0f;
        if (done) {
            fl[v] = len[v]; fr[v] = ret[v];
            if (reset_u) {
                const float4 uu = *reinterpret_cast<const float4*>(reset_u + (size_t)(i0 + v) * 4);
                u[0] = uu.x; u[1] = uu.y; u[2] = uu.z; u[3] = uu.w;
            } else {
                reset_uniforms(seed, (uint32_t)(i0 + v), counter, u);
            }
            Env::reset(s_, u);
            len[v] = 0; ret[v] = 0.0f;
        }
        dn[v] = done ? 1.0f : 0.0f;
#pragma unroll
        for (int s = 0; s < S; s++) st[s][v] = s_[s];
    }
#pragma unroll
    for (int s = 0; s < S; s++) stv<V>(state + (size_t)s * n + i0, st[s]);
    stv<V>(ep_len + i0, len);
    stv<V>(ep_ret + i0, ret);
    stv<V>(reward_out + i0, rew);
    stv<V>(done_out + i0, dn);
    if (fin_len) stv<V>(fin_len + i0, fl);
    if (fin_ret) stv<V>(fin_ret + i0, fr);
    if (obs_out) {
#pragma unroll
        for (int v = 0; v < V; v++) {
            if constexpr (S == 4) {
                *reinterpret_cast<float4*>(obs_out + (size_t)(i0 + v) * 4) =
                    make_float4(st[0][v], st[1][v], st[2][v], st[3][v]);
            } else {
                *reinterpret_cast<float2*>(obs_out + (size_t)(i0 + v) * 2) = make_float2(st[0][v], st[1][v]);
            }
        }
    }
}

template <class Env>
__global__ void __launch_bounds__(256) env_reset_kernel(int n, float* __restrict__ state,
                                                        int32_t* __restrict__ ep_len, float* __restrict__ ep_ret,
                                                        const float* __restrict__ uniforms, uint64_t seed,
                                                        uint64_t counter, float* __restrict__ obs_out) {
    constexpr int S = Env::S;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float u[4];
    if (uniforms) {
#pragma unroll
        for (int k = 0; k < 4; k++) u[k] = uniforms[(size_t)i * 4 + k];
    } else {
        reset_uniforms(seed, (uint32_t)i, counter, u);
    }
    float s[S];
    Env::reset(s, u);
#pragma unroll
    for (int k = 0; k < S; k++) {
        state[(size_t)k * n + i] = s[k];
        if (obs_out) obs_out[(size_t)i * S + k] = s[k];
    }
    ep_len[i] = 0;
    ep_ret[i] = 0.0f;
}

template <class Env>
static int launch_step(int n, float* state, int32_t* ep_len, float* ep_ret, const int32_t* actions,
                       int max_episode_steps, const float* reset_u, uint64_t seed, uint64_t counter,
                       float* obs_out, float* reward_out, float* done_out, int32_t* fin_len, float* fin_ret,
                       cudaStream_t st) {
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
    const bool vec_ok = (n % 4 == 0) && al16(state) && al16(ep_len) && al16(ep_ret) && al16(actions) &&
                        al16(reward_out) && al16(done_out) && al16(fin_len) && al16(fin_ret) && al16(obs_out) &&
                        al16(reset_u);
    if (vec_ok && n >= 4 && Env::HEADS == 1) {
        const int threads = 256, items = n / 4;
        env_step_kernel<Env, 4><<<(items + threads - 1) / threads, threads, 0, st>>>(
            n, 0, state, ep_len, ep_ret, actions, max_episode_steps, reset_u, seed, counter, obs_out, reward_out,
            done_out, fin_len, fin_ret);
    } else {
        const int threads = 256;
        env_step_kernel<Env, 1><<<(n + threads - 1) / threads, threads, 0, st>>>(
            n, 0, state, ep_len, ep_ret, actions, max_episode_steps, reset_u, seed, counter, obs_out, reward_out,
            done_out, fin_len, fin_ret);
    }
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

}  // namespace ppo

using namespace ppo;

extern "C" int ppo_env_state_dim(int kind) {
    return env_dispatch(kind, [](auto tag) { return (int)decltype(tag)::type::S; }, -1);
}
extern "C" int ppo_env_num_actions(int kind) {
    return env_dispatch(kind, [](auto tag) { return (int)decltype(tag)::type::ACTIONS; }, -1);
}
extern "C" int ppo_env_num_heads(int kind) {
    return env_dispatch(kind, [](auto tag) { return (int)decltype(tag)::type::HEADS; }, -1);
}
// the registry's names: what `[b200] env = "<name>"` in PPOConfig.toml resolves through
extern "C" int ppo_env_kind_from_name(const char* name) {
    if (!name) return -1;
#define PPO_ENV_NAME(id, nm, Env) \
    if (!strcmp(name, nm)) return id;
    PPO_ENV_LIST(PPO_ENV_NAME)
#undef PPO_ENV_NAME
    return -1;
}
extern "C" const char* ppo_env_name(int kind) {
#define PPO_ENV_NAME(id, nm, Env) \
    if (kind == id) return nm;
    PPO_ENV_LIST(PPO_ENV_NAME)
#undef PPO_ENV_NAME
    return nullptr;
}

extern "C" int ppo_env_reset(int kind, int n, float* state_soa, int32_t* ep_len, float* ep_ret,
                             const float* uniforms, uint64_t seed, uint64_t counter, float* obs_out,
                             ppo_stream_t stream) {
    PPO_REQUIRE(n >= 0 && state_soa && ep_len && ep_ret, "ppo_env_reset: null buffer or negative n");
    if (n == 0) return PPO_OK;
    const int threads = 256, blocks = (n + threads - 1) / threads;
    const int rc = env_dispatch(kind, [&](auto tag) {
        using Env = typename decltype(tag)::type;
        env_reset_kernel<Env><<<blocks, threads, 0, as_stream(stream)>>>(n, state_soa, ep_len, ep_ret, uniforms, seed, counter,
                                                                        obs_out);
        return PPO_OK;
    }, PPO_ERR_ARG);
    PPO_REQUIRE(rc == PPO_OK, "ppo_env_reset: unknown env kind %d", kind);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

extern "C" int ppo_env_step(int kind, int n, float* state_soa, int32_t* ep_len, float* ep_ret,
                            const int32_t* actions, int max_episode_steps, const float* reset_uniforms,
                            uint64_t seed, uint64_t counter, float* obs_out, float* reward_out, float* done_out,
                            int32_t* fin_len, float* fin_ret, ppo_stream_t stream) {
    PPO_REQUIRE(n >= 0 && state_soa && ep_len && ep_ret && actions && reward_out && done_out,
                "ppo_env_step: null buffer or negative n");
    if (n == 0) return PPO_OK;
    const int rc = env_dispatch(kind, [&](auto tag) {
        using Env = typename decltype(tag)::type;
        return launch_step<Env>(n, state_soa, ep_len, ep_ret, actions, max_episode_steps, reset_uniforms, seed, counter,
                                obs_out, reward_out, done_out, fin_len, fin_ret, as_stream(stream));
    }, -12345);
    if (rc != -12345) return rc;
    PPO_REQUIRE(false, "ppo_env_step: unknown env kind %d", kind);
    return PPO_ERR_ARG;
}
