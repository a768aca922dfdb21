// env.cuh — device-side environment functors (the B200 form of the reference's duck-typed
// env interface: reset() / step(action) / getActionMask(), Environments/CartPole.h:43-74,
// MountainCar.h:33-46).  A new environment is a struct with the same members, instantiated
// into the K3 kernels (env.cu) and the fused rollout (rollout.cu).
//
// Arithmetic uses explicit round-to-nearest intrinsics (no FMA contraction) so a step is
// the same sequence of fp32 roundings as the reference's x86-64 build; only sinf/cosf can
// differ (<= 2 ulp), which the 1e-6 relative parity bound of BASELINE.json absorbs.
#pragma once
#include "common.cuh"

namespace ppo {

struct StepOut {
    float reward;
    bool terminated;
};

// Environments/CartPole.cpp:6-17 (constants), :34-45 (reset), :47-94 (step)
struct CartPoleEnv {
    static constexpr int S = 4;        // state floats
    static constexpr int OBS = 4;      // observation floats (= state)
    static constexpr int ACTIONS = 2;
    static constexpr int RESET_DRAWS = 4;

    __device__ static inline void reset(float (&s)[S], const float (&u)[4]) {
        // uniform_real_distribution<float>(-0.05f, 0.05f): u*(b-a)+a   (CartPole.cpp:97-101)
        const float a = -0.05f, b = 0.05f;
        const float w = __fsub_rn(b, a);
#pragma unroll
        for (int i = 0; i < 4; i++) s[i] = __fadd_rn(__fmul_rn(u[i], w), a);
    }

    __device__ static inline StepOut step(float (&s)[S], int action) {
        const float gravity = 9.8f, mass_pole = 0.1f, total_mass = 1.1f /* 0.1f + 1.0f */;
        const float length = 0.5f, polemass_length = 0.05f /* 0.1f * 0.5f */;
        const float force_mag = 10.0f, tau = 0.02f;
        const float theta_thr = 0.20943951f;  // float(12*2*pi/360)  (CartPole.cpp:16)
        const float x_thr = 2.4f;
        float x = s[0], x_dot = s[1], theta = s[2], theta_dot = s[3];
        const float force = (action == 0) ? -force_mag : force_mag;
        const float cos_t = cosf(theta), sin_t = sinf(theta);
        // temp = (force + pml * theta_dot^2 * sin) / total_mass
        const float t0 = __fmul_rn(__fmul_rn(__fmul_rn(polemass_length, theta_dot), theta_dot), sin_t);
        const float temp = __fdiv_rn(__fadd_rn(force, t0), total_mass);
        // theta_acc = (g*sin - cos*temp) / (l * (4/3 - mp*cos^2/M))
        const float num = __fsub_rn(__fmul_rn(gravity, sin_t), __fmul_rn(cos_t, temp));
        const float den = __fmul_rn(
            length, __fsub_rn(4.0f / 3.0f, __fdiv_rn(__fmul_rn(__fmul_rn(mass_pole, cos_t), cos_t), total_mass)));
        const float theta_acc = __fdiv_rn(num, den);
        const float x_acc =
            __fsub_rn(temp, __fdiv_rn(__fmul_rn(__fmul_rn(polemass_length, theta_acc), cos_t), total_mass));
        x = __fadd_rn(x, __fmul_rn(tau, x_dot));
        x_dot = __fadd_rn(x_dot, __fmul_rn(tau, x_acc));
        theta = __fadd_rn(theta, __fmul_rn(tau, theta_dot));
        theta_dot = __fadd_rn(theta_dot, __fmul_rn(tau, theta_acc));
        s[0] = x; s[1] = x_dot; s[2] = theta; s[3] = theta_dot;
        StepOut o;
        o.terminated = (x < -x_thr) || (x > x_thr) || (theta < -theta_thr) || (theta > theta_thr);
        o.reward = o.terminated ? -1.0f : 1.0f;  // CartPole.cpp:80-88
        return o;
    }
    __device__ static inline uint32_t action_mask(const float (&)[S]) { return 0x3u; }
    static constexpr int HEADS = 1;   // sub-actions consumed per step
    __device__ static inline StepOut step_multi(float (&s)[S], const int32_t* acts) { return step(s, acts[0]); }
};

// Environments/MountainCar.cpp:6-13 (constants), :29-57 (step), :60-67 (reset), :69-77 (mask)
struct MountainCarEnv {
    static constexpr int S = 2;
    static constexpr int OBS = 2;
    static constexpr int ACTIONS = 3;
    static constexpr int RESET_DRAWS = 1;

    __device__ static inline void reset(float (&s)[S], const float (&u)[4]) {
        const float a = -0.6f, b = -0.4f;
        s[0] = __fadd_rn(__fmul_rn(u[0], __fsub_rn(b, a)), a);
        s[1] = 0.0f;
    }
    __device__ static inline StepOut step(float (&s)[S], int action) {
        const float min_position = -1.2f, max_position = 0.6f, max_speed = 0.07f;
        const float goal_position = 0.5f, goal_velocity = 0.0f, force = 0.001f, gravity = 0.0025f;
        float position = s[0], velocity = s[1];
        const float push = __fmul_rn(__fsub_rn((float)action, 1.0f), force);
        const float grav = __fmul_rn(cosf(__fmul_rn(3.0f, position)), -gravity);
        velocity = __fadd_rn(velocity, __fadd_rn(push, grav));
        velocity = velocity < -max_speed ? -max_speed : (max_speed < velocity ? max_speed : velocity);
        position = __fadd_rn(position, velocity);
        position = position < min_position ? min_position : (max_position < position ? max_position : position);
        if (position == min_position && velocity < 0.0f) velocity = 0.0f;
        s[0] = position; s[1] = velocity;
        StepOut o;
        o.terminated = (position >= goal_position) && (velocity >= goal_velocity);
        o.reward = -1.0f;
        return o;
    }
    __device__ static inline uint32_t action_mask(const float (&)[S]) { return 0x7u; }
    static constexpr int HEADS = 1;
    __device__ static inline StepOut step_multi(float (&s)[S], const int32_t* acts) { return step(s, acts[0]); }
};

// MultiWalk: a synthetic MultiDiscrete environment with a STATE-DEPENDENT action mask.  No reference counterpart (the
// reference's two environments return all-ones masks and its MultiDiscrete trainer feeds a single sub-action to the env,
// SURVEY.md Q9): it exists to exercise the generic masked multi-head path end to end with the action
// space the reference keeps commented at PPO/Agent.cpp:16, nvec = {3, 3, 3, 3, 3, 2, 2, 2} (21 logits, 8 heads).
//   state = four positions in [-1, 1]; heads 0..3 move position i by (a_i - 1) * gain, head 4 picks gain = 0.05 (1 + a_4),
//   heads 5..7 are binary flags worth 0.01 each; reward = flags - 0.25 * sum |pos|, or +1 when sum |pos| < 0.2 (terminal).
//   mask: a move that would leave [-0.95, 0.95] is invalid; the largest gain needs every |pos| <= 0.7; flag h may be
//   raised only while pos[h - 5] >= 0.  Every head always keeps at least one valid option.
// The CPU restatement is oracle/ppo_oracle.c: orc_multiwalk_* (same operations in the same order).
struct MultiWalkEnv {
    static constexpr int S = 4;
    static constexpr int OBS = 4;
    static constexpr int ACTIONS = 3;
    static constexpr int RESET_DRAWS = 4;
    static constexpr int HEADS = 8;

    __device__ static inline void reset(float (&s)[S], const float (&u)[4]) {
#pragma unroll
        for (int i = 0; i < 4; i++) s[i] = __fadd_rn(__fmul_rn(u[i], 1.6f), -0.8f);
    }
    __device__ static inline StepOut step_multi(float (&s)[S], const int32_t* acts) {
        const float gain = __fmul_rn(0.05f, (float)(1 + acts[4]));
        float dist = 0.0f;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            float p = __fadd_rn(s[i], __fmul_rn((float)(acts[i] - 1), gain));
            p = p < -1.0f ? -1.0f : (p > 1.0f ? 1.0f : p);
            s[i] = p;
            dist = __fadd_rn(dist, fabsf(p));
        }
        const float bonus = __fmul_rn(0.01f, (float)(acts[5] + acts[6] + acts[7]));
        StepOut o;
        o.terminated = dist < 0.2f;
        o.reward = o.terminated ? 1.0f : __fsub_rn(bonus, __fmul_rn(0.25f, dist));
        return o;
    }
    __device__ static inline StepOut step(float (&s)[S], int action) {   // single sub-action: the other heads at rest
        const int32_t acts[8] = {action, 1, 1, 1, 0, 0, 0, 0};
        return step_multi(s, acts);
    }
    __device__ static inline uint32_t action_mask(const float (&s)[S]) {
        uint32_t m = 0u;
        bool small = true;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            m |= (s[i] > -0.95f ? 1u : 0u) << (3 * i);
            m |= 1u << (3 * i + 1);
            m |= (s[i] < 0.95f ? 1u : 0u) << (3 * i + 2);
            small = small && fabsf(s[i]) <= 0.7f;
        }
        m |= 3u << 12;
        m |= (small ? 1u : 0u) << 14;
#pragma unroll
        for (int h = 0; h < 3; h++) {
            m |= 1u << (15 + 2 * h);
            m |= (s[h] >= 0.0f ? 1u : 0u) << (16 + 2 * h);
        }
        return m;
    }
};

// ---- the environment registry: ONE list.  A new environment = a functor above (S, OBS, ACTIONS, HEADS, RESET_DRAWS,
// reset, step, step_multi, action_mask) + one X(...) line here + its PPO_ENV_* enumerator in include/ppo_b200.h; every
// dispatch site (env.cu, rollout.cu, rollout_tc.cu, rollout_h256.cu) and the `[b200] env = "<name>"` key go through it.
#define PPO_ENV_LIST(X)                                  \
    X(PPO_ENV_CARTPOLE, "cartpole", CartPoleEnv)         \
    X(PPO_ENV_MOUNTAINCAR, "mountaincar", MountainCarEnv) \
    X(PPO_ENV_MULTIWALK, "multiwalk", MultiWalkEnv)
template <class E>
struct EnvTag {
    using type = E;
};
// f(EnvTag<Env>{}) for the functor registered under `kind`; `otherwise` when the kind is unknown
template <class F>
inline int env_dispatch(int kind, F&& f, int otherwise) {
#define PPO_ENV_CASE(id, name, Env) \
    if (kind == id) return f(EnvTag<Env>{});
    PPO_ENV_LIST(PPO_ENV_CASE)
#undef PPO_ENV_CASE
    return otherwise;
}

// One vectorised transition of env `i` (PPO/PPO_Discrete.cpp:435-466): step, truncate at
// episode_length == max_episode_steps (:443), on done record (len, ret) and reset (:453-458).
// u_reset supplies the reset draws.  Returns done.
template <class Env>
__device__ inline bool env_transition(float (&s)[Env::S], int32_t& ep_len, float& ep_ret, int action,
                                      int max_episode_steps, const float (&u_reset)[4], float& reward,
                                      int32_t& fin_len, float& fin_ret) {
    StepOut o = Env::step(s, action);
    ep_len += 1;
    ep_ret = __fadd_rn(ep_ret, o.reward);
    bool done = o.terminated || (ep_len == max_episode_steps);
    reward = o.reward;
    fin_len = 0;
    fin_ret = 0.0f;
    if (done) {
        fin_len = ep_len;
        fin_ret = ep_ret;
        Env::reset(s, u_reset);
        ep_len = 0;
        ep_ret = 0.0f;
    }
    return done;
}

// Philox reset draws for env `i` at (counter): one 4x32 block
__device__ inline void reset_uniforms(uint64_t seed, uint32_t env, uint64_t counter, float (&u)[4]) {
    uint32_t r[4];
    Philox::gen(seed, env, RNG_STREAM_RESET, (uint32_t)counter, (uint32_t)(counter >> 32), r);
#pragma unroll
    for (int k = 0; k < 4; k++) u[k] = Philox::u01(r[k]);
}

}  // namespace ppo
