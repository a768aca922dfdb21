/* oracle/ppo_oracle.c — TEST INFRASTRUCTURE, not product code.  See ppo_oracle.h.
 *
 * Plain fp32 C, compiled with -ffp-contract=off so that a*b+c is two roundings like the
 * reference's x86-64 build (no -mfma).  Reductions that LibTorch performs with pairwise /
 * double accumulation are accumulated in double here.
 * All citations are file:line under /root/reference.
 */
#include "ppo_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ======================= a1/a2: CartPole ======================= */

/* Environments/CartPole.cpp:6-17 (constants), :47-94 (step) */
void orc_cartpole_step(float* s, int64_t action, float* reward, int* terminated) {
    const float gravity = 9.8f, mass_cart = 1.0f, mass_pole = 0.1f;
    const float total_mass = mass_pole + mass_cart;
    const float length = 0.5f;
    const float polemass_length = mass_pole * length;
    const float force_mag = 10.0f, tau = 0.02f;
    /* :16 computed in double, narrowed to the float member */
    const float theta_thr = (float)(12 * 2 * M_PI / 360);
    const float x_thr = 2.4f;

    float x = s[0], x_dot = s[1], theta = s[2], theta_dot = s[3];
    float force = force_mag;
    if (action == 0) force = -force;
    float cos_theta = cosf(theta);
    float sin_theta = sinf(theta);
    float temp = (force + polemass_length * theta_dot * theta_dot * sin_theta) / total_mass;
    float theta_acc = (gravity * sin_theta - cos_theta * temp) /
                      (length * (4.0f / 3.0f - mass_pole * cos_theta * cos_theta / total_mass));
    float x_acc = temp - polemass_length * theta_acc * cos_theta / total_mass;
    x = x + tau * x_dot;
    x_dot = x_dot + tau * x_acc;
    theta = theta + tau * theta_dot;
    theta_dot = theta_dot + tau * theta_acc;
    s[0] = x; s[1] = x_dot; s[2] = theta; s[3] = theta_dot;
    int term = (x < -x_thr || x > x_thr || theta < -theta_thr || theta > theta_thr);
    *terminated = term;
    *reward = term ? -1.0f : 1.0f;   /* :80-88 (Q6) */
}

/* Environments/CartPole.cpp:34-45; dis(-0.05f,0.05f) => u*(b-a)+a in fp32 */
void orc_cartpole_reset(float* s, const float* u4) {
    const float a = -0.05f, b = 0.05f;
    for (int i = 0; i < 4; i++) s[i] = u4[i] * (b - a) + a;
}

/* libstdc++: std::mt19937 + generate_canonical<float,24>: one 32-bit draw per value,
 * (float)raw / 2^32 (both in float; a result of 1.0f is nudged to nextafter(1,0)). */
void orc_mt19937_canonical_f32(uint32_t seed, int n, float* u_out) {
    uint32_t mt[624];
    int idx = 624;
    mt[0] = seed;
    for (int i = 1; i < 624; i++) mt[i] = 1812433253u * (mt[i - 1] ^ (mt[i - 1] >> 30)) + (uint32_t)i;
    for (int k = 0; k < n; k++) {
        if (idx >= 624) {
            for (int i = 0; i < 624; i++) {
                uint32_t y = (mt[i] & 0x80000000u) | (mt[(i + 1) % 624] & 0x7fffffffu);
                mt[i] = mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            }
            idx = 0;
        }
        uint32_t y = mt[idx++];
        y ^= (y >> 11);
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= (y >> 18);
        float r = (float)y / 4294967296.0f;
        if (r >= 1.0f) r = nextafterf(1.0f, 0.0f);
        u_out[k] = r;
    }
}

/* ======================= a3/a4: MountainCar ======================= */

static float clampf(float v, float lo, float hi) { return v < lo ? lo : (hi < v ? hi : v); }

/* Environments/MountainCar.cpp:6-13 (constants), :29-57 (step) */
void orc_mountaincar_step(float* s, int64_t action, float* reward, int* terminated) {
    const float min_position = -1.2f, max_position = 0.6f, max_speed = 0.07f;
    const float goal_position = 0.5f, goal_velocity = 0.0f, force = 0.001f, gravity = 0.0025f;
    float position = s[0], velocity = s[1];
    velocity += ((float)action - 1.0f) * force + cosf(3.0f * position) * (-gravity);
    velocity = clampf(velocity, -max_speed, max_speed);
    position += velocity;
    position = clampf(position, min_position, max_position);
    if (position == min_position && velocity < 0.0f) velocity = 0.0f;
    *terminated = (position >= goal_position && velocity >= goal_velocity);
    *reward = -1.0f;
    s[0] = position; s[1] = velocity;
}

/* Environments/MountainCar.cpp:60-67,79-88: p ~ U(-0.6,-0.4), v = 0 */
void orc_mountaincar_reset(float* s, float u) {
    const float a = -0.6f, b = -0.4f;
    s[0] = u * (b - a) + a;
    s[1] = 0.0f;
}

/* ======================= a5: vectorised step ======================= */

/* PPO/PPO_Discrete.cpp:429-466: step, truncate at episode_length == max_episode_steps (:443),
 * on done record (len, ret) and reset (:453-458); reward is the step's reward, done -> m_dones. */
/* ---- MultiWalk: synthetic MultiDiscrete env with a state-dependent mask (no reference counterpart; restates
 * ppo-libtorch_b200/csrc/env.cuh: MultiWalkEnv operation by operation, -ffp-contract=off keeps the roundings) ---- */
void orc_multiwalk_reset(float* s, const float* u4) {
    for (int i = 0; i < 4; i++) s[i] = u4[i] * 1.6f + -0.8f;
}
void orc_multiwalk_step(float* s, const int32_t* acts, float* reward, int* terminated) {
    const float gain = 0.05f * (float)(1 + acts[4]);
    float dist = 0.0f;
    for (int i = 0; i < 4; i++) {
        float p = s[i] + (float)(acts[i] - 1) * gain;
        p = p < -1.0f ? -1.0f : (p > 1.0f ? 1.0f : p);
        s[i] = p;
        dist = dist + fabsf(p);
    }
    const float bonus = 0.01f * (float)(acts[5] + acts[6] + acts[7]);
    *terminated = dist < 0.2f;
    *reward = *terminated ? 1.0f : bonus - 0.25f * dist;
}
/* mask over the 21 logits of nvec {3,3,3,3,3,2,2,2}, one byte per logit */
void orc_multiwalk_mask(const float* s, uint8_t* mask21) {
    int small = 1;
    for (int i = 0; i < 4; i++) {
        mask21[3 * i + 0] = s[i] > -0.95f;
        mask21[3 * i + 1] = 1;
        mask21[3 * i + 2] = s[i] < 0.95f;
        small = small && fabsf(s[i]) <= 0.7f;
    }
    mask21[12] = 1; mask21[13] = 1; mask21[14] = (uint8_t)small;
    for (int h = 0; h < 3; h++) {
        mask21[15 + 2 * h] = 1;
        mask21[16 + 2 * h] = s[h] >= 0.0f;
    }
}

void orc_env_step(int kind, int n, float* state, int32_t* ep_len, float* ep_ret,
                  const int32_t* actions, int max_episode_steps, const float* reset_u,
                  float* obs_out, float* reward_out, float* done_out,
                  int32_t* fin_len, float* fin_ret) {
    const int S = (kind == ORC_ENV_MOUNTAINCAR) ? 2 : 4;
    for (int i = 0; i < n; i++) {
        float* s = state + (size_t)i * S;
        float r; int term;
        if (kind == ORC_ENV_CARTPOLE) orc_cartpole_step(s, actions[i], &r, &term);
        else if (kind == ORC_ENV_MULTIWALK) orc_multiwalk_step(s, actions + (size_t)i * 8, &r, &term);   /* actions [n, 8] */
        else orc_mountaincar_step(s, actions[i], &r, &term);
        ep_len[i] += 1;              /* CartPole.cpp:90-91 / MountainCar.cpp:51-52 */
        ep_ret[i] += r;
        if (ep_len[i] == max_episode_steps) term = 1;
        if (fin_len) fin_len[i] = 0;
        if (fin_ret) fin_ret[i] = 0.0f;
        if (term) {
            if (fin_len) fin_len[i] = ep_len[i];
            if (fin_ret) fin_ret[i] = ep_ret[i];
            if (kind == ORC_ENV_CARTPOLE) orc_cartpole_reset(s, reset_u + (size_t)i * 4);
            else if (kind == ORC_ENV_MULTIWALK) orc_multiwalk_reset(s, reset_u + (size_t)i * 4);
            else orc_mountaincar_reset(s, reset_u[(size_t)i * 4]);
            ep_len[i] = 0;
            ep_ret[i] = 0.0f;
        }
        for (int k = 0; k < S; k++) obs_out[(size_t)i * S + k] = s[k];
        reward_out[i] = r;
        done_out[i] = term ? 1.0f : 0.0f;
    }
}

/* ======================= a6: MLP ======================= */

int64_t orc_num_params(int O, int H, int K) {
    int64_t critic = (int64_t)H * O + H + (int64_t)H * H + H + H + 1;
    int64_t actor = (int64_t)H * O + H + (int64_t)H * H + H + (int64_t)K * H + K;
    return critic + actor;
}

typedef struct {
    const float *W1, *b1, *W2, *b2, *W3, *b3;
} net_t;

static void split_params(const float* p, int O, int H, int K, net_t* critic, net_t* actor) {
    critic->W1 = p; p += (size_t)H * O;
    critic->b1 = p; p += H;
    critic->W2 = p; p += (size_t)H * H;
    critic->b2 = p; p += H;
    critic->W3 = p; p += H;
    critic->b3 = p; p += 1;
    actor->W1 = p; p += (size_t)H * O;
    actor->b1 = p; p += H;
    actor->W2 = p; p += (size_t)H * H;
    actor->b2 = p; p += H;
    actor->W3 = p; p += (size_t)K * H;
    actor->b3 = p;
}

/* Linear->Tanh->Linear->Tanh->Linear, y = x W^T + b  (PPO/Agent.cpp:25-59) */
static void net_forward(const net_t* nt, int O, int H, int Kout, const float* x,
                        float* h1, float* h2, float* out) {
    for (int j = 0; j < H; j++) {
        float acc = nt->b1[j];
        for (int i = 0; i < O; i++) acc += x[i] * nt->W1[(size_t)j * O + i];
        h1[j] = tanhf(acc);
    }
    for (int j = 0; j < H; j++) {
        float acc = nt->b2[j];
        for (int k = 0; k < H; k++) acc += h1[k] * nt->W2[(size_t)j * H + k];
        h2[j] = tanhf(acc);
    }
    for (int j = 0; j < Kout; j++) {
        float acc = nt->b3[j];
        for (int k = 0; k < H; k++) acc += h2[k] * nt->W3[(size_t)j * H + k];
        out[j] = acc;
    }
}

void orc_mlp_forward(const float* params, int O, int H, int K, int n, const float* obs,
                     float* logits, float* value) {
    net_t critic, actor;
    split_params(params, O, H, K, &critic, &actor);
    float* h1 = (float*)malloc(sizeof(float) * H);
    float* h2 = (float*)malloc(sizeof(float) * H);
    for (int s = 0; s < n; s++) {
        const float* x = obs + (size_t)s * O;
        if (logits) net_forward(&actor, O, H, K, x, h1, h2, logits + (size_t)s * K);
        if (value) net_forward(&critic, O, H, 1, x, h1, h2, value + s);
    }
    free(h1); free(h2);
}

/* ======================= a7/a8: Categorical(Masked) ======================= */

/* One head of K logits.  Categorical.cpp:33-35 / CategoricalMasked.cpp:34-41:
 *   x' = where(mask, x, -1e8f); m_logits = x' - logsumexp(x'); m_probs = softmax(x'). */
static void head_dist(int K, const float* x, const uint8_t* mask, float* xm, float* nl, float* p) {
    float mx = -INFINITY;
    for (int j = 0; j < K; j++) {
        xm[j] = (mask == NULL || mask[j]) ? x[j] : -1e8f;
        if (xm[j] > mx) mx = xm[j];
    }
    float sum = 0.0f;
    for (int j = 0; j < K; j++) { p[j] = expf(xm[j] - mx); sum += p[j]; }
    float lse = mx + logf(sum);
    for (int j = 0; j < K; j++) { nl[j] = xm[j] - lse; p[j] = p[j] / sum; }
}

static float head_entropy(int K, const float* nl, const float* p, const uint8_t* mask, int mode) {
    float acc = 0.0f;
    if (mask == NULL && mode == ORC_ENTROPY_REFERENCE) {
        /* Categorical.cpp:114-117 (Q1): clamp(m_logits, min=FLT_MIN) * m_probs */
        for (int j = 0; j < K; j++) {
            float l = nl[j] < FLT_MIN ? FLT_MIN : nl[j];
            acc += l * p[j];
        }
    } else {
        /* CategoricalMasked.cpp:140-142; unmasked exact = same without the where */
        for (int j = 0; j < K; j++) {
            float t = nl[j] * p[j];
            if (mask != NULL && !mask[j]) t = 0.0f;
            if (mask == NULL && p[j] == 0.0f) t = 0.0f;   /* exact mode: 0*log0 := 0 */
            acc += t;
        }
    }
    return -acc;
}

void orc_categorical(int n, int heads, const int32_t* nvec, const float* logits,
                     const uint8_t* mask, const int32_t* actions_in, const float* uniforms,
                     int entropy_mode, int32_t* actions_out, float* logprob, float* entropy,
                     float* probs_out, float* normlogits_out) {
    int sumK = 0, maxK = 0;
    for (int h = 0; h < heads; h++) { sumK += nvec[h]; if (nvec[h] > maxK) maxK = nvec[h]; }
    float* xm = (float*)malloc(sizeof(float) * maxK);
    float* nl = (float*)malloc(sizeof(float) * maxK);
    float* p = (float*)malloc(sizeof(float) * maxK);
    for (int r = 0; r < n; r++) {
        float lp_sum = 0.0f, ent_sum = 0.0f;
        int off = 0;
        for (int h = 0; h < heads; h++) {
            int K = nvec[h];
            const float* x = logits + (size_t)r * sumK + off;
            const uint8_t* mk = mask ? mask + (size_t)r * sumK + off : NULL;
            head_dist(K, x, mk, xm, nl, p);
            int a;
            if (actions_in) {
                a = actions_in[(size_t)r * heads + h];
            } else {
                /* inverse-CDF on an injected uniform */
                float u = uniforms[(size_t)r * heads + h];
                float c = 0.0f;
                a = -1;
                int last_valid = 0;
                for (int j = 0; j < K; j++) {
                    if (p[j] > 0.0f) last_valid = j;
                    c += p[j];
                    if (a < 0 && u < c) a = j;
                }
                if (a < 0) a = last_valid;
            }
            if (actions_out) actions_out[(size_t)r * heads + h] = a;
            lp_sum += nl[a];                                   /* Categorical.cpp:92-101 gather */
            ent_sum += head_entropy(K, nl, p, mk, entropy_mode);
            if (probs_out) memcpy(probs_out + (size_t)r * sumK + off, p, sizeof(float) * K);
            if (normlogits_out) memcpy(normlogits_out + (size_t)r * sumK + off, nl, sizeof(float) * K);
            off += K;
        }
        if (logprob) logprob[r] = lp_sum;                      /* Agent.cpp:168 sum over heads */
        if (entropy) entropy[r] = ent_sum;
    }
    free(xm); free(nl); free(p);
}

/* ======================= a9: GAE ======================= */

/* PPO/PPO_Discrete.cpp:289-305 (GAE) and :309-329 (plain returns).  Time-major [T,N]. */
void orc_gae(int T, int N, const float* rewards, const float* values, const float* dones,
             const float* next_value, const float* next_done, float gamma, float lam,
             int use_gae, float* adv_out, float* ret_out) {
    const float gl = gamma * lam;   /* m_gamma * m_gae_lambda: float*float */
    for (int n = 0; n < N; n++) {
        float last = 0.0f;          /* lastgaelam, or next_return in the non-GAE branch */
        for (int t = T - 1; t >= 0; t--) {
            size_t i = (size_t)t * N + n;
            float nnt, nv;
            if (t == T - 1) { nnt = 1.0f - next_done[n]; nv = next_value[n]; }
            else { nnt = 1.0f - dones[i + N]; nv = values[i + N]; }
            if (use_gae) {
                float delta = rewards[i] + gamma * nv * nnt - values[i];
                last = delta + gl * nnt * last;
                adv_out[i] = last;
                ret_out[i] = last + values[i];
            } else {
                float nr = (t == T - 1) ? nv : last;
                last = rewards[i] + gamma * nnt * nr;
                ret_out[i] = last;
                adv_out[i] = last - values[i];
            }
        }
    }
}

/* ======================= a10/a11: minibatch loss + backward ======================= */

typedef struct {
    float *W1, *b1, *W2, *b2, *W3, *b3;
} gnet_t;

static void split_grads(float* p, int O, int H, int K, gnet_t* critic, gnet_t* actor) {
    critic->W1 = p; p += (size_t)H * O;
    critic->b1 = p; p += H;
    critic->W2 = p; p += (size_t)H * H;
    critic->b2 = p; p += H;
    critic->W3 = p; p += H;
    critic->b3 = p; p += 1;
    actor->W1 = p; p += (size_t)H * O;
    actor->b1 = p; p += H;
    actor->W2 = p; p += (size_t)H * H;
    actor->b2 = p; p += H;
    actor->W3 = p; p += (size_t)K * H;
    actor->b3 = p;
}

/* Backward of one net for one sample, accumulating into double grads. */
static void net_backward(const net_t* nt, int O, int H, int Kout, const float* x,
                         const float* h1, const float* h2, const float* dout,
                         double* gW1, double* gb1, double* gW2, double* gb2, double* gW3, double* gb3,
                         float* dh2, float* dh1) {
    for (int k = 0; k < H; k++) dh2[k] = 0.0f;
    for (int j = 0; j < Kout; j++) {
        gb3[j] += dout[j];
        for (int k = 0; k < H; k++) {
            gW3[(size_t)j * H + k] += (double)dout[j] * h2[k];
            dh2[k] += dout[j] * nt->W3[(size_t)j * H + k];
        }
    }
    for (int k = 0; k < H; k++) dh1[k] = 0.0f;
    for (int j = 0; j < H; j++) {
        float dz = dh2[j] * (1.0f - h2[j] * h2[j]);
        gb2[j] += dz;
        for (int k = 0; k < H; k++) {
            gW2[(size_t)j * H + k] += (double)dz * h1[k];
            dh1[k] += dz * nt->W2[(size_t)j * H + k];
        }
    }
    for (int j = 0; j < H; j++) {
        float dz = dh1[j] * (1.0f - h1[j] * h1[j]);
        gb1[j] += dz;
        for (int i = 0; i < O; i++) gW1[(size_t)j * O + i] += (double)dz * x[i];
    }
}

void orc_minibatch_grads(const float* params, int O, int H, int heads, const int32_t* nvec,
                         int n, const float* obs, const uint8_t* mask, const int32_t* actions,
                         const float* old_logprob, const float* adv, const float* ret,
                         const float* old_val, const orc_hyper* hp, float* grads, float* diag) {
    int K = 0, maxK = 0;
    for (int h = 0; h < heads; h++) { K += nvec[h]; if (nvec[h] > maxK) maxK = nvec[h]; }
    const int64_t P = orc_num_params(O, H, K);
    net_t critic, actor;
    split_params(params, O, H, K, &critic, &actor);
    double* g = (double*)calloc((size_t)P, sizeof(double));
    /* double mirror of the flat layout */
    double *cW1 = g, *cb1 = cW1 + (size_t)H * O, *cW2 = cb1 + H, *cb2 = cW2 + (size_t)H * H,
           *cW3 = cb2 + H, *cb3 = cW3 + H;
    double *aW1 = cb3 + 1, *ab1 = aW1 + (size_t)H * O, *aW2 = ab1 + H, *ab2 = aW2 + (size_t)H * H,
           *aW3 = ab2 + H, *ab3 = aW3 + (size_t)K * H;

    /* :591-594 advantage normalisation: mean, unbiased std (LibTorch accumulates in double) */
    double mean = 0.0, var = 0.0;
    if (hp->norm_adv) {
        for (int i = 0; i < n; i++) mean += adv[i];
        mean /= n;
        for (int i = 0; i < n; i++) { double d = adv[i] - mean; var += d * d; }
        var /= (n - 1);
    }
    const float meanf = (float)mean, stdf = (float)sqrt(var);
    const float c = hp->clip_coef;
    const float lo = 1.0f - c, hi = 1.0f + c;

    float* ah1 = (float*)malloc(sizeof(float) * H); float* ah2 = (float*)malloc(sizeof(float) * H);
    float* ch1 = (float*)malloc(sizeof(float) * H); float* ch2 = (float*)malloc(sizeof(float) * H);
    float* dh1 = (float*)malloc(sizeof(float) * H); float* dh2 = (float*)malloc(sizeof(float) * H);
    float* logits = (float*)malloc(sizeof(float) * K);
    float* dlogits = (float*)malloc(sizeof(float) * K);
    float* xm = (float*)malloc(sizeof(float) * maxK);
    float* nl = (float*)malloc(sizeof(float) * K);
    float* p = (float*)malloc(sizeof(float) * K);
    float* hent = (float*)malloc(sizeof(float) * heads);

    double s_pg = 0, s_v = 0, s_ent = 0, s_kl = 0, s_clip = 0;
    const float inv_n = 1.0f / (float)n;
    for (int s = 0; s < n; s++) {
        const float* x = obs + (size_t)s * O;
        float value;
        net_forward(&actor, O, H, K, x, ah1, ah2, logits);
        net_forward(&critic, O, H, 1, x, ch1, ch2, &value);
        /* distributions */
        float newlogp = 0.0f, ent = 0.0f;
        int off = 0;
        for (int h = 0; h < heads; h++) {
            int Kh = nvec[h];
            const uint8_t* mk = mask ? mask + (size_t)s * K + off : NULL;
            head_dist(Kh, logits + off, mk, xm, nl + off, p + off);
            newlogp += nl[off + actions[(size_t)s * heads + h]];
            hent[h] = head_entropy(Kh, nl + off, p + off, mk, hp->entropy_mode);
            ent += hent[h];
            off += Kh;
        }
        float logratio = newlogp - old_logprob[s];   /* :585 */
        float ratio = expf(logratio);                /* :586 */
        s_kl += (double)((ratio - 1.0f) - logratio); /* :352 */
        if (fabsf(ratio - 1.0f) > c) s_clip += 1.0;  /* :348 */
        float A = adv[s];
        if (hp->norm_adv) A = (A - meanf) / (stdf + 1e-8f);
        float rc = ratio < lo ? lo : (ratio > hi ? hi : ratio);
        float pg1 = -A * ratio, pg2 = -A * rc;       /* :597-598 */
        s_pg += (double)(pg1 > pg2 ? pg1 : pg2);
        /* d pg / d ratio (ATen max: ties split 1/2-1/2; clamp mask inclusive) */
        float dpg_dr;
        int inside = (ratio >= lo && ratio <= hi);
        if (inside) dpg_dr = -A;
        else if (pg1 > pg2) dpg_dr = -A;
        else if (pg1 == pg2) dpg_dr = -0.5f * A;
        else dpg_dr = 0.0f;
        float g_logp = inv_n * dpg_dr * ratio;
        /* value loss :603-625 */
        float R = ret[s], dv;
        float u = (value - R) * (value - R);
        if (hp->clip_vloss) {
            float dvo = value - old_val[s];
            float dcl = dvo < -c ? -c : (dvo > c ? c : dvo);
            float vc = old_val[s] + dcl;
            float w = (vc - R) * (vc - R);
            s_v += (double)(u > w ? u : w);
            int in_v = (dvo >= -c && dvo <= c);
            float du = (value - R), dw = in_v ? (vc - R) : 0.0f;
            if (u > w) dv = du; else if (w > u) dv = dw; else dv = 0.5f * du + 0.5f * dw;
        } else {
            s_v += (double)u;
            dv = (value - R);
        }
        float g_v = hp->vf_coef * inv_n * dv;
        s_ent += (double)ent;
        /* dL/dlogits */
        off = 0;
        for (int h = 0; h < heads; h++) {
            int Kh = nvec[h];
            int a = actions[(size_t)s * heads + h];
            int exact = (mask != NULL) || (hp->entropy_mode == ORC_ENTROPY_EXACT);
            for (int j = 0; j < Kh; j++) {
                float d = g_logp * ((j == a ? 1.0f : 0.0f) - p[off + j]);
                if (exact && hp->ent_coef != 0.0f && p[off + j] > 0.0f)
                    d += hp->ent_coef * inv_n * p[off + j] * (nl[off + j] + hent[h]);
                if (mask && !mask[(size_t)s * K + off + j]) d = 0.0f;
                dlogits[off + j] = d;
            }
            off += Kh;
        }
        net_backward(&actor, O, H, K, x, ah1, ah2, dlogits, aW1, ab1, aW2, ab2, aW3, ab3, dh2, dh1);
        net_backward(&critic, O, H, 1, x, ch1, ch2, &g_v, cW1, cb1, cW2, cb2, cW3, cb3, dh2, dh1);
    }
    for (int64_t i = 0; i < P; i++) grads[i] = (float)g[i];
    float pg_loss = (float)(s_pg / n);
    float v_loss = 0.5f * (float)(s_v / n);
    float ent_mean = (float)(s_ent / n);
    diag[0] = pg_loss;
    diag[1] = v_loss;
    diag[2] = ent_mean;
    diag[3] = pg_loss - hp->ent_coef * ent_mean + v_loss * hp->vf_coef;   /* :631 */
    diag[4] = (float)(s_kl / n);
    diag[5] = (float)(s_clip / n);
    diag[6] = meanf;
    diag[7] = stdf;
    free(g); free(ah1); free(ah2); free(ch1); free(ch2); free(dh1); free(dh2);
    free(logits); free(dlogits); free(xm); free(nl); free(p); free(hent);
}

/* ======================= a12: clip_grad_norm_ + AdamW ======================= */

/* torch/nn/utils/clip_grad.h: coef = max_norm/(norm+1e-6), applied iff < 1.
 * torch/csrc/api/src/optim/adamw.cpp: p *= 1-lr*wd; m,v EMA; p -= lr/bc1 * m/(sqrt(v)/sqrt(bc2)+eps).
 * Scalars are doubles narrowed to float at the tensor op, as ATen does. */
float orc_clip_adamw(float* params, float* grads, float* m, float* v, int64_t n_params,
                     float lr, float beta1, float beta2, float eps, float weight_decay,
                     int64_t step, float max_grad_norm) {
    double ss = 0.0;
    for (int64_t i = 0; i < n_params; i++) ss += (double)grads[i] * grads[i];
    double norm = sqrt(ss);
    double coef = (double)max_grad_norm / (norm + 1e-6);
    if (coef < 1.0) {
        float cf = (float)coef;
        for (int64_t i = 0; i < n_params; i++) grads[i] *= cf;
    }
    const double b1 = beta1, b2 = beta2;
    const double bc1 = 1.0 - pow(b1, (double)step);
    const double bc2 = 1.0 - pow(b2, (double)step);
    const float decay = (float)(1.0 - (double)lr * (double)weight_decay);
    const float omb1 = (float)(1.0 - b1), omb2 = (float)(1.0 - b2);
    const float sq_bc2 = (float)sqrt(bc2);
    const float neg_step_size = (float)(-((double)lr / bc1));
    for (int64_t i = 0; i < n_params; i++) {
        float p = params[i] * decay;
        float g = grads[i];
        m[i] = m[i] * beta1 + g * omb1;
        v[i] = v[i] * beta2 + (g * g) * omb2;
        float denom = sqrtf(v[i]) / sq_bc2 + eps;
        params[i] = p + neg_step_size * (m[i] / denom);
    }
    return (float)norm;
}

/* PPO/PPO_Discrete.cpp:647-648: 1 - var(R - V)/var(R), unbiased */
float orc_explained_variance(int64_t n, const float* returns, const float* values) {
    double mr = 0, md = 0;
    for (int64_t i = 0; i < n; i++) { mr += returns[i]; md += (double)(returns[i] - values[i]); }
    mr /= n; md /= n;
    double vr = 0, vd = 0;
    for (int64_t i = 0; i < n; i++) {
        double a = returns[i] - mr, b = (double)(returns[i] - values[i]) - md;
        vr += a * a; vd += b * b;
    }
    vr /= (n - 1); vd /= (n - 1);
    return (float)(1.0 - vd / vr);
}
