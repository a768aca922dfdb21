// categorical.cuh — per-head Categorical / CategoricalMasked arithmetic shared by the standalone
// K2 kernel, the fused policy forward, the fused rollout and the fused update.
//   Distributions/Categorical.cpp:28-39 (normalise), :73-79 (sample), :92-101 (log_prob), :112-119 (entropy)
//   Distributions/CategoricalMasked.cpp:31-46, :87-94, :107-116, :127-144
// A head's K logits are read through a pointer (shared or global memory, unit stride); the
// validity mask of the whole row travels as a 64-bit word (sumK <= PPO_MAX_LOGITS = 64).
#pragma once
#include <float.h>

#include "common.cuh"

namespace ppo {

struct HeadStats {
    float mx, sum, lse;
};

__device__ inline float masked_logit(float x, bool valid) { return valid ? x : -1e8f; }  // CategoricalMasked.cpp:35

// m_logits = x' - logsumexp(x'), m_probs = softmax(x')  (Categorical.cpp:33-35)
__device__ inline HeadStats head_stats(const float* x, int K, uint64_t mbits, bool masked) {
    HeadStats s;
    s.mx = -INFINITY;
    const uint32_t lo = (uint32_t)mbits, hi = (uint32_t)(mbits >> 32);   // 32-bit tests: no 64-bit variable shifts
    for (int j = 0; j < K; j++)
        s.mx = fmaxf(s.mx, masked_logit(x[j], !masked || (((j < 32 ? lo : hi) >> (j & 31)) & 1u)));
    s.sum = 0.0f;
    for (int j = 0; j < K; j++)
        s.sum = __fadd_rn(s.sum, expf(__fsub_rn(masked_logit(x[j], !masked || (((j < 32 ? lo : hi) >> (j & 31)) & 1u)),
                                                s.mx)));
    s.lse = __fadd_rn(s.mx, logf(s.sum));
    return s;
}
__device__ inline float head_prob(float xm, const HeadStats& s) { return __fdiv_rn(expf(__fsub_rn(xm, s.mx)), s.sum); }
__device__ inline float head_logp(float xm, const HeadStats& s) { return __fsub_rn(xm, s.lse); }

// Inverse-CDF sample on a canonical uniform: first j with u < cumsum(p)_j, else the last j with
// p_j > 0 (the reference's multinomial stream is not reproducible, parity is
// defined on injected uniforms).
__device__ inline int head_sample(const float* x, int K, uint64_t mbits, bool masked, const HeadStats& s, float u) {
    float c = 0.0f;
    int a = -1, last_valid = 0;
    for (int j = 0; j < K; j++) {
        const float p = head_prob(masked_logit(x[j], !masked || ((mbits >> j) & 1ull)), s);
        if (p > 0.0f) last_valid = j;
        c = __fadd_rn(c, p);
        if (a < 0 && u < c) a = j;
    }
    return a < 0 ? last_valid : a;
}

// entropy of one head.  Unmasked + REFERENCE mode reproduces Categorical.cpp:114-117 (clamp with
// min = FLT_MIN); otherwise -sum(where(valid, logp*p, 0)) (CategoricalMasked.cpp:140-142).
__device__ inline float head_entropy(const float* x, int K, uint64_t mbits, bool masked, const HeadStats& s,
                                     int mode) {
    float acc = 0.0f;
    if (!masked && mode == PPO_ENTROPY_REFERENCE) {
        for (int j = 0; j < K; j++) {
            const float l = fmaxf(head_logp(x[j], s), FLT_MIN);
            acc = __fadd_rn(acc, __fmul_rn(l, head_prob(x[j], s)));
        }
    } else {
        for (int j = 0; j < K; j++) {
            const bool valid = !masked || ((mbits >> j) & 1ull);
            const float xm = masked_logit(x[j], valid);
            const float p = head_prob(xm, s);
            float t = __fmul_rn(head_logp(xm, s), p);
            if (!valid || p == 0.0f) t = 0.0f;
            acc = __fadd_rn(acc, t);
        }
    }
    return -acc;
}

// uniform for (row, head) of sampling call `counter`
__device__ inline float sample_uniform(uint64_t seed, uint32_t row, uint64_t counter, int head) {
    uint32_t r[4];
    Philox::gen(seed, row, RNG_STREAM_SAMPLE | ((uint32_t)(head >> 2) << 8), (uint32_t)counter,
                (uint32_t)(counter >> 32), r);
    return Philox::u01(r[head & 3]);
}

// Whole-row evaluation over `heads` heads (PPO/Agent.cpp:137-170: logprob and entropy summed over
// heads).  actions_in == nullptr => sample (u from `uniforms` when given, else Philox).
struct RowOut {
    float logp, entropy;
};
// One pass over a head after head_stats: probabilities once per logit, feeding the inverse-CDF sample (or the
// given action), its log-prob and the entropy together.  Same operations in the same order per quantity as
// head_sample / head_logp / head_entropy above, so results are bit-identical to the separate passes.
__device__ inline void head_eval(const float* x, int K, uint64_t mb, bool masked, const HeadStats& s, int mode,
                                 bool sample, float u, int& a, float& logp, float& ent) {
    const bool ref_quirk = !masked && mode == PPO_ENTROPY_REFERENCE;
    float c = 0.0f, acc = 0.0f, lp_a = 0.0f;
    int pick = sample ? -1 : a, last_valid = 0;
    float lp_last = 0.0f;
    const uint32_t lo = (uint32_t)mb, hi = (uint32_t)(mb >> 32);
    for (int j = 0; j < K; j++) {
        const bool valid = !masked || (((j < 32 ? lo : hi) >> (j & 31)) & 1u);
        const float xm = masked_logit(x[j], valid);
        const float p = head_prob(xm, s);
        const float lp = head_logp(xm, s);
        float t;
        if (ref_quirk) {
            t = __fmul_rn(fmaxf(lp, FLT_MIN), p);
        } else {
            t = __fmul_rn(lp, p);
            if (!valid || p == 0.0f) t = 0.0f;
        }
        acc = __fadd_rn(acc, t);
        if (sample) {
            if (p > 0.0f) { last_valid = j; lp_last = lp; }
            c = __fadd_rn(c, p);
            if (pick < 0 && u < c) { pick = j; lp_a = lp; }
        } else if (j == pick) {
            lp_a = lp;
        }
    }
    if (sample && pick < 0) { pick = last_valid; lp_a = lp_last; }
    a = pick;
    logp = lp_a;
    ent = -acc;
}
__device__ inline RowOut row_eval(const float* x, int heads, const int* nvec, uint64_t mbits, bool masked,
                                  const int32_t* actions_in, const float* uniforms, uint64_t seed, uint32_t row,
                                  uint64_t counter, int entropy_mode, int32_t* actions_out) {
    RowOut o;
    o.logp = 0.0f;
    o.entropy = 0.0f;
    int off = 0;
    uint32_t r[4] = {0, 0, 0, 0};
    const bool sample = actions_in == nullptr;
    for (int h = 0; h < heads; h++) {
        const int K = nvec[h];
        const uint64_t mb = mbits >> off;
        const HeadStats s = head_stats(x + off, K, mb, masked);
        int a = 0;
        float u = 0.0f;
        if (!sample) {
            a = actions_in[h];
        } else if (uniforms) {
            u = uniforms[h];
        } else {   // one Philox block serves four heads (same words as sample_uniform(seed,row,counter,h))
            if ((h & 3) == 0)
                Philox::gen(seed, row, RNG_STREAM_SAMPLE | ((uint32_t)(h >> 2) << 8), (uint32_t)counter,
                            (uint32_t)(counter >> 32), r);
            u = Philox::u01(r[h & 3]);
        }
        float lp, en;
        head_eval(x + off, K, mb, masked, s, entropy_mode, sample, u, a, lp, en);
        if (actions_out) actions_out[h] = a;
        o.logp = __fadd_rn(o.logp, lp);
        o.entropy = __fadd_rn(o.entropy, en);
        off += K;
    }
    return o;
}

// pack a row of the u8 mask [sumK] into bits
__device__ inline uint64_t pack_mask(const uint8_t* m, int K) {
    uint64_t b = 0;
    for (int j = 0; j < K; j++) b |= (uint64_t)(m[j] != 0) << j;
    return b;
}

}  // namespace ppo
