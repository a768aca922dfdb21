// ppo_loss.cuh — per-sample PPO loss terms and their closed-form derivatives, shared by the fp32 FMA update
// kernel (update.cu) and the tensor-core update kernel (update_tc.cu).
// PPO/PPO_Discrete.cpp:585-631 (ratio, approx-KL, clip fraction, advantage normalisation, clipped surrogate,
// clipped value loss, entropy bonus) and what autograd produces for them (SURVEY.md §8 a-bis).
#pragma once
#include "categorical.cuh"

namespace ppo {

template <class T>
struct RowSumsT {  // per-thread partial sums of the minibatch diagnostics
    T pg = 0, v = 0, ent = 0, kl = 0, clip = 0;
};
using RowSums = RowSumsT<double>;   // fp32 FMA path: double like the oracle; the tensor-core path sums a few dozen rows
using RowSumsF = RowSumsT<float>;   // per thread in float and switches to double in the cross-thread reduction

// Actor side of one sample.  x[0..K) holds the logits on entry and dL/dlogits * (1/scale_inv) on exit:
// every gradient is multiplied by `gscale` (1/n for the FMA path, 1 for the tensor-core path, which applies
// 1/n once when the accumulators are drained).
template <class T>
__device__ inline void actor_row_loss(float* x, const Dims& d, bool masked, uint64_t mbits, const int32_t* act,
                                      float old_logp, float adv_raw, const ppo_hyper& hp, float meanf, float stdf,
                                      float gscale, RowSumsT<T>& s) {
    const int heads = d.heads;
    const float c = hp.clip_coef, lo = 1.0f - c, hi = 1.0f + c;
    // pass 1: log-prob of the stored actions + entropy, summed over heads (Agent.cpp:168)
    float newlogp = 0.0f, ent = 0.0f;
    float hent[PPO_MAX_HEADS];
    HeadStats hs[PPO_MAX_HEADS];
    int off = 0;
    for (int h = 0; h < heads; h++) {
        const int Kh = d.nvec[h];
        const uint64_t mb = mbits >> off;
        hs[h] = head_stats(x + off, Kh, mb, masked);
        const int aa = act[h];
        newlogp = __fadd_rn(newlogp, head_logp(masked_logit(x[off + aa], !masked || ((mb >> aa) & 1ull)), hs[h]));
        hent[h] = head_entropy(x + off, Kh, mb, masked, hs[h], hp.entropy_mode);
        ent = __fadd_rn(ent, hent[h]);
        off += Kh;
    }
    const float logratio = newlogp - old_logp;                      // :585
    const float ratio = expf(logratio);                             // :586
    s.kl += (T)((ratio - 1.0f) - logratio);                         // :352
    if (fabsf(ratio - 1.0f) > c) s.clip += (T)1;                    // :348
    float A = adv_raw;
    if (hp.norm_adv) A = __fdiv_rn(A - meanf, stdf + 1e-8f);        // :591-594
    const float rc = ratio < lo ? lo : (ratio > hi ? hi : ratio);
    const float pg1 = -A * ratio, pg2 = -A * rc;                    // :597-598
    s.pg += (T)(pg1 > pg2 ? pg1 : pg2);
    s.ent += (T)ent;
    // d pg / d ratio: ATen max splits ties 1/2-1/2, clamp's mask is inclusive (SURVEY §8 a-bis)
    float dpg_dr;
    if (ratio >= lo && ratio <= hi) dpg_dr = -A;
    else if (pg1 > pg2) dpg_dr = -A;
    else if (pg1 == pg2) dpg_dr = -0.5f * A;
    else dpg_dr = 0.0f;
    const float g_logp = gscale * dpg_dr * ratio;
    const bool exact = masked || hp.entropy_mode == PPO_ENTROPY_EXACT;
    const float ge = hp.ent_coef * gscale;
    off = 0;
    for (int h = 0; h < heads; h++) {
        const int Kh = d.nvec[h];
        const uint64_t mb = mbits >> off;
        const int aa = act[h];
        for (int j = 0; j < Kh; j++) {
            const bool valid = !masked || ((mb >> j) & 1ull);
            const float xm = masked_logit(x[off + j], valid);
            const float p = head_prob(xm, hs[h]);
            float dz = g_logp * ((j == aa ? 1.0f : 0.0f) - p);
            if (exact && ge != 0.0f && p > 0.0f) dz += ge * p * (head_logp(xm, hs[h]) + hent[h]);
            if (!valid) dz = 0.0f;
            x[off + j] = dz;
        }
        off += Kh;
    }
}

// Critic side of one sample: accumulates the (un-halved) value-loss term, returns dL/dvalue * gscale.
template <class T>
__device__ inline float critic_row_loss(float value, float R, float vo, const ppo_hyper& hp, float gscale,
                                        RowSumsT<T>& s) {
    const float c = hp.clip_coef;
    const float du = value - R;
    const float u = du * du;
    float dv;
    if (hp.clip_vloss) {  // :603-618
        const float dvo = value - vo;
        const float dcl = dvo < -c ? -c : (dvo > c ? c : dvo);
        const float vc = vo + dcl;
        const float w = (vc - R) * (vc - R);
        s.v += (T)(u > w ? u : w);
        const float dw = (dvo >= -c && dvo <= c) ? (vc - R) : 0.0f;
        dv = (u > w) ? du : ((w > u) ? dw : 0.5f * du + 0.5f * dw);
    } else {  // :621-625
        s.v += (T)u;
        dv = du;
    }
    return hp.vf_coef * gscale * dv;
}

// Single unmasked head (PPO_Discrete) with the fast-math intrinsics: the tensor-core path's forward values carry
// ~1e-3 relative error already, so __expf / __logf / __fdividef (2 ulp-class) change nothing measurable, and the
// whole row is one short dependent chain instead of ~10 libm-grade calls.  Same formulas as actor_row_loss.
__device__ inline void actor_row_loss_fast1(float* x, int K, int aa, float old_logp, float adv_raw, const ppo_hyper& hp,
                                            float meanf, float stdf, RowSumsF& s) {
    const float c = hp.clip_coef, lo = 1.0f - c, hi = 1.0f + c;
    float mx = x[0];
    for (int j = 1; j < K; j++) mx = fmaxf(mx, x[j]);
    float sum = 0.0f;
    for (int j = 0; j < K; j++) sum += __expf(x[j] - mx);
    const float lse = mx + __logf(sum), rs = __fdividef(1.0f, sum);
    const float newlogp = x[aa] - lse;
    const bool exact = hp.entropy_mode == PPO_ENTROPY_EXACT;
    float ent = 0.0f;
    if (exact) {
        for (int j = 0; j < K; j++) {
            const float lp = x[j] - lse, p = __expf(x[j] - mx) * rs;
            if (p > 0.0f) ent -= lp * p;
        }
    } else {
        ent = -FLT_MIN;   // Categorical.cpp:114-117: every log-prob clamps up to FLT_MIN, sum(p) = 1
    }
    const float logratio = newlogp - old_logp;
    const float ratio = __expf(logratio);
    s.kl += (ratio - 1.0f) - logratio;
    if (fabsf(ratio - 1.0f) > c) s.clip += 1.0f;
    float A = adv_raw;
    if (hp.norm_adv) A = __fdividef(A - meanf, stdf + 1e-8f);
    const float rc = ratio < lo ? lo : (ratio > hi ? hi : ratio);
    const float pg1 = -A * ratio, pg2 = -A * rc;
    s.pg += pg1 > pg2 ? pg1 : pg2;
    s.ent += ent;
    float dpg_dr;
    if (ratio >= lo && ratio <= hi) dpg_dr = -A;
    else if (pg1 > pg2) dpg_dr = -A;
    else if (pg1 == pg2) dpg_dr = -0.5f * A;
    else dpg_dr = 0.0f;
    const float g_logp = dpg_dr * ratio;
    const float ge = hp.ent_coef;
    for (int j = 0; j < K; j++) {
        const float p = __expf(x[j] - mx) * rs;
        float dz = g_logp * ((j == aa ? 1.0f : 0.0f) - p);
        if (exact && ge != 0.0f && p > 0.0f) dz += ge * p * ((x[j] - lse) + ent);
        x[j] = dz;
    }
}


// actor_row_loss for the tensor-core path (any head layout, optional mask) with the fast intrinsics: the forward values
// carry ~1e-3 relative error from the 16-bit operands already.  One exponential per logit and pass, no IEEE division, and
// no per-head arrays (runtime-indexed arrays live in local memory): pass 1 overwrites every logit with its log-prob
// (-inf marks a masked one), pass 2 turns the log-probs into dL/dlogits.  Same formulas as actor_row_loss.
// apack (heads <= 8): the row's actions as 8-bit fields, prefetched with the row; else they are read through `act`.
__device__ inline void actor_row_loss_fast_multi(float* x, const Dims& d, bool masked, uint64_t mbits, const int32_t* act,
                                                 uint64_t apack, float old_logp, float adv_raw, const ppo_hyper& hp,
                                                 float meanf, float rstd, RowSumsF& s) {
    const bool packed_act = d.heads <= 8;
    auto action_of = [&](int h) { return packed_act ? (int)((apack >> (8 * h)) & 0xFFull) : act[h]; };
    const int heads = d.heads;
    const float c = hp.clip_coef, lo = 1.0f - c, hi = 1.0f + c;
    const bool exact = masked || hp.entropy_mode == PPO_ENTROPY_EXACT;
    const uint32_t mlo = (uint32_t)mbits, mhi = (uint32_t)(mbits >> 32);
    auto valid_at = [&](int g) { return !masked || (((g < 32 ? mlo : mhi) >> (g & 31)) & 1u); };
    float newlogp = 0.0f, ent = 0.0f;
    int off = 0;
    for (int h = 0; h < heads; h++) {
        const int Kh = d.nvec[h];
        float mx = -INFINITY;
        for (int j = 0; j < Kh; j++)
            if (valid_at(off + j)) mx = fmaxf(mx, x[off + j]);
        float sum = 0.0f;
        for (int j = 0; j < Kh; j++)
            if (valid_at(off + j)) sum += __expf(x[off + j] - mx);
        const float lse = mx + __logf(sum);
        float eh = 0.0f;
        for (int j = 0; j < Kh; j++) {
            const bool v = valid_at(off + j);
            const float lp = v ? x[off + j] - lse : -INFINITY;
            x[off + j] = lp;
            if (exact && v) {
                const float p = __expf(lp);
                if (p > 0.0f) eh -= lp * p;
            }
        }
        const float lpa = x[off + action_of(h)];
        newlogp += lpa == -INFINITY ? -1e8f - lse : lpa;   // CategoricalMasked.cpp:35: a masked logit reads -1e8
        ent += exact ? eh : -FLT_MIN;                        // Categorical.cpp:114-117
        off += Kh;
    }
    const float logratio = newlogp - old_logp;
    const float ratio = __expf(logratio);
    s.kl += (ratio - 1.0f) - logratio;
    if (fabsf(ratio - 1.0f) > c) s.clip += 1.0f;
    float A = adv_raw;
    if (hp.norm_adv) A = (A - meanf) * rstd;
    const float rc = ratio < lo ? lo : (ratio > hi ? hi : ratio);
    const float pg1 = -A * ratio, pg2 = -A * rc;
    s.pg += pg1 > pg2 ? pg1 : pg2;
    s.ent += ent;
    float dpg_dr;
    if (ratio >= lo && ratio <= hi) dpg_dr = -A;
    else if (pg1 > pg2) dpg_dr = -A;
    else if (pg1 == pg2) dpg_dr = -0.5f * A;
    else dpg_dr = 0.0f;
    const float g_logp = dpg_dr * ratio;
    const float ge = hp.ent_coef;
    const bool ent_grad = exact && ge != 0.0f;
    off = 0;
    for (int h = 0; h < heads; h++) {
        const int Kh = d.nvec[h];
        const int aa = action_of(h);
        float eh = 0.0f;
        if (ent_grad)
            for (int j = 0; j < Kh; j++) {
                const float lp = x[off + j];
                if (lp != -INFINITY) {
                    const float p = __expf(lp);
                    if (p > 0.0f) eh -= lp * p;
                }
            }
        for (int j = 0; j < Kh; j++) {
            const float lp = x[off + j];
            float dz = 0.0f;
            if (lp != -INFINITY) {
                const float p = __expf(lp);
                dz = g_logp * ((j == aa ? 1.0f : 0.0f) - p);
                if (ent_grad && p > 0.0f) dz += ge * p * (lp + eh);
            }
            x[off + j] = dz;
        }
        off += Kh;
    }
}

// actor_row_loss_fast1 with the row held in registers (KMAX = compile-time bound of K, loops fully unrolled and predicated):
// one exponential per logit, reused for the probabilities of the backward pass; no shared-memory round trips inside the
// dependent chain.  x[0..K) = logits on entry, dL/dlogits on exit.  Same formulas and the same intrinsics.
template <int KMAX>
__device__ __forceinline__ void actor_row_loss_fast_reg(float (&x)[KMAX], int K, int aa, float old_logp, float adv_raw,
                                                        const ppo_hyper& hp, float meanf, float rstd, RowSumsF& s) {
    const float c = hp.clip_coef, lo = 1.0f - c, hi = 1.0f + c;
    float mx = x[0];
#pragma unroll
    for (int j = 1; j < KMAX; j++)
        if (j < K) mx = fmaxf(mx, x[j]);
    float e[KMAX], sum = 0.0f, xa = x[0];
#pragma unroll
    for (int j = 0; j < KMAX; j++) {
        e[j] = j < K ? __expf(x[j] - mx) : 0.0f;
        sum += e[j];
        if (j == aa) xa = x[j];
    }
    const float lse = mx + __logf(sum), rs = __fdividef(1.0f, sum);
    const float newlogp = xa - lse;
    const bool exact = hp.entropy_mode == PPO_ENTROPY_EXACT;
    float ent = 0.0f;
    if (exact) {
#pragma unroll
        for (int j = 0; j < KMAX; j++) {
            const float p = e[j] * rs;
            if (j < K && p > 0.0f) ent -= (x[j] - lse) * p;
        }
    } else {
        ent = -FLT_MIN;   // Categorical.cpp:114-117
    }
    const float logratio = newlogp - old_logp;
    const float ratio = __expf(logratio);
    s.kl += (ratio - 1.0f) - logratio;
    if (fabsf(ratio - 1.0f) > c) s.clip += 1.0f;
    float A = adv_raw;
    if (hp.norm_adv) A = (A - meanf) * rstd;     // rstd = 1 / (std + 1e-8), formed once per kernel with __fdividef
    const float rc = ratio < lo ? lo : (ratio > hi ? hi : ratio);
    const float pg1 = -A * ratio, pg2 = -A * rc;
    s.pg += pg1 > pg2 ? pg1 : pg2;
    s.ent += ent;
    float dpg_dr;
    if (ratio >= lo && ratio <= hi) dpg_dr = -A;
    else if (pg1 > pg2) dpg_dr = -A;
    else if (pg1 == pg2) dpg_dr = -0.5f * A;
    else dpg_dr = 0.0f;
    const float g_logp = dpg_dr * ratio;
    const float ge = hp.ent_coef;
#pragma unroll
    for (int j = 0; j < KMAX; j++) {
        const float p = e[j] * rs;
        float dz = g_logp * ((j == aa ? 1.0f : 0.0f) - p);
        if (exact && ge != 0.0f && p > 0.0f) dz += ge * p * ((x[j] - lse) + ent);
        x[j] = j < K ? dz : 0.0f;
    }
}

}  // namespace ppo
