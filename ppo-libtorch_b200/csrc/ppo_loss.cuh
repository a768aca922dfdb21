// ppo_loss.cuh — per-sample PPO loss terms and their closed-form derivatives, shared by the fp32 FMA update
// kernel (update.cu) and the tensor-core update kernel (update_tc.cu).
// PPO/PPO_Discrete.cpp:585-631 (ratio, approx-KL, clip fraction, advantage normalisation, clipped surrogate,
// clipped value loss, entropy bonus) and what autograd produces for them (SURVEY.md §8 a-bis).
#pragma once
#include "categorical.cuh"

namespace ppo {

struct RowSums {  // per-thread partial sums of the minibatch diagnostics
    double pg = 0.0, v = 0.0, ent = 0.0, kl = 0.0, clip = 0.0;
};

// Actor side of one sample.  x[0..K) holds the logits on entry and dL/dlogits * (1/scale_inv) on exit:
// every gradient is multiplied by `gscale` (1/n for the FMA path, 1 for the tensor-core path, which applies
// 1/n once when the accumulators are drained).
__device__ inline void actor_row_loss(float* x, const Dims& d, bool masked, uint64_t mbits, const int32_t* act,
                                      float old_logp, float adv_raw, const ppo_hyper& hp, float meanf, float stdf,
                                      float gscale, RowSums& s) {
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
    s.kl += (double)((ratio - 1.0f) - logratio);                    // :352
    if (fabsf(ratio - 1.0f) > c) s.clip += 1.0;                     // :348
    float A = adv_raw;
    if (hp.norm_adv) A = __fdiv_rn(A - meanf, stdf + 1e-8f);        // :591-594
    const float rc = ratio < lo ? lo : (ratio > hi ? hi : ratio);
    const float pg1 = -A * ratio, pg2 = -A * rc;                    // :597-598
    s.pg += (double)(pg1 > pg2 ? pg1 : pg2);
    s.ent += (double)ent;
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
__device__ inline float critic_row_loss(float value, float R, float vo, const ppo_hyper& hp, float gscale,
                                        RowSums& s) {
    const float c = hp.clip_coef;
    const float du = value - R;
    const float u = du * du;
    float dv;
    if (hp.clip_vloss) {  // :603-618
        const float dvo = value - vo;
        const float dcl = dvo < -c ? -c : (dvo > c ? c : dvo);
        const float vc = vo + dcl;
        const float w = (vc - R) * (vc - R);
        s.v += (double)(u > w ? u : w);
        const float dw = (dvo >= -c && dvo <= c) ? (vc - R) : 0.0f;
        dv = (u > w) ? du : ((w > u) ? dw : 0.5f * du + 0.5f * dw);
    } else {  // :621-625
        s.v += (double)u;
        dv = du;
    }
    return hp.vf_coef * gscale * dv;
}

}  // namespace ppo
