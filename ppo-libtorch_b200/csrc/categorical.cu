// categorical.cu — K2 standalone: Categorical / CategoricalMasked sample + log_prob + entropy over
// [n, sumK] logits split into `heads` heads.  HBM-bound (K*4 + K*1 B read, heads*4 + 8 B written per
// row): rows are staged through shared memory with fully coalesced global traffic, then one thread
// evaluates one row.  Replaces Distributions/Categorical.cpp / CategoricalMasked.cpp as used by
// Agent::getActionAndValue{Discrete,Masked} (PPO/Agent.cpp:117-170).
#include "categorical.cuh"

namespace ppo {

constexpr int CAT_ROWS = 128;

struct CatArgs {
    int n, heads, K;
    int nvec[PPO_MAX_HEADS];
    const float* logits;
    const uint8_t* mask;
    const int32_t* actions_in;
    const float* uniforms;
    uint64_t seed, counter;
    int entropy_mode;
    int32_t* actions_out;
    float* logprob;
    float* entropy;
    float* probs_out;
};

__global__ void __launch_bounds__(CAT_ROWS) categorical_kernel(const CatArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int K = a.K, ld = K | 1;  // odd stride: conflict-free row-per-thread access
    float* xs = reinterpret_cast<float*>(smem_raw);                       // [CAT_ROWS][ld]
    int32_t* acts = reinterpret_cast<int32_t*>(xs + CAT_ROWS * ld);       // [CAT_ROWS][heads]
    float* us = reinterpret_cast<float*>(acts + CAT_ROWS * a.heads);      // [CAT_ROWS][heads]
    uint8_t* ms = reinterpret_cast<uint8_t*>(us + CAT_ROWS * a.heads);    // [CAT_ROWS][K]

    const int64_t r0 = (int64_t)blockIdx.x * CAT_ROWS;
    const int rows = (int)min((int64_t)CAT_ROWS, (int64_t)a.n - r0);
    const int t = threadIdx.x;

    // coalesced staging of the tile (contiguous in global memory)
    {
        const float* src = a.logits + r0 * K;
        const int total = rows * K;
        // (row, col) of element e tracked incrementally: no per-element division
        const int qs = CAT_ROWS / K, rs = CAT_ROWS % K;
        int row = t / K, col = t % K;
        for (int e = t; e < total; e += CAT_ROWS) {
            xs[row * ld + col] = __ldg(src + e);
            row += qs; col += rs;
            if (col >= K) { col -= K; row++; }
        }
        if (a.mask) {
            const uint8_t* msrc = a.mask + r0 * K;   // tile base is 128*K bytes in: word-aligned when the buffer is
            if (rows == CAT_ROWS && (reinterpret_cast<uintptr_t>(msrc) & 3) == 0) {
                const uint32_t* m4 = reinterpret_cast<const uint32_t*>(msrc);
                uint32_t* d4 = reinterpret_cast<uint32_t*>(ms);
                for (int e = t; e < total / 4; e += CAT_ROWS) d4[e] = __ldg(m4 + e);
            } else {
                for (int e = t; e < total; e += CAT_ROWS) ms[e] = __ldg(msrc + e);
            }
        }
        const int th = rows * a.heads;
        if (a.actions_in)
            for (int e = t; e < th; e += CAT_ROWS) acts[e] = __ldg(a.actions_in + r0 * a.heads + e);
        else if (a.uniforms)
            for (int e = t; e < th; e += CAT_ROWS) us[e] = __ldg(a.uniforms + r0 * a.heads + e);
    }
    __syncthreads();

    if (t < rows) {
        float* x = xs + t * ld;
        const bool masked = a.mask != nullptr;
        const uint64_t mbits = masked ? pack_mask(ms + t * K, K) : ~0ull;
        const RowOut o = row_eval(x, a.heads, a.nvec, mbits, masked, a.actions_in ? acts + t * a.heads : nullptr,
                                  a.uniforms ? us + t * a.heads : nullptr, a.seed, (uint32_t)(r0 + t), a.counter,
                                  a.entropy_mode, acts + t * a.heads);
        if (a.logprob) a.logprob[r0 + t] = o.logp;
        if (a.entropy) a.entropy[r0 + t] = o.entropy;
        if (a.probs_out) {  // overwrite the staged logits of this row with probabilities
            int off = 0;
            for (int h = 0; h < a.heads; h++) {
                const int Kh = a.nvec[h];
                const uint64_t mb = mbits >> off;
                const HeadStats s = head_stats(x + off, Kh, mb, masked);
                float p[PPO_MAX_LOGITS];
                for (int j = 0; j < Kh; j++)
                    p[j] = head_prob(masked_logit(x[off + j], !masked || ((mb >> j) & 1ull)), s);
                for (int j = 0; j < Kh; j++) x[off + j] = p[j];
                off += Kh;
            }
        }
    }
    __syncthreads();
    if (a.actions_out) {
        const int th = rows * a.heads;
        for (int e = t; e < th; e += CAT_ROWS) a.actions_out[r0 * a.heads + e] = acts[e];
    }
    if (a.probs_out) {
        const int total = rows * K;
        const int qs = CAT_ROWS / K, rs = CAT_ROWS % K;
        int row = t / K, col = t % K;
        for (int e = t; e < total; e += CAT_ROWS) {
            a.probs_out[r0 * K + e] = xs[row * ld + col];
            row += qs; col += rs;
            if (col >= K) { col -= K; row++; }
        }
    }
}

}  // namespace ppo

using namespace ppo;

extern "C" int ppo_categorical(int n, int heads, const int32_t* h_nvec, const float* logits, const uint8_t* mask,
                               const int32_t* actions_in, const float* uniforms, uint64_t seed, uint64_t counter,
                               int entropy_mode, int32_t* actions_out, float* logprob_out, float* entropy_out,
                               float* probs_out, ppo_stream_t stream) {
    PPO_REQUIRE(n >= 0 && logits && h_nvec, "ppo_categorical: null logits/nvec or negative n");
    PPO_REQUIRE(heads >= 1 && heads <= PPO_MAX_HEADS, "ppo_categorical: heads=%d out of [1,%d]", heads,
                PPO_MAX_HEADS);
    CatArgs a{};
    a.n = n; a.heads = heads; a.K = 0;
    for (int h = 0; h < heads; h++) {
        PPO_REQUIRE(h_nvec[h] >= 1, "ppo_categorical: nvec[%d]=%d < 1", h, h_nvec[h]);
        a.nvec[h] = h_nvec[h];
        a.K += h_nvec[h];
    }
    PPO_REQUIRE(a.K <= PPO_MAX_LOGITS, "ppo_categorical: sum(nvec)=%d exceeds %d", a.K, PPO_MAX_LOGITS);
    if (n == 0) return PPO_OK;
    a.logits = logits; a.mask = mask; a.actions_in = actions_in; a.uniforms = uniforms;
    a.seed = seed; a.counter = counter; a.entropy_mode = entropy_mode;
    a.actions_out = actions_out; a.logprob = logprob_out; a.entropy = entropy_out; a.probs_out = probs_out;
    const int ld = a.K | 1;
    const size_t smem = (size_t)CAT_ROWS * ld * 4 + (size_t)CAT_ROWS * heads * 8 + (size_t)CAT_ROWS * a.K + 16;
    const int blocks = (n + CAT_ROWS - 1) / CAT_ROWS;
    categorical_kernel<<<blocks, CAT_ROWS, smem, as_stream(stream)>>>(a);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}
