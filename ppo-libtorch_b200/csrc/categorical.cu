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

// ---- warp-cooperative kernel: one (row, head) per lane --------------------------------------------------------------------
// The row-per-thread kernel above spends ~4 300 thread-instructions per 21-logit row (IEEE expf / logf / division per
// logit, twice) and runs at 16 % of the HBM roofline.  Here a lane owns ONE head of one row (heads padded to HP = 1, 2, 4 or
// 8 lanes, 32 / HP rows per pass, 32 rows per warp), its <= 8 logits stay in registers, the row's log-prob and entropy are
// xor-shuffle sums over the row's lanes, and no shared-memory staging is needed: the lanes of a pass read one contiguous
// span of logits.  Arithmetic: ex2 / lg2 / rcp approximations (2 ulp: probabilities within 1e-6, log-probs within 1e-6 of
// the IEEE path - inside the 4e-6 / 1e-5 parity bounds).  The ONE place where a last-bit difference is visible is the
// inverse-CDF compare u < cumsum(p): a draw closer than CAT_BAND to a CDF edge is re-evaluated with the IEEE routines
// (categorical.cuh: the oracle's operation order), so sampled actions are those of the row-per-thread kernel.
// Philox: lane l draws the blocks of row (base + l) once per 32 rows; a pass fetches its uniform with shuffles.
constexpr float CAT_BAND = 4e-6f;
constexpr int CAT_MAXKH_LIMIT = 8;

// HP = lanes per row (heads padded to a power of two), CAT_MAXKH = compile-time bound of the logits per head (the
// per-logit loops are fully unrolled and predicated on the lane's own head size)
template <int HP, int CAT_MAXKH>
__global__ void __launch_bounds__(256) categorical_lane_kernel(const CatArgs a) {
    constexpr int RPW = 32 / HP;
    __shared__ uint32_t us[8][32][PPO_MAX_HEADS + 1];   // per warp: the Philox words of its 32 rows (odd stride: no conflicts)
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t row_base = ((int64_t)blockIdx.x * (blockDim.x >> 5) + wib) * 32;
    if (row_base >= a.n) return;
    const int h = lane % HP, rl = lane / HP;
    const int heads = a.heads, K = a.K;
    const bool hon = h < heads;
    int off = 0;
#pragma unroll
    for (int hh = 0; hh < PPO_MAX_HEADS; hh++)
        if (hh < h && hh < heads) off += a.nvec[hh];
    const int Kh = hon ? a.nvec[h] : 0;
    const bool masked = a.mask != nullptr;
    const bool sample = a.actions_in == nullptr;
    const bool philox = sample && a.uniforms == nullptr;
    const bool ref_quirk = !masked && a.entropy_mode == PPO_ENTROPY_REFERENCE;

    if (philox) {   // lane l draws the blocks of row (row_base + l) once; a pass reads its word back from shared memory
        uint32_t r[4] = {0u, 0u, 0u, 0u};
        const bool live = row_base + lane < a.n;
        if (live)
            Philox::gen(a.seed, (uint32_t)(row_base + lane), RNG_STREAM_SAMPLE, (uint32_t)a.counter, (uint32_t)(a.counter >> 32), r);
        us[wib][lane][0] = r[0]; us[wib][lane][1] = r[1]; us[wib][lane][2] = r[2]; us[wib][lane][3] = r[3];
        if (heads > 4) {
            if (live)
                Philox::gen(a.seed, (uint32_t)(row_base + lane), RNG_STREAM_SAMPLE | (1u << 8), (uint32_t)a.counter,
                            (uint32_t)(a.counter >> 32), r);
            us[wib][lane][4] = r[0]; us[wib][lane][5] = r[1]; us[wib][lane][6] = r[2]; us[wib][lane][7] = r[3];
        }
        __syncwarp();
    }
    // the per-logit loops below are branch-free: a slot beyond the lane's head size carries x = -inf (e = p = 0) and
    // every conditional is a select, so lanes with 2- and 3-logit heads run the same instruction stream
#pragma unroll 1
    for (int p = 0; p < HP; p++) {
        const int r_local = p * RPW + rl;
        const int64_t row = row_base + r_local;
        const bool on = hon && row < a.n;
        const int64_t base = on ? row * K + off : 0;
        float u = 0.0f;
        if (philox) u = Philox::u01(us[wib][r_local][hon ? h : 0]);
        else if (sample && on) u = __ldg(a.uniforms + row * heads + h);
        const int given = (!sample && on) ? __ldg(a.actions_in + row * heads + h) : -1;
        float xm[CAT_MAXKH], e[CAT_MAXKH];
        uint32_t vbits = 0u;
        float mx = -INFINITY;
#pragma unroll
        for (int j = 0; j < CAT_MAXKH; j++) {
            const bool in = on && j < Kh;
            const int64_t idx = base + (in ? j : 0);
            const float x = __ldg(a.logits + idx);
            const bool v = in && (!masked || __ldg(a.mask + idx) != 0);
            xm[j] = in ? (v ? x : -1e8f) : -INFINITY;      // CategoricalMasked.cpp:35
            vbits |= (uint32_t)v << j;
            mx = fmaxf(mx, xm[j]);
        }
        mx = on ? mx : 0.0f;
        float sum = 0.0f;
#pragma unroll
        for (int j = 0; j < CAT_MAXKH; j++) {
            asm("ex2.approx.f32 %0, %1;" : "=f"(e[j]) : "f"((xm[j] - mx) * 1.4426950408889634f));
            sum += e[j];
        }
        sum = on ? sum : 1.0f;
        float rs, lg;
        asm("rcp.approx.f32 %0, %1;" : "=f"(rs) : "f"(sum));
        asm("lg2.approx.f32 %0, %1;" : "=f"(lg) : "f"(sum));
        const float lse = fmaf(lg, 0.6931471805599453f, mx);
        float c = 0.0f, acc = 0.0f, chi = 2.0f, clo = -1.0f;
        int last_valid = 0, apick = -1;
#pragma unroll
        for (int j = 0; j < CAT_MAXKH; j++) {
            const float pj = e[j] * rs, lp = xm[j] - lse;
            const bool pos = pj > 0.0f;
            float t = ref_quirk ? __fmul_rn(fmaxf(lp, FLT_MIN), pj) : lp * pj;
            t = (ref_quirk ? j < Kh : (pos && ((vbits >> j) & 1u))) ? t : 0.0f;
            acc += t;
            last_valid = pos ? j : last_valid;
            const float cn = c + pj;
            const bool hit = apick < 0 && u < cn && j < Kh;
            clo = hit ? c : clo;
            chi = hit ? cn : chi;
            apick = hit ? j : apick;
            c = cn;
        }
        if (a.probs_out && on) {
#pragma unroll
            for (int j = 0; j < CAT_MAXKH; j++)
                if (j < Kh) a.probs_out[base + j] = e[j] * rs;
        }
        int pick = given;
        if (sample) {
            pick = apick < 0 ? last_valid : apick;
            // the draw decided between clo <= u < chi: closer than the band to either edge (or past the last edge), the
            // oracle's IEEE operation order decides
            const bool near = apick < 0 || chi - u < CAT_BAND || u - clo < CAT_BAND;
            if (near && on) {
                float mxe = -INFINITY;
#pragma unroll
                for (int j = 0; j < CAT_MAXKH; j++) mxe = fmaxf(mxe, xm[j]);
                float se = 0.0f;
#pragma unroll
                for (int j = 0; j < CAT_MAXKH; j++)
                    if (j < Kh) se = __fadd_rn(se, expf(__fsub_rn(xm[j], mxe)));
                float ce = 0.0f;
                int ae = -1, lv = 0;
#pragma unroll
                for (int j = 0; j < CAT_MAXKH; j++) {
                    if (j < Kh) {
                        const float pe = __fdiv_rn(expf(__fsub_rn(xm[j], mxe)), se);
                        if (pe > 0.0f) lv = j;
                        ce = __fadd_rn(ce, pe);
                        if (ae < 0 && u < ce) ae = j;
                    }
                }
                pick = ae < 0 ? lv : ae;
            }
        }
        float lpa = 0.0f;
#pragma unroll
        for (int j = 0; j < CAT_MAXKH; j++) lpa = j == pick ? xm[j] - lse : lpa;
        float lpr = on ? lpa : 0.0f, enr = on ? -acc : 0.0f;
#pragma unroll
        for (int o = 1; o < HP; o <<= 1) {
            lpr += __shfl_xor_sync(0xffffffffu, lpr, o);
            enr += __shfl_xor_sync(0xffffffffu, enr, o);
        }
        if (on && a.actions_out) a.actions_out[row * heads + h] = pick;
        if (h == 0 && row < a.n) {
            if (a.logprob) a.logprob[row] = lpr;
            if (a.entropy) a.entropy[row] = enr;
        }
    }
}

// m_logits = x' - logsumexp(x') and m_probs = softmax(x') (Categorical.cpp:33-35, CategoricalMasked.cpp:35-40) for the
// distribution classes' public members: one thread per (row, head), IEEE routines of categorical.cuh
__global__ void categorical_normalize_kernel(const CatArgs a, float* __restrict__ logits_out, float* __restrict__ probs_out) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int heads = a.heads;
    if (t >= (int64_t)a.n * heads) return;
    const int64_t row = t / heads;
    const int h = (int)(t - row * heads);
    int off = 0;
    for (int hh = 0; hh < h; hh++) off += a.nvec[hh];
    const int Kh = a.nvec[h];
    const float* x = a.logits + row * a.K + off;
    const bool masked = a.mask != nullptr;
    const uint64_t mb = masked ? pack_mask(a.mask + row * a.K + off, Kh) : ~0ull;
    const HeadStats s = head_stats(x, Kh, mb, masked);
    for (int j = 0; j < Kh; j++) {
        const float xm = masked_logit(x[j], !masked || ((mb >> j) & 1ull));
        if (logits_out) logits_out[row * a.K + off + j] = head_logp(xm, s);
        if (probs_out) probs_out[row * a.K + off + j] = head_prob(xm, s);
    }
}

static int g_cat_impl = 1;   // ppo_debug_categorical_impl: 0 = row-per-thread kernel for every shape

}  // namespace ppo

using namespace ppo;

extern "C" int ppo_debug_categorical_impl(int impl) {
    g_cat_impl = impl;
    return PPO_OK;
}


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
    int maxk = 0;
    for (int h = 0; h < heads; h++) maxk = a.nvec[h] > maxk ? a.nvec[h] : maxk;
    if (g_cat_impl && maxk <= CAT_MAXKH_LIMIT) {   // a (row, head) per lane; heads padded to a power of two
        const int blocks = (n + 255) / 256;
        const int hp = heads <= 1 ? 1 : (heads <= 2 ? 2 : (heads <= 4 ? 4 : 8));
        const int mk = maxk <= 2 ? 2 : (maxk <= 3 ? 3 : (maxk <= 4 ? 4 : 8));
        cudaStream_t st = as_stream(stream);
#define CAT_LAUNCH(HPV, MKV) categorical_lane_kernel<HPV, MKV><<<blocks, 256, 0, st>>>(a)
#define CAT_BY_MK(HPV) (mk == 2 ? CAT_LAUNCH(HPV, 2) : mk == 3 ? CAT_LAUNCH(HPV, 3) : mk == 4 ? CAT_LAUNCH(HPV, 4) : CAT_LAUNCH(HPV, 8))
        if (hp == 1) CAT_BY_MK(1);
        else if (hp == 2) CAT_BY_MK(2);
        else if (hp == 4) CAT_BY_MK(4);
        else CAT_BY_MK(8);
#undef CAT_BY_MK
#undef CAT_LAUNCH
        PPO_LAUNCH_CHECK();
        return PPO_OK;
    }
    const int ld = a.K | 1;
    const size_t smem = (size_t)CAT_ROWS * ld * 4 + (size_t)CAT_ROWS * heads * 8 + (size_t)CAT_ROWS * a.K + 16;
    const int blocks = (n + CAT_ROWS - 1) / CAT_ROWS;
    categorical_kernel<<<blocks, CAT_ROWS, smem, as_stream(stream)>>>(a);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

extern "C" int ppo_categorical_normalize(int n, int heads, const int32_t* h_nvec, const float* logits, const uint8_t* mask,
                                         float* logits_norm_out, float* probs_out, ppo_stream_t stream) {
    PPO_REQUIRE(n >= 0 && logits && h_nvec && (logits_norm_out || probs_out), "ppo_categorical_normalize: null argument");
    PPO_REQUIRE(heads >= 1 && heads <= PPO_MAX_HEADS, "ppo_categorical_normalize: heads=%d out of [1,%d]", heads, PPO_MAX_HEADS);
    CatArgs a{};
    a.n = n; a.heads = heads; a.K = 0;
    for (int h = 0; h < heads; h++) {
        PPO_REQUIRE(h_nvec[h] >= 1 && h_nvec[h] <= 64, "ppo_categorical_normalize: nvec[%d]=%d", h, h_nvec[h]);
        a.nvec[h] = h_nvec[h];
        a.K += h_nvec[h];
    }
    PPO_REQUIRE(a.K <= PPO_MAX_LOGITS, "ppo_categorical_normalize: sum(nvec)=%d exceeds %d", a.K, PPO_MAX_LOGITS);
    if (n == 0) return PPO_OK;
    a.logits = logits; a.mask = mask;
    const int64_t total = (int64_t)n * heads;
    categorical_normalize_kernel<<<(unsigned)((total + 255) / 256), 256, 0, as_stream(stream)>>>(a, logits_norm_out, probs_out);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}
