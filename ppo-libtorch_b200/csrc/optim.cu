// optim.cu — K6 and the small reductions around the update:
//   ppo_clip_adamw          clip_grad_norm_ + AdamW in one launch   (PPO/PPO_Discrete.cpp:640-641)
//   ppo_adv_stats(_finalize) per-minibatch advantage mean / unbiased std (:591-594)
//   ppo_explained_variance  1 - var(R - V)/var(R)                   (:647-648)
//   ppo_make_permutation    counter-based replacement of torch::randperm (:569)
#include <mutex>

#include "common.cuh"

namespace ppo {

// ------------------------------------------------------------------------------------------------
// block-wide sum of doubles in a fixed order (deterministic); result valid in every thread
template <int THREADS>
__device__ inline double block_sum(double v, double* sh /* [THREADS/32] */) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < THREADS / 32; w++) t += sh[w];
    return t;
}

// ------------------------------------------------------------------------------------------------
// K6.  Every CTA recomputes the full gradient norm in the same order (P <= 134k floats, L2-resident),
// so all CTAs agree bit-for-bit on the clip coefficient without a grid barrier; each then applies
// clip + decoupled weight decay + Adam to its own slice.  `grads` is read-only here (other CTAs are
// still reading it for their norm), so unlike clip_grad_norm_ the stored gradient stays PRE-clip.
//   torch/nn/utils/clip_grad.h: coef = max_norm / (norm + 1e-6), applied iff < 1
//   torch/csrc/api/src/optim/adamw.cpp: p *= 1 - lr*wd; m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2;
//                                       p -= lr/bc1 * m / (sqrt(v)/sqrt(bc2) + eps)
constexpr int ADAM_THREADS = 512;
__global__ void __launch_bounds__(ADAM_THREADS) clip_adamw_kernel(
    float* __restrict__ params, const float* __restrict__ grads, float* __restrict__ m, float* __restrict__ v, int64_t n,
    float lr_host, const float* __restrict__ d_lr, float beta1, float beta2, float eps, float weight_decay,
    int64_t step_host, const int64_t* __restrict__ d_step, float max_grad_norm, float grad_scale,
    float* __restrict__ norm_out, const float* __restrict__ diag_src, float* __restrict__ diag_dst) {
    __shared__ double sh[ADAM_THREADS / 32];
    if (diag_dst && blockIdx.x == 0 && threadIdx.x < 8) diag_dst[threadIdx.x] = diag_src[threadIdx.x];
    double ss = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += ADAM_THREADS) {
        const float gsc = grads[i] * grad_scale;
        ss += (double)gsc * (double)gsc;
    }
    const double norm = sqrt(block_sum<ADAM_THREADS>(ss, sh));
    if (norm_out && blockIdx.x == 0 && threadIdx.x == 0) *norm_out = (float)norm;
    const double coef = (double)max_grad_norm / (norm + 1e-6);
    const float cf = (max_grad_norm > 0.0f && coef < 1.0) ? (float)coef : 1.0f;

    const float lr = d_lr ? *d_lr : lr_host;
    const int64_t step = step_host + (d_step ? *d_step : 0);
    const double b1 = (double)beta1, b2 = (double)beta2;
    const double bc1 = 1.0 - pow(b1, (double)step);
    const double bc2 = 1.0 - pow(b2, (double)step);
    const float decay = (float)(1.0 - (double)lr * (double)weight_decay);
    const float omb1 = (float)(1.0 - b1), omb2 = (float)(1.0 - b2);
    const float sq_bc2 = (float)sqrt(bc2);
    const float neg_step = (float)(-((double)lr / bc1));
    for (int64_t i = (int64_t)blockIdx.x * ADAM_THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * ADAM_THREADS) {
        float g = __fmul_rn(grads[i], grad_scale);
        if (cf != 1.0f) g = __fmul_rn(g, cf);
        const float p = __fmul_rn(params[i], decay);
        const float mi = __fadd_rn(__fmul_rn(m[i], beta1), __fmul_rn(g, omb1));
        const float vi = __fadd_rn(__fmul_rn(v[i], beta2), __fmul_rn(__fmul_rn(g, g), omb2));
        m[i] = mi;
        v[i] = vi;
        const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(vi), sq_bc2), eps);
        params[i] = __fadd_rn(p, __fmul_rn(neg_step, __fdiv_rn(mi, denom)));
    }
}

// ------------------------------------------------------------------------------------------------
// advantage statistics: one CTA per minibatch, two passes in double like LibTorch's mean()/std()
constexpr int STAT_THREADS = 1024;
__global__ void __launch_bounds__(STAT_THREADS) adv_stats_kernel(const float* __restrict__ adv,
                                                                 const int32_t* __restrict__ perm, int64_t batch,
                                                                 int mb_size, float* __restrict__ stats_out,
                                                                 double* __restrict__ partial_out) {
    __shared__ double sh[STAT_THREADS / 32];
    const int mb = blockIdx.x;
    const int64_t start = (int64_t)mb * mb_size;
    int64_t cnt = batch - start;
    if (cnt > mb_size) cnt = mb_size;
    if (cnt < 0) cnt = 0;
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < cnt; i += STAT_THREADS) {
        const int64_t r = perm ? (int64_t)perm[start + i] : start + i;
        s += (double)__ldg(adv + r);
    }
    const double sum = block_sum<STAT_THREADS>(s, sh);
    const double mean = cnt > 0 ? sum / (double)cnt : 0.0;
    if (partial_out) {
        // raw moments for the data-parallel merge
        double q = 0.0;
        for (int64_t i = threadIdx.x; i < cnt; i += STAT_THREADS) {
            const int64_t r = perm ? (int64_t)perm[start + i] : start + i;
            const double x = (double)__ldg(adv + r);
            q += x * x;
        }
        const double sq = block_sum<STAT_THREADS>(q, sh);
        if (threadIdx.x == 0) {
            partial_out[mb * 3 + 0] = (double)cnt;
            partial_out[mb * 3 + 1] = sum;
            partial_out[mb * 3 + 2] = sq;
        }
    }
    if (stats_out) {
        double q = 0.0;
        for (int64_t i = threadIdx.x; i < cnt; i += STAT_THREADS) {
            const int64_t r = perm ? (int64_t)perm[start + i] : start + i;
            const double dlt = (double)__ldg(adv + r) - mean;
            q += dlt * dlt;
        }
        const double ssd = block_sum<STAT_THREADS>(q, sh);
        if (threadIdx.x == 0) {
            stats_out[mb * 2 + 0] = (float)mean;
            stats_out[mb * 2 + 1] = (float)sqrt(ssd / (double)(cnt - 1));
        }
    }
}

// Sliced variant: ADV_SLICES CTAs per minibatch accumulate raw moments in double; the last CTA of a minibatch to
// finish (ticket) sums the slices in fixed order and writes the statistics.  workspace = [n_mb][ADV_SLICES][2] doubles
// + [n_mb] uint32 tickets, zeroed once by the caller and left zeroed.
constexpr int ADV_SLICES = 32, ADV_THREADS = 256;
// Workspace layout shared by the three sliced kernels below (a trainer hands the same buffer to all of them): a fixed
// header of ADV_TICKETS uint32 arrival tickets, then the double partials.  (Until round 2 every kernel put its tickets
// BEHIND its own partials, i.e. inside another kernel's partials: after an all-epoch launch the per-epoch kernel found
// garbage tickets, never elected a last CTA and left the statistics of the previous launch in place.)
constexpr int ADV_TICKETS = 1024;
__device__ __forceinline__ unsigned int* adv_tickets(double* ws) { return reinterpret_cast<unsigned int*>(ws); }
__device__ __forceinline__ double* adv_partials(double* ws) { return ws + ADV_TICKETS / 2; }
__global__ void __launch_bounds__(ADV_THREADS) adv_moments_kernel(const float* __restrict__ adv,
                                                                  const int32_t* __restrict__ perm, int64_t batch,
                                                                  int mb_size, int n_mb, double* __restrict__ ws,
                                                                  float* __restrict__ stats_out,
                                                                  double* __restrict__ partial_out) {
    __shared__ double sh[ADV_THREADS / 32];
    __shared__ bool last;
    const int mb = blockIdx.x, sl = blockIdx.y;
    const int64_t start = (int64_t)mb * mb_size;
    int64_t cnt = batch - start;
    if (cnt > mb_size) cnt = mb_size;
    if (cnt < 0) cnt = 0;
    const int64_t per = (cnt + ADV_SLICES - 1) / ADV_SLICES;
    const int64_t i0 = sl * per, i1 = i0 + per < cnt ? i0 + per : cnt;
    double s = 0.0, q = 0.0;
    for (int64_t i = i0 + threadIdx.x; i < i1; i += ADV_THREADS) {
        const int64_t r = perm ? (int64_t)perm[start + i] : start + i;
        const double x = (double)__ldg(adv + r);
        s += x;
        q += x * x;
    }
    s = block_sum<ADV_THREADS>(s, sh);
    q = block_sum<ADV_THREADS>(q, sh);
    unsigned int* tickets = adv_tickets(ws);
    ws = adv_partials(ws);
    if (threadIdx.x == 0) {
        ws[((size_t)mb * ADV_SLICES + sl) * 2 + 0] = s;
        ws[((size_t)mb * ADV_SLICES + sl) * 2 + 1] = q;
        __threadfence();
        last = (atomicAdd(tickets + mb, 1u) == ADV_SLICES - 1);
    }
    __syncthreads();
    if (!last || threadIdx.x != 0) return;
    __threadfence();
    double S = 0.0, Q = 0.0;
    for (int k = 0; k < ADV_SLICES; k++) {
        S += ws[((size_t)mb * ADV_SLICES + k) * 2 + 0];
        Q += ws[((size_t)mb * ADV_SLICES + k) * 2 + 1];
    }
    tickets[mb] = 0;
    const double n = (double)cnt;
    if (partial_out) {
        partial_out[mb * 3 + 0] = n; partial_out[mb * 3 + 1] = S; partial_out[mb * 3 + 2] = Q;
    }
    if (stats_out) {
        const double mean = cnt > 0 ? S / n : 0.0;
        double var = (Q - n * mean * mean) / (n - 1.0);
        if (var < 0.0) var = 0.0;
        stats_out[mb * 2 + 0] = (float)mean;
        stats_out[mb * 2 + 1] = (float)sqrt(var);
    }
}

__global__ void adv_stats_finalize_kernel(const double* __restrict__ partial, int n_mb, float* __restrict__ stats_out) {
    const int mb = blockIdx.x * blockDim.x + threadIdx.x;
    if (mb >= n_mb) return;
    const double n = partial[mb * 3 + 0], s = partial[mb * 3 + 1], q = partial[mb * 3 + 2];
    const double mean = s / n;
    double var = (q - n * mean * mean) / (n - 1.0);
    if (var < 0.0) var = 0.0;
    stats_out[mb * 2 + 0] = (float)mean;
    stats_out[mb * 2 + 1] = (float)sqrt(var);
}

// ------------------------------------------------------------------------------------------------
// explained variance: grid-wide raw moments in double, finished by the last CTA to arrive.
// workspace: 4*EV_MAX_BLOCKS doubles + one uint32 ticket, zero-initialised; left zeroed on exit.
constexpr int EV_THREADS = 256;
constexpr int EV_MAX_BLOCKS = 1024;
__global__ void __launch_bounds__(EV_THREADS) explained_variance_kernel(int64_t n, const float* __restrict__ ret,
                                                                        const float* __restrict__ val,
                                                                        double* __restrict__ part,
                                                                        unsigned int* __restrict__ ticket,
                                                                        float* __restrict__ out) {
    __shared__ double sh[EV_THREADS / 32];
    __shared__ bool last;
    double sr = 0, srr = 0, sd = 0, sdd = 0;
    for (int64_t i = (int64_t)blockIdx.x * EV_THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * EV_THREADS) {
        const float r = __ldg(ret + i);
        const double rr = (double)r, d = (double)(r - __ldg(val + i));
        sr += rr; srr += rr * rr; sd += d; sdd += d * d;
    }
    sr = block_sum<EV_THREADS>(sr, sh);
    srr = block_sum<EV_THREADS>(srr, sh);
    sd = block_sum<EV_THREADS>(sd, sh);
    sdd = block_sum<EV_THREADS>(sdd, sh);
    if (threadIdx.x == 0) {
        part[blockIdx.x * 4 + 0] = sr; part[blockIdx.x * 4 + 1] = srr;
        part[blockIdx.x * 4 + 2] = sd; part[blockIdx.x * 4 + 3] = sdd;
        __threadfence();
        last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    double a[4] = {0, 0, 0, 0};
    for (int b = threadIdx.x; b < (int)gridDim.x; b += EV_THREADS)
#pragma unroll
        for (int k = 0; k < 4; k++) a[k] += part[b * 4 + k];
#pragma unroll
    for (int k = 0; k < 4; k++) a[k] = block_sum<EV_THREADS>(a[k], sh);
    if (threadIdx.x == 0) {
        const double N = (double)n;
        const double vr = (a[1] - a[0] * a[0] / N) / (N - 1.0);
        const double vd = (a[3] - a[2] * a[2] / N) / (N - 1.0);
        *out = (float)(1.0 - vd / vr);
        *ticket = 0;
    }
}

// ------------------------------------------------------------------------------------------------
// Permutation of [0, batch): a 4-round Feistel network on the smallest even-width bit domain that
// covers `batch`, with cycle walking back into range (a bijection for every batch size).
__global__ void __launch_bounds__(256) permutation_kernel(int64_t batch, int half_bits, uint64_t seed, uint64_t epoch,
                                                          const uint64_t* __restrict__ d_epoch,
                                                          int32_t* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= batch) return;
    if (d_epoch) epoch += *d_epoch;
    uint32_t r[4];
    perm_keys(seed, epoch, r);
    out[i] = (int32_t)perm_index(i, batch, half_bits, r);
}

// adv_moments_kernel for ALL epochs of an update at once: the advantages are fixed during the update and the
// permutation of epoch e is a function of (seed, e), so the E x MB minibatch statistics (and, under data parallelism,
// their ONE allreduce) can be produced before the first optimiser step.  blockIdx.x = e * n_mb + mb; same slicing and
// summation order as adv_moments_kernel => the same numbers as the per-epoch path.
__global__ void __launch_bounds__(ADV_THREADS) adv_moments_all_kernel(const float* __restrict__ adv, int64_t batch,
                                                                      int mb_size, int n_mb, int n_epochs, int half_bits,
                                                                      uint64_t seed, uint64_t epoch0,
                                                                      const uint64_t* __restrict__ d_epoch,
                                                                      double* __restrict__ ws, float* __restrict__ stats_out,
                                                                      double* __restrict__ partial_out) {
    __shared__ double sh[ADV_THREADS / 32];
    __shared__ bool last;
    const int emb = blockIdx.x, e = emb / n_mb, mb = emb % n_mb, sl = blockIdx.y;
    uint32_t key[4];
    perm_keys(seed, epoch0 + (uint64_t)e + (d_epoch ? *d_epoch : 0ull), key);
    const int64_t start = (int64_t)mb * mb_size;
    int64_t cnt = batch - start;
    if (cnt > mb_size) cnt = mb_size;
    if (cnt < 0) cnt = 0;
    const int64_t per = (cnt + ADV_SLICES - 1) / ADV_SLICES;
    const int64_t i0 = sl * per, i1 = i0 + per < cnt ? i0 + per : cnt;
    double s = 0.0, q = 0.0;
    // four independent index evaluations + gathers in flight per thread; the sums are added in the same element order
    // as the one-at-a-time loop (x0, x1, x2, x3 of a thread are consecutive terms of its chain)
    int64_t i = i0 + threadIdx.x;
    for (; i + 3 * ADV_THREADS < i1; i += 4 * ADV_THREADS) {
        float x[4];
#pragma unroll
        for (int u = 0; u < 4; u++) x[u] = __ldg(adv + perm_index(start + i + u * ADV_THREADS, batch, half_bits, key));
#pragma unroll
        for (int u = 0; u < 4; u++) {
            s += (double)x[u];
            q += (double)x[u] * (double)x[u];
        }
    }
    for (; i < i1; i += ADV_THREADS) {
        const double x = (double)__ldg(adv + perm_index(start + i, batch, half_bits, key));
        s += x;
        q += x * x;
    }
    s = block_sum<ADV_THREADS>(s, sh);
    q = block_sum<ADV_THREADS>(q, sh);
    unsigned int* tickets = adv_tickets(ws);
    ws = adv_partials(ws);
    if (threadIdx.x == 0) {
        ws[((size_t)emb * ADV_SLICES + sl) * 2 + 0] = s;
        ws[((size_t)emb * ADV_SLICES + sl) * 2 + 1] = q;
        __threadfence();
        last = (atomicAdd(tickets + emb, 1u) == ADV_SLICES - 1);
    }
    __syncthreads();
    if (!last || threadIdx.x != 0) return;
    __threadfence();
    double S = 0.0, Q = 0.0;
    for (int k = 0; k < ADV_SLICES; k++) {
        S += ws[((size_t)emb * ADV_SLICES + k) * 2 + 0];
        Q += ws[((size_t)emb * ADV_SLICES + k) * 2 + 1];
    }
    tickets[emb] = 0;
    const double n = (double)cnt;
    if (partial_out) {
        partial_out[emb * 3 + 0] = n; partial_out[emb * 3 + 1] = S; partial_out[emb * 3 + 2] = Q;
    }
    if (stats_out) {
        const double mean = cnt > 0 ? S / n : 0.0;
        double var = (Q - n * mean * mean) / (n - 1.0);
        if (var < 0.0) var = 0.0;
        stats_out[emb * 2 + 0] = (float)mean;
        stats_out[emb * 2 + 1] = (float)sqrt(var);
    }
}


// The same statistics from ONE coalesced pass per epoch over the advantages: row j sits at position perm_position(j) of
// epoch e's order (the inverse of the keyed bijection), i.e. in minibatch position / mb_size.  adv_moments_all_kernel's
// permuted 4-byte gathers are fine while `adv` is L2-resident (< ~30 M rows) but at the 134 M rows of BASELINE configs[4]
// on one GPU every gather is a 32-byte DRAM sector: 34.8 ms of a 373 ms iteration, against E coalesced reads of the array
// here.  Grid (epoch, chunk): a CTA sums its chunk's rows into per-minibatch registers (double), ws[e][mb][chunk]; the
// last CTA of an epoch adds the chunks in fixed order => run-to-run identical.
constexpr int ADV_MBMAX = 8;
constexpr int ADV_CSLICES_MAX = 128;   // slices per epoch of the coalesced kernel (grid.y, chosen at launch): workspace bound
__global__ void __launch_bounds__(ADV_THREADS) adv_moments_all_coalesced_kernel(
    const float* __restrict__ adv, int64_t batch, int mb_size, int n_mb, int n_epochs, int half_bits, uint64_t seed,
    uint64_t epoch0, const uint64_t* __restrict__ d_epoch, double* __restrict__ ws, float* __restrict__ stats_out,
    double* __restrict__ partial_out) {
    __shared__ double sh[ADV_THREADS / 32];
    __shared__ bool last;
    const int e = blockIdx.x, c = blockIdx.y;
    uint32_t key[4];
    perm_keys(seed, epoch0 + (uint64_t)e + (d_epoch ? *d_epoch : 0ull), key);
    const int nsl = (int)gridDim.y;
    const int64_t per = (((batch + nsl - 1) / nsl) + 3) & ~(int64_t)3;
    const int64_t i0 = c * per, i1 = i0 + per < batch ? i0 + per : batch;
    // fp64 adds issue at ~2 per clock and SM on this part (the first version, 16 predicated double adds per row, ran at
    // 7 clocks per row): rows go into fp32 per-minibatch registers, flushed into the double accumulators every 32 rows
    double s[ADV_MBMAX], q[ADV_MBMAX];
    float sf[ADV_MBMAX], qf[ADV_MBMAX];
#pragma unroll
    for (int m = 0; m < ADV_MBMAX; m++) { s[m] = 0.0; q[m] = 0.0; sf[m] = 0.0f; qf[m] = 0.0f; }
    const float inv_mb = 1.0f / (float)mb_size;
    auto add = [&](float x, int64_t j) {
        const int64_t pos = perm_position(j, batch, half_bits, key);
        int mb = (int)((float)pos * inv_mb);                 // pos / mb_size: float estimate, corrected by one either way
        if ((int64_t)mb * mb_size > pos) mb--;
        else if ((int64_t)(mb + 1) * mb_size <= pos) mb++;
        const float xx = x * x;
#pragma unroll
        for (int m = 0; m < ADV_MBMAX; m++) {
            sf[m] += m == mb ? x : 0.0f;
            qf[m] += m == mb ? xx : 0.0f;
        }
    };
    auto flush = [&]() {
#pragma unroll
        for (int m = 0; m < ADV_MBMAX; m++) {
            s[m] += (double)sf[m]; q[m] += (double)qf[m];
            sf[m] = 0.0f; qf[m] = 0.0f;
        }
    };
    int64_t i = i0 + threadIdx.x * 4;
    int since = 0;
    for (; i + 3 < i1; i += 4 * ADV_THREADS) {   // i0 and per are multiples of 4: 16-byte aligned vector loads
        const float4 v = __ldg(reinterpret_cast<const float4*>(adv + i));
        add(v.x, i); add(v.y, i + 1); add(v.z, i + 2); add(v.w, i + 3);
        if (++since == 8) { flush(); since = 0; }
    }
    for (int64_t t = i; t < i1 && t < i + 4; t++) add(__ldg(adv + t), t);   // the ragged tail of the last chunk
    flush();
    unsigned int* tickets = adv_tickets(ws);
    ws = adv_partials(ws);
#pragma unroll
    for (int m = 0; m < ADV_MBMAX; m++) {
        if (m < n_mb) {
            const double S = block_sum<ADV_THREADS>(s[m], sh);
            const double Q = block_sum<ADV_THREADS>(q[m], sh);
            if (threadIdx.x == 0) {
                ws[((size_t)(e * n_mb + m) * nsl + c) * 2 + 0] = S;
                ws[((size_t)(e * n_mb + m) * nsl + c) * 2 + 1] = Q;
            }
        }
    }
    if (threadIdx.x == 0) {
        __threadfence();
        last = (atomicAdd(tickets + e, 1u) == (unsigned)nsl - 1u);
    }
    __syncthreads();
    if (!last || threadIdx.x >= n_mb) return;
    __threadfence();
    const int m = threadIdx.x, emb = e * n_mb + m;
    double S = 0.0, Q = 0.0;
    for (int k = 0; k < nsl; k++) {
        S += ws[((size_t)emb * nsl + k) * 2 + 0];
        Q += ws[((size_t)emb * nsl + k) * 2 + 1];
    }
    int64_t cnt = batch - (int64_t)m * mb_size;
    if (cnt > mb_size) cnt = mb_size;
    if (cnt < 0) cnt = 0;
    const double n = (double)cnt;
    if (partial_out) {
        partial_out[emb * 3 + 0] = n; partial_out[emb * 3 + 1] = S; partial_out[emb * 3 + 2] = Q;
    }
    if (stats_out) {
        const double mean = cnt > 0 ? S / n : 0.0;
        double var = (Q - n * mean * mean) / (n - 1.0);
        if (var < 0.0) var = 0.0;
        stats_out[emb * 2 + 0] = (float)mean;
        stats_out[emb * 2 + 1] = (float)sqrt(var);
    }
    if (threadIdx.x == 0) tickets[e] = 0;   // leave the workspace zeroed for the next launch
}
int g_adv_impl = 1;   // ppo_debug_adv_impl: 0 = permuted-gather kernel always, 1 = coalesced from 2^25 rows, 2 = coalesced always

}  // namespace ppo

using namespace ppo;

extern "C" int ppo_debug_adv_impl(int impl) {
    g_adv_impl = impl;
    return PPO_OK;
}

namespace ppo {
int clip_adamw_launch(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t n_params, float lr,
                      const float* d_lr, float beta1, float beta2, float eps, float weight_decay, int64_t step,
                      const int64_t* d_step, float max_grad_norm, float grad_scale, float* norm_out,
                      const float* diag_src, float* diag_dst, cudaStream_t st) {
    PPO_REQUIRE(params && grads && exp_avg && exp_avg_sq && n_params >= 0, "ppo_clip_adamw: null buffer");
    if (n_params == 0) return PPO_OK;
    int blocks = (int)((n_params + ADAM_THREADS * 4 - 1) / (ADAM_THREADS * 4));
    if (blocks > 64) blocks = 64;
    clip_adamw_kernel<<<blocks, ADAM_THREADS, 0, st>>>(params, grads, exp_avg, exp_avg_sq, n_params, lr, d_lr, beta1,
                                                       beta2, eps, weight_decay, step, d_step, max_grad_norm,
                                                       grad_scale, norm_out, diag_src, diag_dst);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}
}  // namespace ppo

extern "C" int ppo_clip_adamw(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t n_params,
                              float lr, const float* d_lr, float beta1, float beta2, float eps, float weight_decay,
                              int64_t step, const int64_t* d_step, float max_grad_norm, float grad_scale,
                              float* norm_out, ppo_stream_t stream) {
    return clip_adamw_launch(params, grads, exp_avg, exp_avg_sq, n_params, lr, d_lr, beta1, beta2, eps, weight_decay,
                             step, d_step, max_grad_norm, grad_scale, norm_out, nullptr, nullptr, as_stream(stream));
}

extern "C" int64_t ppo_adv_stats_workspace_bytes(int n_mb) {
    return n_mb < 0 ? -1 : (int64_t)ADV_TICKETS * 4 + (int64_t)n_mb * ADV_CSLICES_MAX * 2 * (int64_t)sizeof(double);
}

extern "C" int ppo_adv_stats(const float* adv, const int32_t* perm, int64_t batch, int mb_size, int n_mb,
                             float* stats_out, double* partial_out, void* workspace, ppo_stream_t stream) {
    PPO_REQUIRE(adv && mb_size >= 1 && n_mb >= 0 && batch >= 0, "ppo_adv_stats: bad arguments");
    PPO_REQUIRE(stats_out || partial_out, "ppo_adv_stats: no output requested");
    if (n_mb == 0) return PPO_OK;
    if (workspace && mb_size >= 4 * ADV_SLICES * ADV_THREADS && n_mb <= ADV_TICKETS)
        adv_moments_kernel<<<dim3(n_mb, ADV_SLICES), ADV_THREADS, 0, as_stream(stream)>>>(
            adv, perm, batch, mb_size, n_mb, reinterpret_cast<double*>(workspace), stats_out, partial_out);
    else
        adv_stats_kernel<<<n_mb, STAT_THREADS, 0, as_stream(stream)>>>(adv, perm, batch, mb_size, stats_out, partial_out);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

extern "C" int ppo_adv_stats_finalize(const double* partial, int n_mb, float* stats_out, ppo_stream_t stream) {
    PPO_REQUIRE(partial && stats_out && n_mb >= 0, "ppo_adv_stats_finalize: bad arguments");
    if (n_mb == 0) return PPO_OK;
    adv_stats_finalize_kernel<<<(n_mb + 63) / 64, 64, 0, as_stream(stream)>>>(partial, n_mb, stats_out);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

extern "C" int64_t ppo_explained_variance_workspace_bytes(void) {
    return (int64_t)EV_MAX_BLOCKS * 4 * sizeof(double) + 16;
}

extern "C" int ppo_explained_variance(int64_t n, const float* returns, const float* values, void* workspace,
                                      float* out, ppo_stream_t stream) {
    PPO_REQUIRE(n >= 2 && returns && values && workspace && out, "ppo_explained_variance: bad arguments");
    const int sms = sm_count() > 0 ? sm_count() : 148;
    int64_t blocks = (n + EV_THREADS * 8 - 1) / (EV_THREADS * 8);
    if (blocks > 4 * sms) blocks = 4 * sms;
    if (blocks > EV_MAX_BLOCKS) blocks = EV_MAX_BLOCKS;
    double* part = reinterpret_cast<double*>(workspace);
    unsigned int* ticket = reinterpret_cast<unsigned int*>(part + (size_t)EV_MAX_BLOCKS * 4);
    explained_variance_kernel<<<(int)blocks, EV_THREADS, 0, as_stream(stream)>>>(n, returns, values, part, ticket, out);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

// ------------------------------------------------------------------------------------------------
// ppo_explained_variance_static: the same kernel with its scratch (per-CTA partial moments + the arrival ticket) in
// a file-scope device array instead of a caller-allocated workspace, for callers that may not allocate (stream
// capture, serving loops).  No reference counterpart.
// Every GPU that loads the library holds these 32 784 bytes until exit.  ALL calls on a device share the array:
// calls that overlap on the GPU (different streams) race on it.
constexpr size_t EV_STATIC_BYTES = (size_t)EV_MAX_BLOCKS * 4 * sizeof(double) + 16;
static __device__ __align__(256) unsigned char g_ev_static_scratch[EV_STATIC_BYTES];
constexpr int EV_STATIC_MAX_DEVICES = 64;

// In host code `g_ev_static_scratch` is only a stand-in for the symbol: its device address comes from
// cudaGetSymbolAddress, resolved once per device with that device current.
static int ev_static_scratch_ptr(void** out) {
    static std::mutex mu;
    static void* cached[EV_STATIC_MAX_DEVICES] = {};
    int dev = -1;
    PPO_CUDA(cudaGetDevice(&dev));
    PPO_REQUIRE(dev >= 0 && dev < EV_STATIC_MAX_DEVICES, "ppo_explained_variance_static: device ordinal %d out of [0,%d)",
                dev, EV_STATIC_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[dev]) {
        void* p = nullptr;
        PPO_CUDA(cudaGetSymbolAddress(&p, g_ev_static_scratch));
        PPO_REQUIRE(p != nullptr, "ppo_explained_variance_static: cudaGetSymbolAddress returned NULL");
        cached[dev] = p;
    }
    *out = cached[dev];
    return PPO_OK;
}

extern "C" int64_t ppo_explained_variance_static_max_n(void) { return PPO_EV_STATIC_MAX_N; }

extern "C" int ppo_explained_variance_static(int64_t n, const float* returns, const float* values, float* out,
                                             ppo_stream_t stream) {
    PPO_REQUIRE(n >= 2 && returns && values && out, "ppo_explained_variance_static: bad arguments");
    PPO_REQUIRE(n <= PPO_EV_STATIC_MAX_N,
                "ppo_explained_variance_static: n=%lld exceeds the static-scratch variant's maximum of %lld rows "
                "(use ppo_explained_variance with a caller-allocated workspace)",
                (long long)n, (long long)PPO_EV_STATIC_MAX_N);
    void* scratch = nullptr;
    const int rc = ev_static_scratch_ptr(&scratch);
    if (rc != PPO_OK) return rc;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    int64_t blocks = (n + EV_THREADS * 8 - 1) / (EV_THREADS * 8);
    if (blocks > 4 * sms) blocks = 4 * sms;
    if (blocks > EV_MAX_BLOCKS) blocks = EV_MAX_BLOCKS;
    double* part = reinterpret_cast<double*>(scratch);
    unsigned int* ticket = reinterpret_cast<unsigned int*>(part + (size_t)EV_MAX_BLOCKS * 4);
    // the array keeps the last call's contents: the partials are overwritten before they are read, the ticket must
    // start at zero (the kernel re-arms it, but a call that was aborted mid-flight would not have)
    PPO_CUDA(cudaMemsetAsync(ticket, 0, 16, as_stream(stream)));
    explained_variance_kernel<<<(int)blocks, EV_THREADS, 0, as_stream(stream)>>>(n, returns, values, part, ticket, out);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

extern "C" int ppo_make_permutation(int64_t batch, uint64_t seed, uint64_t epoch, const uint64_t* d_epoch,
                                    int32_t* perm_out, ppo_stream_t stream) {
    PPO_REQUIRE(batch >= 0 && batch <= 0x7fffffffLL && (perm_out || batch == 0), "ppo_make_permutation: bad arguments");
    if (batch == 0) return PPO_OK;
    const int64_t blocks = (batch + 255) / 256;
    permutation_kernel<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(batch, perm_half_bits(batch), seed, epoch, d_epoch,
                                                                        perm_out);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

// The newest max_n finished episodes of an episode log (ppo_episode_tail): one CTA walks the [T,n] log backwards, 4096
// rows per round (four independent loads per thread in flight; thread 0 / slot 0 = newest row), ranks the hits with ballot
// prefixes and stops as soon as max_n are found - one round whenever episodes are shorter than ~40 steps on average.
__global__ void __launch_bounds__(1024) episode_tail_kernel(const int32_t* __restrict__ fin_len,
                                                            const float* __restrict__ fin_ret, int64_t rows, int max_n,
                                                            int32_t* __restrict__ out_count, int32_t* __restrict__ out_len,
                                                            float* __restrict__ out_ret) {
    constexpr int U = 4;
    __shared__ int32_t s_len[1024];
    __shared__ float s_ret[1024];
    __shared__ int s_warp[U * 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int found = 0;   // CTA-uniform
    for (int64_t hi = rows; hi > 0 && found < max_n; hi -= U * 1024) {
        int32_t l[U];
        float v[U];
        bool any = false;
#pragma unroll
        for (int k = 0; k < U; k++) {
            const int64_t r = hi - 1 - (k * 1024 + threadIdx.x);
            l[k] = r >= 0 ? __ldg(fin_len + r) : 0;
        }
#pragma unroll
        for (int k = 0; k < U; k++) {
            const int64_t r = hi - 1 - (k * 1024 + threadIdx.x);
            v[k] = l[k] ? __ldg(fin_ret + r) : 0.0f;
            any |= l[k] != 0;
        }
        if (!__syncthreads_or(any)) continue;
        unsigned b[U];
#pragma unroll
        for (int k = 0; k < U; k++) {
            b[k] = __ballot_sync(0xffffffffu, l[k] != 0);
            if (lane == 0) s_warp[k * 32 + w] = __popc(b[k]);
        }
        __syncthreads();
        int base = found;   // rank of (slot k, warp w, lane 0); slots are ordered newest first
#pragma unroll
        for (int k = 0; k < U; k++) {
            int off = 0, total = 0;
            for (int i = 0; i < 32; i++) {
                const int c = s_warp[k * 32 + i];
                if (i < w) off += c;
                total += c;
            }
            const int pos = base + off + __popc(b[k] & ((1u << lane) - 1u));
            if (l[k] != 0 && pos < max_n) { s_len[pos] = l[k]; s_ret[pos] = v[k]; }
            base += total;
        }
        found = base;
        __syncthreads();
    }
    const int cnt = found < max_n ? found : max_n;
    for (int i = threadIdx.x; i < cnt; i += 1024) {   // oldest first
        out_len[i] = s_len[cnt - 1 - i];
        out_ret[i] = s_ret[cnt - 1 - i];
    }
    if (threadIdx.x == 0) *out_count = cnt;
}

extern "C" int ppo_episode_tail(const int32_t* fin_len, const float* fin_ret, int64_t rows, int max_n, int32_t* out_count,
                                int32_t* out_len, float* out_ret, ppo_stream_t stream) {
    PPO_REQUIRE(rows >= 0 && max_n >= 1 && max_n <= 1024, "ppo_episode_tail: rows=%lld max_n=%d (1..1024)",
                (long long)rows, max_n);
    PPO_REQUIRE(out_count && out_len && out_ret && (rows == 0 || (fin_len && fin_ret)), "ppo_episode_tail: null buffer");
    episode_tail_kernel<<<1, 1024, 0, as_stream(stream)>>>(fin_len, fin_ret, rows, max_n, out_count, out_len, out_ret);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

// One 64-byte record per rollout row for the update kernel's gather (see ppo_pack_rows in the header)
__global__ void pack_rows_kernel(int64_t batch, int O, const float* __restrict__ obs, const int32_t* __restrict__ actions,
                                 const float* __restrict__ logp, const float* __restrict__ adv,
                                 const float* __restrict__ ret, const float* __restrict__ val, float4* __restrict__ rec) {
    // thread = (row, quarter): the four float4 of a record are written by four adjacent lanes (coalesced 64 B)
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = i >> 2;
    const int q = (int)(i & 3);
    if (r >= batch) return;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (q == 0) {
        if (O == 4) v = __ldg(reinterpret_cast<const float4*>(obs) + r);
        else {
            float o[4] = {0.f, 0.f, 0.f, 0.f};
            for (int k = 0; k < O; k++) o[k] = __ldg(obs + r * O + k);
            v = make_float4(o[0], o[1], o[2], o[3]);
        }
    } else if (q == 1) {
        v = make_float4(__ldg(logp + r), __ldg(adv + r), __ldg(ret + r), val ? __ldg(val + r) : 0.0f);
    } else if (q == 2) {
        v.x = __int_as_float(__ldg(actions + r));
    }
    rec[i] = v;
}

extern "C" int ppo_pack_rows(int64_t batch, int obs_dim, const float* obs, const int32_t* actions, const float* old_logprob,
                             const float* adv, const float* ret, const float* old_val, float* rec_out,
                             ppo_stream_t stream) {
    PPO_REQUIRE(batch >= 0 && obs_dim >= 1 && obs_dim <= 4, "ppo_pack_rows: batch=%lld obs_dim=%d (1..4)", (long long)batch,
                obs_dim);
    if (batch == 0) return PPO_OK;
    PPO_REQUIRE(obs && actions && old_logprob && adv && ret && rec_out, "ppo_pack_rows: null buffer");
    PPO_REQUIRE((reinterpret_cast<uintptr_t>(rec_out) & 63) == 0, "ppo_pack_rows: rec_out must be 64-byte aligned");
    const int64_t n = batch * 4;
    pack_rows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, as_stream(stream)>>>(
        batch, obs_dim, obs, actions, old_logprob, adv, ret, old_val, reinterpret_cast<float4*>(rec_out));
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

extern "C" int ppo_adv_stats_all_epochs(const float* adv, int64_t batch, int mb_size, int n_mb, int n_epochs, uint64_t seed,
                                        uint64_t epoch0, const uint64_t* d_epoch, float* stats_out, double* partial_out,
                                        void* workspace, ppo_stream_t stream) {
    PPO_REQUIRE(adv && workspace && mb_size >= 1 && n_mb >= 0 && n_epochs >= 0 && batch >= 0 && batch <= 0x7fffffffLL,
                "ppo_adv_stats_all_epochs: bad arguments");
    PPO_REQUIRE(stats_out || partial_out, "ppo_adv_stats_all_epochs: no output requested");
    if (n_mb == 0 || n_epochs == 0) return PPO_OK;
    PPO_REQUIRE((int64_t)n_mb * n_epochs <= ADV_TICKETS, "ppo_adv_stats_all_epochs: more than %d minibatches per update", ADV_TICKETS);
    // one coalesced pass per epoch once the advantages outgrow L2 (the permuted gathers turn into DRAM sectors); rows of
    // a partial minibatch (batch % n_mb != 0) are handled alike: position / mb_size may reach n_mb - 1 only
    // (below ~31 M rows `adv` stays L2-resident and the gathers are cheaper: 0.83 vs 1.05 ms at 16.8 M rows)
    if (n_mb <= ADV_MBMAX && n_epochs <= 1024 && (g_adv_impl == 2 || (g_adv_impl == 1 && batch >= (1 << 25)))) {
        double* ws = reinterpret_cast<double*>(workspace);
        // slices per epoch: 5 CTAs per SM over the whole grid (10 epochs x 32 slices = 320 CTAs left 24 of 148 SMs with three
        // CTAs and the rest with two: the launch ran at the pace of the three)
        const int sms = sm_count() > 0 ? sm_count() : 148;
        int nsl = (5 * sms + n_epochs / 2) / n_epochs;
        nsl = nsl < 16 ? 16 : (nsl > ADV_CSLICES_MAX ? ADV_CSLICES_MAX : nsl);
        adv_moments_all_coalesced_kernel<<<dim3(n_epochs, nsl), ADV_THREADS, 0, as_stream(stream)>>>(
            adv, batch, mb_size, n_mb, n_epochs, perm_half_bits(batch), seed, epoch0, d_epoch, ws, stats_out, partial_out);
        PPO_LAUNCH_CHECK();
        return PPO_OK;
    }
    adv_moments_all_kernel<<<dim3(n_mb * n_epochs, ADV_SLICES), ADV_THREADS, 0, as_stream(stream)>>>(
        adv, batch, mb_size, n_mb, n_epochs, perm_half_bits(batch), seed, epoch0, d_epoch,
        reinterpret_cast<double*>(workspace), stats_out, partial_out);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}
