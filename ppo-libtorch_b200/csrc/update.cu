// update.cu — K5: fused minibatch forward + PPO loss + closed-form backward (fp32 FMA path).
// Replaces PPO/PPO_Discrete.cpp:579-638 (forward with stored actions, ratio, clipped surrogate,
// clipped value loss, entropy bonus, autograd backward): ~100 LibTorch op dispatches + the autograd
// graph become two launches (tile kernel + deterministic cross-CTA reduction).
//
// Each persistent CTA walks tiles of TM minibatch rows gathered through the permutation:
//   actor : forward -> per-row Categorical(Masked) log-prob / entropy / ratio -> dL/dlogits -> backward
//   critic: forward -> per-row (clipped) value loss -> dL/dv -> backward
// (the two nets are disjoint graphs, so they reuse the same two activation buffers).
// Weight-gradient partials accumulate in a per-CTA slab of `workspace`; reduce_kernel sums the
// slabs in CTA order (no atomics: results are run-to-run identical) and finishes the diagnostics.
#include "policy.cuh"
#include "ppo_loss.cuh"
#include "update.cuh"

namespace ppo {

template <int TM, int H>
__device__ inline void net_backward(float* sm, const TileLayout& L, int O, const float* __restrict__ params,
                                    const NetOff& off, int K3, const float* dOut, int ldo, int rows,
                                    float* __restrict__ g) {
    // dOut rows >= `rows` are zero, so every dZ below is zero there too.
    layer3_bwd_dw<TM, H>(sm + L.a2, K3, dOut, ldo, rows, g + off.W3, g + off.b3);
    __syncthreads();
    layer3_bwd_dx<TM, H>(sm + L.a2, params + off.W3, K3, dOut, ldo);
    __syncthreads();
    layer2_bwd_dw<TM, H>(sm + L.a2, sm + L.a1, g + off.W2, g + off.b2);
    layer2_bwd_dx<TM, H>(sm + L.a2, params + off.W2, sm + L.a1, sm + L.ws);
    layer1_bwd_dw<TM, H>(sm + L.a1, sm + L.xs, O, g + off.W1, g + off.b1);
    __syncthreads();
}

template <int TM, int H>
__global__ void __launch_bounds__(MLP_THREADS) update_kernel(const UpdArgs a) {
    extern __shared__ __align__(16) float sm[];
    const int O = a.d.O, K = a.d.K, heads = a.d.heads;
    const TileLayout L = tile_layout(TM, H, O, K);
    int64_t* sidx = reinterpret_cast<int64_t*>(sm + ((L.end + 1) & ~1));  // [TM] gathered row index
    const NetOff co = critic_off(O, H), ao = actor_off(O, H, K);
    const int P = num_params(O, H, K);
    float* g = a.slabs + (size_t)blockIdx.x * slab_stride(P);
    for (int i = threadIdx.x; i < P; i += MLP_THREADS) g[i] = 0.0f;

    const bool masked = a.mask != nullptr;
    float meanf = 0.0f, stdf = 1.0f;
    if (a.hp.norm_adv) { meanf = a.adv_stats[0]; stdf = a.adv_stats[1]; }
    RowSums rs;

    const int ntiles = (a.mb_size + TM - 1) / TM;
    const int t = threadIdx.x;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int r0 = tile * TM;
        const int rows = min(TM, a.mb_size - r0);
        __syncthreads();
        if (t < TM) {
            int64_t idx = -1;
            if (t < rows) idx = a.perm ? (int64_t)a.perm[a.mb_start + r0 + t] : a.mb_start + r0 + t;
            sidx[t] = idx;
        }
        __syncthreads();
        for (int e = t; e < TM * O; e += MLP_THREADS) {
            const int m = e / O, i = e % O;
            sm[L.xs + e] = (m < rows) ? __ldg(a.obs + sidx[m] * O + i) : 0.0f;
        }
        __syncthreads();

        // ---------------- actor: forward, loss terms, dL/dlogits (in place over the logits) ----------
        net_forward<TM, H>(sm, L, O, a.params, ao, K, sm + L.ls, L.ldo);
        if (t < TM) {
            float* x = sm + L.ls + t * L.ldo;
            if (t < rows) {
                const int64_t r = sidx[t];
                const uint64_t mbits = masked ? pack_mask(a.mask + r * K, K) : ~0ull;
                actor_row_loss(x, a.d, masked, mbits, a.actions + r * heads, __ldg(a.old_logprob + r), __ldg(a.adv + r),
                               a.hp, meanf, stdf, a.inv_n, rs);
            } else {
                for (int j = 0; j < K; j++) x[j] = 0.0f;
            }
        }
        __syncthreads();
        net_backward<TM, H>(sm, L, O, a.params, ao, K, sm + L.ls, L.ldo, rows, g);

        // ---------------- critic: forward, value loss, dL/dv ----------------------------------------
        net_forward<TM, H>(sm, L, O, a.params, co, 1, sm + L.vs, 1);
        if (t < TM) {
            float gv = 0.0f;
            if (t < rows) {
                const int64_t r = sidx[t];
                gv = critic_row_loss(sm[L.vs + t], __ldg(a.ret + r), a.hp.clip_vloss ? __ldg(a.old_val + r) : 0.0f, a.hp,
                                     a.inv_n, rs);
            }
            sm[L.vs + t] = gv;
        }
        __syncthreads();
        net_backward<TM, H>(sm, L, O, a.params, co, 1, sm + L.vs, 1, rows, g);
    }

    // per-CTA diagnostic partial sums (only threads < TM hold non-zero values)
    __syncthreads();
    double* red = reinterpret_cast<double*>(sm);  // reuse: 5 x TM doubles
    if (t < TM) {
        red[0 * TM + t] = rs.pg; red[1 * TM + t] = rs.v; red[2 * TM + t] = rs.ent;
        red[3 * TM + t] = rs.kl; red[4 * TM + t] = rs.clip;
    }
    __syncthreads();
    if (t < 5) {
        double acc = 0.0;
        for (int m = 0; m < TM; m++) acc += red[t * TM + m];
        a.dpart[(size_t)blockIdx.x * UPD_NDIAG + t] = acc;
    }
}

// grads_out[i] = sum over CTAs (fixed order, double accumulation) of slab[cta][i]; block 0 also turns the diagnostic
// partials into diag_out = {pg_loss, v_loss, entropy, loss, approx_kl, clipfrac, 0, 0}.
// A block owns RED_COLS consecutive parameters; its 256 threads split the slabs RED_PARTS ways (coalesced 256-byte
// rows, independent loads in flight), then combine the RED_PARTS partial sums in fixed order.
constexpr int RED_COLS = 64, RED_PARTS = 4;
__global__ void __launch_bounds__(RED_COLS * RED_PARTS) reduce_kernel(
    const float* __restrict__ slabs, const double* __restrict__ dpart, int nslabs, int P, int64_t stride, float inv_n,
    float ent_coef, float vf_coef, float* __restrict__ grads_out, float* __restrict__ diag_out) {
    __shared__ double part[RED_PARTS][RED_COLS];
    const int col = threadIdx.x & (RED_COLS - 1), pt = threadIdx.x / RED_COLS;
    const int i = blockIdx.x * RED_COLS + col;
    double acc = 0.0;
    if (i < P) {
        const int per = (nslabs + RED_PARTS - 1) / RED_PARTS;
        const int s0 = pt * per, s1 = min(nslabs, s0 + per);
        const float* p = slabs + (size_t)s0 * stride + i;
        int s = s0;
        for (; s + 4 <= s1; s += 4, p += 4 * stride) {
            const float a0 = p[0], a1 = p[stride], a2 = p[2 * stride], a3 = p[3 * stride];
            acc += (double)a0; acc += (double)a1; acc += (double)a2; acc += (double)a3;
        }
        for (; s < s1; s++, p += stride) acc += (double)p[0];
    }
    part[pt][col] = acc;
    __syncthreads();
    if (pt == 0 && i < P) {
        double t = part[0][col];
#pragma unroll
        for (int q = 1; q < RED_PARTS; q++) t += part[q][col];
        grads_out[i] = (float)t;
    }
    if (blockIdx.x == 0 && threadIdx.x < 32 && diag_out) {
        double v = 0.0;
        if (threadIdx.x < 5)
            for (int s = 0; s < nslabs; s++) v += dpart[(size_t)s * UPD_NDIAG + threadIdx.x];
        const double n_inv = (double)inv_n;
        const float pg = (float)(__shfl_sync(0xffffffffu, v, 0) * n_inv);
        const float vl = 0.5f * (float)(__shfl_sync(0xffffffffu, v, 1) * n_inv);
        const float en = (float)(__shfl_sync(0xffffffffu, v, 2) * n_inv);
        const float kl = (float)(__shfl_sync(0xffffffffu, v, 3) * n_inv);
        const float cf = (float)(__shfl_sync(0xffffffffu, v, 4) * n_inv);
        if (threadIdx.x == 0) {
            diag_out[0] = pg; diag_out[1] = vl; diag_out[2] = en;
            diag_out[3] = pg - ent_coef * en + vl * vf_coef;  // :631
            diag_out[4] = kl; diag_out[5] = cf; diag_out[6] = 0.0f; diag_out[7] = 0.0f;
        }
    }
}
static inline void launch_reduce(const UpdArgs& a, int grid, int P, float* grads_out, float* diag_out, cudaStream_t st) {
    reduce_kernel<<<(P + RED_COLS - 1) / RED_COLS, RED_COLS * RED_PARTS, 0, st>>>(
        a.slabs, a.dpart, grid, P, slab_stride(P), a.inv_n, a.hp.ent_coef, a.hp.vf_coef, grads_out, diag_out);
}

template <int TM, int H>
static int launch_update(const UpdArgs& a, float* grads_out, float* diag_out, cudaStream_t st) {
    const TileLayout L = tile_layout(TM, H, a.d.O, a.d.K);
    const size_t smem = (size_t)(((L.end + 1) & ~1) + 2 * TM) * sizeof(float);
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(update_kernel<TM, H>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_done = true;
    }
    PPO_REQUIRE(smem <= 200 * 1024, "update: tile needs %zu B of shared memory", smem);
    const int ntiles = (a.mb_size + TM - 1) / TM;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    int grid = 2 * sms < UPD_MAX_GRID ? 2 * sms : UPD_MAX_GRID;
    if (ntiles < grid) grid = ntiles;
    update_kernel<TM, H><<<grid, MLP_THREADS, smem, st>>>(a);
    PPO_LAUNCH_CHECK();
    const int P = num_params(a.d.O, a.d.H, a.d.K);
    launch_reduce(a, grid, P, grads_out, diag_out, st);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

}  // namespace ppo

using namespace ppo;

extern "C" int64_t ppo_update_workspace_bytes(const ppo_dims* dims) {
    Dims d;
    if (make_dims(dims, &d)) return -1;
    const int P = num_params(d.O, d.H, d.K);
    return (int64_t)UPD_MAX_GRID * (slab_stride(P) * (int64_t)sizeof(float) + UPD_NDIAG * (int64_t)sizeof(double));
}

extern "C" int ppo_update_minibatch(const float* params, const ppo_dims* dims, const float* obs, const uint8_t* mask,
                                    const int32_t* actions, const float* old_logprob, const float* adv,
                                    const float* ret, const float* old_val, const int32_t* perm, int64_t mb_start,
                                    int mb_size, float inv_n, const float* adv_stats, const ppo_hyper* h_hyper,
                                    void* workspace, float* grads_out, float* diag_out, ppo_stream_t stream) {
    UpdArgs a{};
    if (int rc = make_dims(dims, &a.d)) return rc;
    PPO_REQUIRE(params && obs && actions && old_logprob && adv && ret && h_hyper && workspace && grads_out,
                "ppo_update_minibatch: null buffer");
    PPO_REQUIRE(mb_size >= 1 && mb_start >= 0, "ppo_update_minibatch: mb_start=%lld mb_size=%d",
                (long long)mb_start, mb_size);
    PPO_REQUIRE(!h_hyper->clip_vloss || old_val, "ppo_update_minibatch: clip_vloss needs old_val");
    PPO_REQUIRE(!h_hyper->norm_adv || adv_stats, "ppo_update_minibatch: norm_adv needs adv_stats");
    a.params = params; a.obs = obs; a.mask = mask; a.actions = actions; a.old_logprob = old_logprob;
    a.adv = adv; a.ret = ret; a.old_val = old_val; a.perm = perm; a.mb_start = mb_start; a.mb_size = mb_size;
    a.inv_n = inv_n; a.adv_stats = adv_stats; a.hp = *h_hyper;
    const int P = num_params(a.d.O, a.d.H, a.d.K);
    a.slabs = reinterpret_cast<float*>(workspace);
    a.dpart = reinterpret_cast<double*>(reinterpret_cast<char*>(workspace) +
                                        (size_t)UPD_MAX_GRID * slab_stride(P) * sizeof(float));
    cudaStream_t st = as_stream(stream);
    const int mode = math_mode();
    if (mode != PPO_MATH_FP32) {  // tensor-core path when the shape fits its envelope, else the FMA kernels below
        int grid = 0;
        const int rc = launch_update_tc(a, mode == PPO_MATH_BF16 ? 1 : 0, &grid, st);
        if (rc == PPO_OK) {
            launch_reduce(a, grid, P, grads_out, diag_out, st);
            PPO_LAUNCH_CHECK();
            return PPO_OK;
        }
        if (rc != PPO_ERR_UNSUPPORTED) return rc;
    }
    switch (a.d.H) {
        case 64: return launch_update<64, 64>(a, grads_out, diag_out, st);
        case 128: return launch_update<32, 128>(a, grads_out, diag_out, st);
        case 256: return launch_update<32, 256>(a, grads_out, diag_out, st);
    }
    set_error("hidden=%d unsupported", a.d.H);
    return PPO_ERR_UNSUPPORTED;
}
