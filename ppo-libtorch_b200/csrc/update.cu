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
    uint32_t pkey[4] = {0u, 0u, 0u, 0u};
    if (!a.perm && a.pf.inline_on) perm_keys(a.pf.seed, a.pf.epoch + (a.pf.d_epoch ? *a.pf.d_epoch : 0ull), pkey);
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int r0 = tile * TM;
        const int rows = min(TM, a.mb_size - r0);
        __syncthreads();
        if (t < TM) {
            int64_t idx = -1;
            if (t < rows) {
                if (a.perm) idx = (int64_t)a.perm[a.mb_start + r0 + t];
                else if (a.pf.inline_on) idx = perm_index(a.mb_start + r0 + t, a.pf.batch, a.pf.half_bits, pkey);
                else idx = a.mb_start + r0 + t;
            }
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
// reduce_kernel + clip_grad_norm_ + AdamW in one launch.  Every CTA reduces its RED_COLS columns as above and leaves the
// sum of their squares in sq_part[block].  spin (grid <= SM count, so every CTA is resident): each CTA waits until all
// tickets are taken, adds the partials in block order (the norm is run-to-run identical and the same in every CTA) and
// applies clip + decoupled weight decay + Adam (optim.cu's arithmetic) to its own columns.  Otherwise the last CTA to
// take a ticket steps all P parameters.  The ticket pair (arrivals, departures) is re-armed before the kernel ends.
__global__ void __launch_bounds__(RED_COLS * RED_PARTS) reduce_apply_kernel(
    const float* __restrict__ slabs, const double* __restrict__ dpart, int nslabs, int P, int64_t stride, float inv_n,
    float ent_coef, float vf_coef, float* grads_out, float* __restrict__ diag_out, const FuseAdam f, unsigned* ticket,
    double* sq_part, const bool spin) {
    __shared__ double part[RED_PARTS][RED_COLS];
    __shared__ double shn[RED_COLS * RED_PARTS / 32];
    __shared__ unsigned s_last;
    if (threadIdx.x == 0) s_last = 0u;
    const int col = threadIdx.x & (RED_COLS - 1), pt = threadIdx.x / RED_COLS;
    const int i = blockIdx.x * RED_COLS + col;
    pdl_trigger();   // the next update kernel may be scheduled behind this grid; it waits for our completion itself
    pdl_wait();      // the update kernel's slabs are complete and visible
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
    double sq = 0.0;
    if (pt == 0 && i < P) {
        double t = part[0][col];
#pragma unroll
        for (int q = 1; q < RED_PARTS; q++) t += part[q][col];
        const float g = (float)t;
        grads_out[i] = g;
        sq = (double)g * (double)g;
    }
    if (threadIdx.x < RED_COLS) {   // warps 0, 1 hold the columns: fixed xor tree, then warp 0 + warp 1
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
        if ((threadIdx.x & 31) == 0) shn[threadIdx.x >> 5] = sq;
    }
    if (blockIdx.x == 0 && threadIdx.x >= 64 && threadIdx.x < 96) {   // diagnostics (warp 2 of block 0)
        const int l = threadIdx.x - 64;
        double v = 0.0;
        if (l < 5)
            for (int s = 0; s < nslabs; s++) v += dpart[(size_t)s * UPD_NDIAG + l];
        const double n_inv = (double)inv_n;
        const float pg = (float)(__shfl_sync(0xffffffffu, v, 0) * n_inv);
        const float vl = 0.5f * (float)(__shfl_sync(0xffffffffu, v, 1) * n_inv);
        const float en = (float)(__shfl_sync(0xffffffffu, v, 2) * n_inv);
        const float kl = (float)(__shfl_sync(0xffffffffu, v, 3) * n_inv);
        const float cf = (float)(__shfl_sync(0xffffffffu, v, 4) * n_inv);
        if (l < 8) {
            const float d8[8] = {pg, vl, en, pg - ent_coef * en + vl * vf_coef, kl, cf, 0.0f, 0.0f};
            float dv = 0.0f;
#pragma unroll
            for (int q = 0; q < 8; q++)
                if (q == l) dv = d8[q];
            if (diag_out) diag_out[l] = dv;
            if (f.diag_hist) f.diag_hist[l] = dv;
        }
    }
    __threadfence();   // this thread's grads_out / diagnostics stores are visible device-wide before the ticket is taken
    __syncthreads();
    int j0, j1;   // the parameter range this CTA steps
    if (spin) {
        // every CTA is resident (grid <= SM count, checked by the launcher): publish, wait for all, step the own columns
        if (threadIdx.x == 0) {
            sq_part[blockIdx.x] = shn[0] + shn[1];
            __threadfence();
            atomicAdd(ticket, 1u);
            const long long t0 = clock64();   // bounded: a CTA that never sees the others gives up (no GPU hang) ...
            const long long lim = f.spin_cycles > 0 ? f.spin_cycles : 40000000000ll;
            while (*reinterpret_cast<volatile unsigned*>(ticket) < gridDim.x) {
                if (clock64() - t0 > lim) { if (f.err) atomicExch(f.err, 1u); s_last = 2u; break; }
                __nanosleep(32);
            }
            __threadfence();
        }
        __syncthreads();
        // ... and does NOT step: a partial norm would clip with the wrong coefficient (the host sees PPO_ERR_COMM)
        if (s_last == 2u || (f.err && *reinterpret_cast<volatile unsigned*>(f.err))) return;
        j0 = blockIdx.x * RED_COLS;
        j1 = min(P, j0 + RED_COLS);
    } else {
        if (threadIdx.x == 0) {
            sq_part[blockIdx.x] = shn[0] + shn[1];
            __threadfence();
            s_last = atomicAdd(ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
        }
        __syncthreads();
        if (!s_last) return;
        __threadfence();
        j0 = 0;
        j1 = P;
    }
    // ---- norm (partials added in block order by every stepping CTA alike), clip coefficient, AdamW
    double ss = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += RED_COLS * RED_PARTS) ss += __ldcg(sq_part + b);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if ((threadIdx.x & 31) == 0) shn[threadIdx.x >> 5] = ss;
    __syncthreads();
    double tot = 0.0;
#pragma unroll
    for (int w = 0; w < RED_COLS * RED_PARTS / 32; w++) tot += shn[w];
    const double norm = sqrt(tot);
    if (threadIdx.x == 0 && blockIdx.x == (spin ? 0 : blockIdx.x) && f.norm_out) *f.norm_out = (float)norm;
    const double coef = (double)f.max_grad_norm / (norm + 1e-6);
    const float cf = (f.max_grad_norm > 0.0f && coef < 1.0) ? (float)coef : 1.0f;
    const float lr = *f.d_lr;
    const int64_t step = f.step_host + (f.d_step ? *f.d_step : 0);
    const double b1 = (double)f.beta1, b2 = (double)f.beta2;
    const double bc1 = 1.0 - pow(b1, (double)step);
    const double bc2 = 1.0 - pow(b2, (double)step);
    const float decay = (float)(1.0 - (double)lr * (double)f.weight_decay);
    const float omb1 = (float)(1.0 - b1), omb2 = (float)(1.0 - b2);
    const float sq_bc2 = (float)sqrt(bc2);
    const float neg_step = (float)(-((double)lr / bc1));
    constexpr int NT = RED_COLS * RED_PARTS, U = 4;
    for (int jb = j0 + threadIdx.x; jb < j1; jb += NT * U) {   // U independent elements per thread: loads in flight together
        float g[U], pv[U], mv[U], vv[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int j = jb + u * NT;
            if (j < j1) { g[u] = __ldcg(grads_out + j); pv[u] = f.params[j]; mv[u] = f.exp_avg[j]; vv[u] = f.exp_avg_sq[j]; }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int j = jb + u * NT;
            if (j >= j1) continue;
            float gg = g[u];
            if (cf != 1.0f) gg = __fmul_rn(gg, cf);
            const float p = __fmul_rn(pv[u], decay);
            const float mi = __fadd_rn(__fmul_rn(mv[u], f.beta1), __fmul_rn(gg, omb1));
            const float vi = __fadd_rn(__fmul_rn(vv[u], f.beta2), __fmul_rn(__fmul_rn(gg, gg), omb2));
            f.exp_avg[j] = mi;
            f.exp_avg_sq[j] = vi;
            const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(vi), sq_bc2), f.eps);
            f.params[j] = __fadd_rn(p, __fmul_rn(neg_step, __fdiv_rn(mi, denom)));
        }
    }
    // re-arm the ticket for the next launch
    if (spin) {
        __syncthreads();
        if (threadIdx.x == 0 && atomicAdd(ticket + 1, 1u) == gridDim.x - 1) { ticket[0] = 0u; ticket[1] = 0u; }
    } else if (threadIdx.x == 0) {
        *ticket = 0u;
    }
}
// ---- data-parallel tail over peer memory (see PeerArgs in update.cuh) ------------------------------------------------
__device__ __forceinline__ unsigned ld_sys_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_sys_f32(const float* p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_sys_u32(unsigned* p, unsigned v) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_sys_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_sys_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__global__ void __launch_bounds__(RED_COLS * RED_PARTS) reduce_peer_apply_kernel(
    const float* __restrict__ slabs, const double* __restrict__ dpart, int nslabs, int P, int64_t stride, float inv_n,
    float ent_coef, float vf_coef, float* grads_out, float* __restrict__ diag_out, const FuseAdam f,
    const __grid_constant__ PeerArgs pa,   // indexed with runtime ranks: read from the constant bank, not a local copy
    unsigned* ticket, double* sq_part) {
    __shared__ double part[RED_PARTS][RED_COLS];
    __shared__ double shn[RED_COLS * RED_PARTS / 32];
    __shared__ unsigned s_fail;
    const int col = threadIdx.x & (RED_COLS - 1), pt = threadIdx.x / RED_COLS;
    const int i = blockIdx.x * RED_COLS + col;
    const int64_t step = f.step_host + (f.d_step ? *f.d_step : 0);
    const unsigned seq = (unsigned)step;
    const int64_t px = peer_px(P);
    const int W = pa.world;
    const int64_t slot_off = (int64_t)(seq & 1u) * peer_slot_bytes(P, W);
    if (threadIdx.x == 0) s_fail = 0u;
    pdl_trigger();
    pdl_wait();
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
    // ---- push this rank's column block (and, from block 0, its 8 diagnostics) into EVERY rank's buffer, row [rank].
    // Every value travels as ONE 8-byte word {step number, fp32 bits} (the "LL" protocol): an aligned 8-byte store is
    // single-copy atomic, so the receiver polls the word itself until its tag is this step's - no flag array, no
    // system-scope fence between data and flag (a fence behind remote stores waits for their acknowledgement, i.e. a
    // full NVLink round trip: the first push version with separate flags cost the same ~11 us per step as the pull one).
    const uint64_t tag = (uint64_t)seq << 32;
    if (pt == 0 && i < P) {
        double t = part[0][col];
#pragma unroll
        for (int q = 1; q < RED_PARTS; q++) t += part[q][col];
        const uint64_t word = tag | (uint64_t)__float_as_uint((float)t);
#pragma unroll
        for (int r = 0; r < PEER_MAX; r++)
            if (r < W) st_sys_u64(reinterpret_cast<uint64_t*>(pa.xchg[r] + slot_off) + (int64_t)pa.rank * px + i, word);
    }
    if (blockIdx.x == 0 && threadIdx.x >= 64 && threadIdx.x < 96) {
        const int l = threadIdx.x - 64;
        double v = 0.0;
        if (l < 5)
            for (int s = 0; s < nslabs; s++) v += dpart[(size_t)s * UPD_NDIAG + l];
        const double n_inv = (double)inv_n;
        const float pg = (float)(__shfl_sync(0xffffffffu, v, 0) * n_inv);
        const float vl = 0.5f * (float)(__shfl_sync(0xffffffffu, v, 1) * n_inv);
        const float en = (float)(__shfl_sync(0xffffffffu, v, 2) * n_inv);
        const float kl = (float)(__shfl_sync(0xffffffffu, v, 3) * n_inv);
        const float cf = (float)(__shfl_sync(0xffffffffu, v, 4) * n_inv);
        if (l < 8) {
            const float d8[8] = {pg, vl, en, pg - ent_coef * en + vl * vf_coef, kl, cf, 0.0f, 0.0f};
            float dv = 0.0f;
#pragma unroll
            for (int q = 0; q < 8; q++)
                if (q == l) dv = d8[q];
            const uint64_t word = tag | (uint64_t)__float_as_uint(dv);
#pragma unroll
            for (int r = 0; r < PEER_MAX; r++)
                if (r < W)
                    st_sys_u64(reinterpret_cast<uint64_t*>(pa.xchg[r] + slot_off) + (int64_t)pa.rank * px + px - 64 + l, word);
        }
    }
    // ---- receive, in LOCAL memory: every thread polls the world's words of its own column (bounded spin: a missing peer
    // sets *err and the step is skipped, no GPU hang), then sums them in rank order
    const uint64_t* mine = reinterpret_cast<const uint64_t*>(pa.xchg[pa.rank] + slot_off);
    const long long spin_lim = pa.spin_cycles > 0 ? pa.spin_cycles : 40000000000ll;
    auto receive = [&](int64_t idx) -> float {
        uint64_t w[PEER_MAX];
        const long long t0 = clock64();
        for (;;) {
            bool all = true;
#pragma unroll
            for (int r = 0; r < PEER_MAX; r++) {
                w[r] = r < W ? ld_sys_u64(mine + (int64_t)r * px + idx) : tag;   // the world's loads in flight together
            }
#pragma unroll
            for (int r = 0; r < PEER_MAX; r++) all = all && (uint32_t)(w[r] >> 32) == seq;
            if (all) break;
            if (clock64() - t0 > spin_lim) { atomicExch(pa.err, 1u); s_fail = 1u; break; }
        }
        float g = 0.0f;
#pragma unroll
        for (int r = 0; r < PEER_MAX; r++)
            if (r < W) g += __uint_as_float((uint32_t)w[r]);
        return g;
    };
    double sq = 0.0;
    if (pt == 0 && i < P) {
        const float g = receive(i);
        grads_out[i] = g;
        sq = (double)g * (double)g;
    }
    if (blockIdx.x == 0 && threadIdx.x >= 64 && threadIdx.x < 72) {
        const int l = threadIdx.x - 64;
        const float dv = receive(px - 64 + l);
        if (diag_out) diag_out[l] = dv;
        if (f.diag_hist) f.diag_hist[l] = dv;
    }
    if (threadIdx.x < RED_COLS) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
        if ((threadIdx.x & 31) == 0) shn[threadIdx.x >> 5] = sq;
    }
    __threadfence();
    __syncthreads();
    // ---- intra-GPU: every CTA is resident (launcher checks): publish the partial norm, wait for all, step own columns
    if (threadIdx.x == 0) {
        sq_part[blockIdx.x] = shn[0] + shn[1];
        __threadfence();
        atomicAdd(ticket, 1u);
        const long long t1 = clock64();
        const long long lim = pa.spin_cycles > 0 ? pa.spin_cycles : 40000000000ll;
        while (*reinterpret_cast<volatile unsigned*>(ticket) < gridDim.x) {
            if (clock64() - t1 > lim) { atomicExch(pa.err, 1u); s_fail = 1u; break; }
            __nanosleep(32);
        }
        __threadfence();
    }
    __syncthreads();
    // a rank that missed a peer (this CTA or any other: *err is device-wide and sticky) must not step on a partial sum
    const bool failed = s_fail != 0u || *reinterpret_cast<volatile unsigned*>(pa.err) != 0u;
    double ss = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += RED_COLS * RED_PARTS) ss += __ldcg(sq_part + b);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if ((threadIdx.x & 31) == 0) shn[threadIdx.x >> 5] = ss;
    __syncthreads();
    double tot = 0.0;
#pragma unroll
    for (int w = 0; w < RED_COLS * RED_PARTS / 32; w++) tot += shn[w];
    const double norm = sqrt(tot);
    if (threadIdx.x == 0 && blockIdx.x == 0 && f.norm_out) *f.norm_out = (float)norm;
    const double coef = (double)f.max_grad_norm / (norm + 1e-6);
    const float cf = (f.max_grad_norm > 0.0f && coef < 1.0) ? (float)coef : 1.0f;
    const float lr = *f.d_lr;
    const double b1 = (double)f.beta1, b2 = (double)f.beta2;
    const double bc1 = 1.0 - pow(b1, (double)step);
    const double bc2 = 1.0 - pow(b2, (double)step);
    const float decay = (float)(1.0 - (double)lr * (double)f.weight_decay);
    const float omb1 = (float)(1.0 - b1), omb2 = (float)(1.0 - b2);
    const float sq_bc2 = (float)sqrt(bc2);
    const float neg_step = (float)(-((double)lr / bc1));
    if (pt == 0 && i < P && !failed) {
        float gg = grads_out[i];
        if (cf != 1.0f) gg = __fmul_rn(gg, cf);
        const float p = __fmul_rn(f.params[i], decay);
        const float mi = __fadd_rn(__fmul_rn(f.exp_avg[i], f.beta1), __fmul_rn(gg, omb1));
        const float vi = __fadd_rn(__fmul_rn(f.exp_avg_sq[i], f.beta2), __fmul_rn(__fmul_rn(gg, gg), omb2));
        f.exp_avg[i] = mi;
        f.exp_avg_sq[i] = vi;
        const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(vi), sq_bc2), f.eps);
        f.params[i] = __fadd_rn(p, __fmul_rn(neg_step, __fdiv_rn(mi, denom)));
    }
    __syncthreads();
    if (threadIdx.x == 0 && atomicAdd(ticket + 1, 1u) == gridDim.x - 1) { ticket[0] = 0u; ticket[1] = 0u; }
}
static inline void launch_reduce(const UpdArgs& a, int grid, int P, float* grads_out, float* diag_out, cudaStream_t st,
                                 const FuseAdam* fuse = nullptr, const PeerArgs* peer = nullptr) {
    const int blocks = (P + RED_COLS - 1) / RED_COLS;
    if (fuse && peer) {
        char* tail = reinterpret_cast<char*>(a.dpart) + (size_t)UPD_MAX_GRID * UPD_NDIAG * sizeof(double);
        launch_pdl((g_pdl & 2) != 0, reduce_peer_apply_kernel, dim3(blocks), dim3(RED_COLS * RED_PARTS), 0, st, (const float*)a.slabs,
                   (const double*)a.dpart, grid, P, slab_stride(P), a.inv_n, a.hp.ent_coef, a.hp.vf_coef, grads_out, diag_out,
                   *fuse, *peer, reinterpret_cast<unsigned*>(tail), reinterpret_cast<double*>(tail + 64));
        return;
    }
    if (fuse) {
        char* tail = reinterpret_cast<char*>(a.dpart) + (size_t)UPD_MAX_GRID * UPD_NDIAG * sizeof(double);
        launch_pdl((g_pdl & 1) != 0, reduce_apply_kernel, dim3(blocks), dim3(RED_COLS * RED_PARTS), 0, st, (const float*)a.slabs,
                   (const double*)a.dpart, grid, P, slab_stride(P), a.inv_n, a.hp.ent_coef, a.hp.vf_coef, grads_out, diag_out,
                   *fuse, reinterpret_cast<unsigned*>(tail), reinterpret_cast<double*>(tail + 64),
                   blocks <= (sm_count() > 0 ? sm_count() : 0));
        return;
    }
    reduce_kernel<<<blocks, RED_COLS * RED_PARTS, 0, st>>>(
        a.slabs, a.dpart, grid, P, slab_stride(P), a.inv_n, a.hp.ent_coef, a.hp.vf_coef, grads_out, diag_out);
}

template <int TM, int H>
static int launch_update(const UpdArgs& a, float* grads_out, float* diag_out, cudaStream_t st, const FuseAdam* fuse,
                         const PeerArgs* peer) {
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
    launch_reduce(a, grid, P, grads_out, diag_out, st, fuse, peer);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

}  // namespace ppo

using namespace ppo;

extern "C" int64_t ppo_update_workspace_bytes(const ppo_dims* dims) {
    Dims d;
    if (make_dims(dims, &d)) return -1;
    const int P = num_params(d.O, d.H, d.K);
    return (int64_t)UPD_MAX_GRID * (slab_stride(P) * (int64_t)sizeof(float) + UPD_NDIAG * (int64_t)sizeof(double)) +
           upd_tail_bytes(P);
}

// which kernels serve ppo_update_minibatch for this shape in the current math mode (bench.py labels its roofline with it)
extern "C" int ppo_update_path(const ppo_dims* dims) {
    Dims d;
    if (make_dims(dims, &d)) return -1;
    if (math_mode() == PPO_MATH_FP32) return 0;
    if (update_tc_supported(d)) return 1;
    if (update_h256_supported(d)) return 2;
    return 0;
}

extern "C" int ppo_update_minibatch(const float* params, const ppo_dims* dims, const float* obs, const uint8_t* mask,
                                    const int32_t* actions, const float* old_logprob, const float* adv,
                                    const float* ret, const float* old_val, const int32_t* perm, int64_t mb_start,
                                    int mb_size, float inv_n, const float* adv_stats, const ppo_hyper* h_hyper,
                                    void* workspace, float* grads_out, float* diag_out, ppo_stream_t stream) {
    return update_minibatch_fused(params, dims, obs, mask, actions, old_logprob, adv, ret, old_val, perm, mb_start,
                                  mb_size, inv_n, adv_stats, h_hyper, workspace, grads_out, diag_out, nullptr,
                                  as_stream(stream));
}

int ppo::update_minibatch_fused(const float* params, const ppo_dims* dims, const float* obs, const uint8_t* mask,
                                const int32_t* actions, const float* old_logprob, const float* adv, const float* ret,
                                const float* old_val, const int32_t* perm, int64_t mb_start, int mb_size, float inv_n,
                                const float* adv_stats, const ppo_hyper* h_hyper, void* workspace, float* grads_out,
                                float* diag_out, const FuseAdam* fuse, cudaStream_t st, const PeerArgs* peer,
                                const PermFn* pf, const float* rec) {
    UpdArgs a{};
    if (pf) a.pf = *pf;
    a.rec = reinterpret_cast<const float4*>(rec);
    // programmatic dependent launch of the chain: bit 0 of ppo_debug_pdl on one GPU, bit 1 next to the peer exchange (with
    // the PULL exchange early launches cost 7 % on 2 GPUs, 4.03 vs 3.77 ms; the push exchange polls local memory only)
    a.pdl = peer ? ((g_pdl & 2) != 0) : ((g_pdl & 1) != 0);
    if (peer) {
        Dims dd;
        if (int rc = make_dims(dims, &dd)) return rc;
        PPO_REQUIRE(fuse && (num_params(dd.O, dd.H, dd.K) + RED_COLS - 1) / RED_COLS <= min(PEER_NB, sm_count()),
                    "peer-memory reduction needs the fused optimiser and a grid no larger than the SM count");
    }
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
    if (a.rec) PPO_REQUIRE(a.d.heads == 1 && !mask && a.d.O <= 4, "packed rows: single head, no mask, obs_dim <= 4");
    const int P = num_params(a.d.O, a.d.H, a.d.K);
    a.slabs = reinterpret_cast<float*>(workspace);
    a.dpart = reinterpret_cast<double*>(reinterpret_cast<char*>(workspace) +
                                        (size_t)UPD_MAX_GRID * slab_stride(P) * sizeof(float));
    const int mode = math_mode();
    if (mode != PPO_MATH_FP32) {  // tensor-core path when the shape fits its envelope, else the FMA kernels below
        int grid = 0;
        bool tail_done = false;   // single GPU, enough CTAs: the update kernel reduces and steps by itself
        const int rc = launch_update_tc(a, mode == PPO_MATH_BF16 ? 1 : 0, &grid, st, peer ? nullptr : fuse, grads_out,
                                        diag_out, &tail_done);
        if (rc == PPO_OK) {
            if (!tail_done) launch_reduce(a, grid, P, grads_out, diag_out, st, fuse, peer);
            PPO_LAUNCH_CHECK();
            return PPO_OK;
        }
        if (rc != PPO_ERR_UNSUPPORTED) return rc;
        const int rc2 = launch_update_h256(a, mode == PPO_MATH_BF16 ? 1 : 0, &grid, st);
        if (rc2 == PPO_OK) {
            launch_reduce(a, grid, P, grads_out, diag_out, st, fuse, peer);
            PPO_LAUNCH_CHECK();
            return PPO_OK;
        }
        if (rc2 != PPO_ERR_UNSUPPORTED) return rc2;
    }
    switch (a.d.H) {
        case 64: return launch_update<64, 64>(a, grads_out, diag_out, st, fuse, peer);
        case 128: return launch_update<32, 128>(a, grads_out, diag_out, st, fuse, peer);
        case 256: return launch_update<32, 256>(a, grads_out, diag_out, st, fuse, peer);
    }
    set_error("hidden=%d unsupported", a.d.H);
    return PPO_ERR_UNSUPPORTED;
}
