// rollout_tc.cu — the fused rollout (rollout.cu) for H = 64 with BOTH hidden layers on the tensor core.
// Replaces the rollout phase of PPO_Discrete::train (PPO/PPO_Discrete.cpp:524-554) like rollout.cu does; this variant
// is the one used when there are enough envs to fill 128-row MMA tiles.
//
// One thread owns one env for all T steps (state, counters, done flag in registers); a warpgroup owns a 128-env tile
// = the M = 128 rows of every MMA:
//     G0  Z[net] = XE P1^T          (K = 16: {obs, 1, 0..}; layer-1 pre-activation incl. bias)
//     G1  Z[net] = A1 W2^T          (K = 64)
// The rollout has to reproduce the fp32 reference (sampled actions are compared index by index), so the 16-bit tensor
// core operands are SPLIT: x = hi + lo with hi = fp16(x), lo = fp16(x - hi), and a product is issued as
// hi*hi + lo*hi + hi*lo (fp32 accumulation in tensor memory).  That carries ~22 mantissa bits per operand, i.e. the
// result is as accurate as an fp32 FMA chain; the dropped lo*lo term is ~2^-22 relative.  Three times the MMAs is free
// here: the tensor pipe is otherwise idle and the FMA pipe was the measured limit of the FMA kernel (8192 FMAs per
// env-step for the two 64 x 64 layers against ~400 for everything else).
// What is left per env-step per thread: 256 tanh (ex2 + rcp on the SFU - the new bound), the 16-bit splits, the
// output layer (64 (K+1) FMAs against broadcast shared-memory weights), Categorical sampling and the env step.
#include "categorical.cuh"
#include "rollout.cuh"
#include "umma.cuh"

namespace ppo {
using namespace umma;

namespace {

constexpr int RT_NWG = 4;            // warpgroups (= 128-env tiles in flight) per CTA; one CTA per SM
constexpr int RT_ROWS = 128;
constexpr int RT_H = 64;
constexpr int RT_ACT = RT_ROWS * 128;   // one 16-bit activation tile
constexpr int RT_MAX_K = 8;
constexpr int FMT = FMT_F16;

struct RtLayout {    // byte offsets into the 1024-aligned dynamic shared memory
    int w2h[2], w2l[2];          // W2 hi / lo, 64 rows x 128 B, 128 B swizzle
    int p1h[2];                  // layer-1 weights {hi, hi, lo, 0} chunks (see update_tc.cu)
    int b2[2], w3, b3;           // fp32: b2 [64] per net, W3 [(K + 1)][64] (actor rows then the critic row), b3 [K + 1]
    int wg0, wg_stride;          // per-warpgroup block
    int a1h, a1l, xe, lg;        // inside it: A1 hi / lo tiles, XE {hi chunk, lo chunk}, logits rows
    int bars, tmem_slot, end, ldo;
};
__host__ __device__ inline RtLayout rt_layout(int K) {
    RtLayout L;
    L.ldo = K | 1;
    int o = 0;
    L.w2h[0] = o; o += 8192; L.w2h[1] = o; o += 8192;
    L.w2l[0] = o; o += 8192; L.w2l[1] = o; o += 8192;
    L.p1h[0] = o; o += 4 * RT_H * 16; L.p1h[1] = o; o += 4 * RT_H * 16;
    L.b2[0] = o; o += 256; L.b2[1] = o; o += 256;
    L.w3 = o; o += (K + 1) * 256;
    L.b3 = o; o += ((K + 1 + 3) & ~3) * 4;
    L.bars = o; o += RT_NWG * 2 * 8;
    L.tmem_slot = o; o += 16;
    o = (o + 1023) & ~1023;
    L.wg0 = o;
    L.a1h = 0; L.a1l = RT_ACT; L.xe = 2 * RT_ACT; L.lg = L.xe + 2 * RT_ROWS * 16;
    L.wg_stride = (L.lg + RT_ROWS * L.ldo * 4 + 1023) & ~1023;
    L.end = L.wg0 + RT_NWG * L.wg_stride;
    return L;
}

// tanh(x) = 1 - 2 / (exp(2x) + 1): absolute error ~1e-7 (same routine as rollout64_kernel)
__device__ __forceinline__ float tanh_acc(float x) {
    float t, r;   // bare ex2 / rcp: exp2f() wraps the SFU op in a denormal-range rescale that costs 3 more instructions
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(2.8853900817779268f * x));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(t + 1.0f));   // t = inf -> r = 0 -> +1;  t = 0 -> r = 1 -> -1
    return fmaf(-2.0f, r, 1.0f);
}
// 8 fp32 -> 16-byte chunk of fp16 (hi) and the chunk of the residuals (lo)
__device__ __forceinline__ void split8(const float* v, uint4& hi, uint4& lo) {
    uint32_t h[4], l[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        h[i] = pack2<FMT>(v[2 * i], v[2 * i + 1]);
        l[i] = pack2<FMT>(v[2 * i] - unpack_lo<FMT>(h[i]), v[2 * i + 1] - unpack_hi<FMT>(h[i]));
    }
    hi = make_uint4(h[0], h[1], h[2], h[3]);
    lo = make_uint4(l[0], l[1], l[2], l[3]);
}

// warp 0 of a warpgroup issues; operands are warp-uniform (see update_tc.cu: issue_stage)
__device__ __forceinline__ void issue_g0(uint32_t wb_, uint32_t sb_, uint32_t z_, const RtLayout& L, int net, uint64_t* done) {
    constexpr uint32_t id = idesc_f16(FMT, 128, 64, 0, 0);
    const uint32_t wb = __shfl_sync(0xffffffffu, wb_, 0), sb = __shfl_sync(0xffffffffu, sb_, 0);
    const uint32_t Z = __shfl_sync(0xffffffffu, z_, 0);
    const uint64_t da = narrow_desc_kmajor(wb + L.xe, RT_ROWS * 16);                        // {x.hi, x.lo}
    const uint64_t db = narrow_desc_kmajor(sb + L.p1h[net], RT_H * 16);                     // {P.hi, P.hi}
    const uint64_t db2 = narrow_desc_kmajor(sb + L.p1h[net] + 2 * RT_H * 16, RT_H * 16);    // {P.lo, 0}
    tc_fence_after();
    if (elect_one()) {
        mma_f16(Z, da, db, id, 0u);
        mma_f16(Z, da, db2, id, 1u);
        mma_commit(done);
    }
    __syncwarp();
}
__device__ __forceinline__ void issue_g1(uint32_t wb_, uint32_t sb_, uint32_t z_, const RtLayout& L, int net, uint64_t* done) {
    constexpr uint32_t id = idesc_f16(FMT, 128, 64, 0, 0);
    const uint32_t wb = __shfl_sync(0xffffffffu, wb_, 0), sb = __shfl_sync(0xffffffffu, sb_, 0);
    const uint32_t Z = __shfl_sync(0xffffffffu, z_, 0);
    const uint64_t ah = act_desc_kmajor(wb + L.a1h, 0), al = act_desc_kmajor(wb + L.a1l, 0);
    const uint64_t wh = act_desc_kmajor(sb + L.w2h[net], 0), wl = act_desc_kmajor(sb + L.w2l[net], 0);
    tc_fence_after();
    if (elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 4; ks++) mma_f16(Z, ah + 2u * ks, wh + 2u * ks, id, ks > 0);
#pragma unroll
        for (int ks = 0; ks < 4; ks++) mma_f16(Z, al + 2u * ks, wh + 2u * ks, id, 1u);
#pragma unroll
        for (int ks = 0; ks < 4; ks++) mma_f16(Z, ah + 2u * ks, wl + 2u * ks, id, 1u);
        mma_commit(done);
    }
    __syncwarp();
}

// Z (fp32, tensor memory, this thread's lane) -> h = tanh(Z) -> split 16-bit tiles A1 hi / lo (layer-1 epilogue)
__device__ __forceinline__ void layer1_epilogue(uint32_t zaddr, uint8_t* A1h, uint8_t* A1l, int wt) {
#pragma unroll
    for (int hlf = 0; hlf < 2; hlf++) {
        float h[32];
        tmem_ld16(zaddr + hlf * 32, h);
        tmem_ld16(zaddr + hlf * 32 + 16, h + 16);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i++) h[i] = tanh_acc(h[i]);
        uint4 hi[4], lo[4];
#pragma unroll
        for (int c = 0; c < 4; c++) split8(h + c * 8, hi[c], lo[c]);
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const uint32_t off = act_chunk_off(wt, hlf * 4 + c);
            *reinterpret_cast<uint4*>(A1h + off) = hi[c];
            *reinterpret_cast<uint4*>(A1l + off) = lo[c];
        }
    }
}
// Z -> h2 = tanh(Z + b2) in registers (layer-2 epilogue)
__device__ __forceinline__ void layer2_epilogue(uint32_t zaddr, const float* B2, float (&h2)[64]) {
#pragma unroll
    for (int c = 0; c < 4; c++) tmem_ld16(zaddr + c * 16, h2 + c * 16);
    tmem_ld_wait();
#pragma unroll
    for (int c = 0; c < 16; c++) {
        const float4 b = *reinterpret_cast<const float4*>(B2 + c * 4);
        h2[c * 4 + 0] = tanh_acc(h2[c * 4 + 0] + b.x); h2[c * 4 + 1] = tanh_acc(h2[c * 4 + 1] + b.y);
        h2[c * 4 + 2] = tanh_acc(h2[c * 4 + 2] + b.z); h2[c * 4 + 3] = tanh_acc(h2[c * 4 + 3] + b.w);
    }
}
__device__ __forceinline__ float out_dot(const float (&h2)[64], const float* W3row, float bias) {
    float a0 = bias, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
#pragma unroll
    for (int c = 0; c < 16; c++) {
        const float4 w = *reinterpret_cast<const float4*>(W3row + c * 4);
        a0 = fmaf(h2[c * 4 + 0], w.x, a0); a1 = fmaf(h2[c * 4 + 1], w.y, a1);
        a2 = fmaf(h2[c * 4 + 2], w.z, a2); a3 = fmaf(h2[c * 4 + 3], w.w, a3);
    }
    return (a0 + a1) + (a2 + a3);
}

template <class Env>
__global__ void __launch_bounds__(RT_NWG * 128, 1) rollout_tc_kernel(const RollArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    constexpr int S = Env::S, O = Env::OBS;
    static_assert(S == O && O <= 4, "observation = state, at most 4 floats");
    const int K = a.d.K, heads = a.d.heads;
    const RtLayout L = rt_layout(K);
    const NetOff offs[2] = {actor_off(O, RT_H, K), critic_off(O, RT_H)};
    const int n = a.cfg.n_envs, T = a.cfg.n_steps;
    const bool masked = a.cfg.masked != 0;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int g = warp >> 2, wt = tid & 127, w4 = warp & 3;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + L.bars);   // [g * 2 + net]: that accumulator's MMAs completed
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + L.tmem_slot);
    const uint64_t ctr0 = a.cfg.counter + (a.cfg.d_counter ? *a.cfg.d_counter : 0ull);
    double fin_cnt = 0.0, fin_len = 0.0, fin_ret = 0.0;

    // ---------------- prologue: tensor memory, barriers, weights -> shared memory ----------------
    if (warp == 0) tmem_alloc(tmem_slot, 512);
    if (tid == 0) {
        for (int i = 0; i < RT_NWG * 2; i++) mbar_init(bars + i, 1);
        mbar_init_fence();
    }
    {   // W2 of both nets, hi and lo tiles: 1024 chunks of 8 weights over 512 threads
        float v[2][8];
#pragma unroll
        for (int it = 0; it < 2; it++) {
            const int e = tid + it * RT_NWG * 128;
            const float* src = a.params + offs[e >> 9].W2 + ((e >> 3) & 63) * RT_H + (e & 7) * 8;
#pragma unroll
            for (int i = 0; i < 8; i++) v[it][i] = __ldg(src + i);
        }
#pragma unroll
        for (int it = 0; it < 2; it++) {
            const int e = tid + it * RT_NWG * 128;
            uint4 hi, lo;
            split8(v[it], hi, lo);
            const uint32_t off = act_chunk_off((e >> 3) & 63, e & 7);
            *reinterpret_cast<uint4*>(sm + L.w2h[e >> 9] + off) = hi;
            *reinterpret_cast<uint4*>(sm + L.w2l[e >> 9] + off) = lo;
        }
    }
    if (tid < 2 * RT_H) {   // layer-1 weights P1[j] = {W1[j][0..O), b1[j], 0..} as K chunks {hi, hi, lo, 0}
        const int net = tid >> 6, j = tid & 63;
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            v[i] = 0.0f;
            if (i < O) v[i] = __ldg(a.params + offs[net].W1 + j * O + i);
            if (i == O) v[i] = __ldg(a.params + offs[net].b1 + j);
        }
        uint4 hi, lo;
        split8(v, hi, lo);
        uint8_t* dst = sm + L.p1h[net] + j * 16;
        *reinterpret_cast<uint4*>(dst) = hi;
        *reinterpret_cast<uint4*>(dst + RT_H * 16) = hi;
        *reinterpret_cast<uint4*>(dst + 2 * RT_H * 16) = lo;
        *reinterpret_cast<uint4*>(dst + 3 * RT_H * 16) = make_uint4(0u, 0u, 0u, 0u);
    }
    for (int e = tid; e < 128; e += RT_NWG * 128)
        reinterpret_cast<float*>(sm + L.b2[e >> 6])[e & 63] = __ldg(a.params + offs[e >> 6].b2 + (e & 63));
    for (int e = tid; e < (K + 1) * 64; e += RT_NWG * 128) {
        const int q = e >> 6, k = e & 63;
        reinterpret_cast<float*>(sm + L.w3)[e] = q < K ? __ldg(a.params + offs[0].W3 + q * RT_H + k) : __ldg(a.params + offs[1].W3 + k);
    }
    for (int q = tid; q <= K; q += RT_NWG * 128)
        reinterpret_cast<float*>(sm + L.b3)[q] = q < K ? __ldg(a.params + offs[0].b3 + q) : __ldg(a.params + offs[1].b3);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    uint8_t* wg = sm + L.wg0 + g * L.wg_stride;
    const uint32_t wb = smem_u32(wg), sb = smem_u32(sm);
    float* x = reinterpret_cast<float*>(wg + L.lg) + wt * L.ldo;     // this env's logits
    const float* W3 = reinterpret_cast<const float*>(sm + L.w3);
    const float* B3 = reinterpret_cast<const float*>(sm + L.b3);
    uint64_t* done = bars + g * 2;
    uint32_t ph[2] = {0u, 0u};
    const uint32_t zc[2] = {tmem + (uint32_t)(g * 2) * 64u, tmem + (uint32_t)(g * 2 + 1) * 64u};   // Z[net], lane 0
    const uint32_t zl[2] = {tmem_addr(zc[0], (uint32_t)w4 * 32u, 0), tmem_addr(zc[1], (uint32_t)w4 * 32u, 0)};
    const int bar_id = 1 + g;

    // tiles are dealt CTA-first so that a small env count spreads over the SMs before it doubles up inside one
    const int ntiles = (n + RT_ROWS - 1) / RT_ROWS;
    for (int tile = g * gridDim.x + blockIdx.x; tile < ntiles; tile += RT_NWG * gridDim.x) {
        const int e = tile * RT_ROWS + wt;
        const bool on = e < n;
        float s[S];
        int32_t len = 0;
        float ret = 0.0f, nd = 0.0f;
#pragma unroll
        for (int k = 0; k < S; k++) s[k] = on ? a.state[(size_t)k * n + e] : 0.0f;
        if (on) {
            len = a.ep_len[e];
            ret = a.ep_ret[e];
            nd = a.next_done[e];
        }
        const uint32_t env_id = a.cfg.env_id_base + (uint32_t)e;
        for (int step = 0; step <= T; step++) {
            const size_t row = (size_t)step * n + e;
            // m_obs[step] = next_obs; m_dones[step] = next_done   (:529-530)
            if (on && step < T) {
                if constexpr (O == 4) *reinterpret_cast<float4*>(a.obs_buf + row * 4) = make_float4(s[0], s[1], s[2], s[3]);
                else if constexpr (O == 2) *reinterpret_cast<float2*>(a.obs_buf + row * 2) = make_float2(s[0], s[1]);
                else
                    for (int k = 0; k < O; k++) a.obs_buf[row * O + k] = s[k];
                a.dones_buf[row] = nd;
            }
            {   // XE row = {obs, 1, 0..} split hi / lo: A operand of G0
                float xe[8];
#pragma unroll
                for (int i = 0; i < 8; i++) xe[i] = i < O ? s[i < S ? i : 0] : (i == O ? 1.0f : 0.0f);
                uint4 hi, lo;
                split8(xe, hi, lo);
                *reinterpret_cast<uint4*>(wg + L.xe + wt * 16) = hi;
                *reinterpret_cast<uint4*>(wg + L.xe + RT_ROWS * 16 + wt * 16) = lo;
            }
            tc_fence_before();      // this thread's tensor-memory reads of the previous step are complete
            fence_async_smem();
            bar_sync(bar_id, 128);
            if (w4 == 0) {
                issue_g0(wb, sb, zc[0], L, 0, done + 0);
                issue_g0(wb, sb, zc[1], L, 1, done + 1);
            }
            // ---- actor: layer 1 epilogue -> A1; G1; layer 2 epilogue; logits
            mbar_wait(done + 0, ph[0]); ph[0] ^= 1u;
            tc_fence_after();
            layer1_epilogue(zl[0], wg + L.a1h, wg + L.a1l, wt);
            tc_fence_before();
            fence_async_smem();
            bar_sync(bar_id, 128);
            if (w4 == 0) issue_g1(wb, sb, zc[0], L, 0, done + 0);
            mbar_wait(done + 0, ph[0]); ph[0] ^= 1u;      // G1 (actor) complete: Z[0] = layer-2 pre-activation, A1 free
            tc_fence_after();
            {
                float h2[64];
                layer2_epilogue(zl[0], reinterpret_cast<const float*>(sm + L.b2[0]), h2);
                for (int q = 0; q < K; q++) x[q] = out_dot(h2, W3 + q * RT_H, B3[q]);
            }
            // ---- critic: layer 1 epilogue -> A1; G1
            mbar_wait(done + 1, ph[1]); ph[1] ^= 1u;
            tc_fence_after();
            layer1_epilogue(zl[1], wg + L.a1h, wg + L.a1l, wt);
            tc_fence_before();
            fence_async_smem();
            bar_sync(bar_id, 128);
            if (w4 == 0) issue_g1(wb, sb, zc[1], L, 1, done + 1);
            // ---- while the critic's G1 runs: sample the action from the actor's logits
            int32_t acts[PPO_MAX_HEADS];
            RowOut ro{};
            uint64_t mbits = ~0ull;
            const uint64_t ctr = ctr0 + (uint64_t)step;
            if (on && step < T) {
                mbits = masked ? (uint64_t)Env::action_mask(s) : ~0ull;
                ro = row_eval(x, heads, a.d.nvec, mbits, masked, nullptr, nullptr, a.cfg.seed, env_id, ctr,
                              a.cfg.entropy_mode, acts);
            }
            mbar_wait(done + 1, ph[1]); ph[1] ^= 1u;
            tc_fence_after();
            float value;
            {
                float h2[64];
                layer2_epilogue(zl[1], reinterpret_cast<const float*>(sm + L.b2[1]), h2);
                value = out_dot(h2, W3 + K * RT_H, B3[K]);
            }
            if (step == T) {  // bootstrap: next_value = critic(next_obs)  (calcAdvantage :280)
                if (on) {
                    a.next_value[e] = value;
                    a.next_done[e] = nd;
                }
                break;
            }
            if (on) {
                for (int h = 0; h < heads; h++) a.actions_buf[row * heads + h] = acts[h];
                a.logprobs_buf[row] = ro.logp;
                a.values_buf[row] = value;
                if (a.mask_buf)
                    for (int j = 0; j < K; j++) a.mask_buf[row * K + j] = (uint8_t)((mbits >> j) & 1ull);
                // stepEnvs (:413-483): step, truncate, auto-reset, episode statistics
                StepOut so = Env::step_multi(s, acts);
                len += 1;
                ret = __fadd_rn(ret, so.reward);
                const bool fin = so.terminated || (len == a.cfg.max_episode_steps);
                const int32_t f_len = fin ? len : 0;
                const float f_ret = fin ? ret : 0.0f;
                if (fin) {
                    fin_cnt += 1.0; fin_len += (double)len; fin_ret += (double)ret;
                    float u[4];
                    reset_uniforms(a.cfg.seed, env_id, ctr, u);
                    Env::reset(s, u);
                    len = 0; ret = 0.0f;
                }
                a.rewards_buf[row] = so.reward;
                if (a.cfg.fin_len) { a.cfg.fin_len[row] = f_len; a.cfg.fin_ret[row] = f_ret; }
                nd = fin ? 1.0f : 0.0f;
            }
        }
        if (on) {
#pragma unroll
            for (int k = 0; k < S; k++) a.state[(size_t)k * n + e] = s[k];
            a.ep_len[e] = len;
            a.ep_ret[e] = ret;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (a.ep_stats) {   // finished-episode statistics: warp sums, one atomic per warp and quantity
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            fin_cnt += __shfl_xor_sync(0xffffffffu, fin_cnt, o);
            fin_len += __shfl_xor_sync(0xffffffffu, fin_len, o);
            fin_ret += __shfl_xor_sync(0xffffffffu, fin_ret, o);
        }
        if ((tid & 31) == 0 && fin_cnt != 0.0) {
            atomicAdd(a.ep_stats + 0, fin_cnt);
            atomicAdd(a.ep_stats + 1, fin_len);
            atomicAdd(a.ep_stats + 2, fin_ret);
        }
    }
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem, 512);
    }
}

}  // namespace

bool rollout_tc_supported(const Dims& d) { return d.H == RT_H && d.K <= RT_MAX_K && d.O <= 4; }

template <class Env>
int launch_rollout_tc(const RollArgs& a, cudaStream_t st) {
    const RtLayout L = rt_layout(a.d.K);
    const size_t smem = (size_t)L.end + 1024;
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(rollout_tc_kernel<Env>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_done = true;
    }
    PPO_REQUIRE(smem <= 227 * 1024, "rollout_tc: needs %zu B of shared memory", smem);
    const int ntiles = (a.cfg.n_envs + RT_ROWS - 1) / RT_ROWS;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int grid = ntiles < sms ? ntiles : sms;
    rollout_tc_kernel<Env><<<grid, RT_NWG * 128, smem, st>>>(a);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}
template int launch_rollout_tc<CartPoleEnv>(const RollArgs&, cudaStream_t);
template int launch_rollout_tc<MountainCarEnv>(const RollArgs&, cudaStream_t);
template int launch_rollout_tc<MultiWalkEnv>(const RollArgs&, cudaStream_t);

}  // namespace ppo
