// policy.cu — K1: fused actor-critic forward (+ K2 epilogue) on arbitrary [n, O] observations.
// Replaces Agent::getActionAndValueDiscrete / getActionAndValueMasked / getValue
// (PPO/Agent.cpp:107-170): ~10 addmm/tanh launches + the Categorical op chain become one launch.
#include "policy.cuh"

namespace ppo {

struct FwdArgs {
    const float* params;
    Dims d;
    int n;
    const float* obs;
    const uint8_t* mask;
    const int32_t* actions_in;
    const float* uniforms;
    uint64_t seed, counter;
    int entropy_mode;
    int32_t* actions_out;
    float* logprob;
    float* entropy;
    float* value;
    float* logits_out;
    int want_policy;  // 0 => critic only (getValue)
};

template <int TM, int H>
__global__ void __launch_bounds__(MLP_THREADS) policy_forward_kernel(const FwdArgs a) {
    extern __shared__ __align__(16) float sm[];
    const int O = a.d.O, K = a.d.K;
    const TileLayout L = tile_layout(TM, H, O, K);
    const NetOff co = critic_off(O, H), ao = actor_off(O, H, K);
    const int ntiles = (a.n + TM - 1) / TM;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t r0 = (int64_t)tile * TM;
        const int rows = (int)min((int64_t)TM, (int64_t)a.n - r0);
        __syncthreads();
        for (int e = threadIdx.x; e < TM * O; e += MLP_THREADS)
            sm[L.xs + e] = (e < rows * O) ? __ldg(a.obs + r0 * O + e) : 0.0f;
        __syncthreads();
        if (a.want_policy) net_forward<TM, H>(sm, L, O, a.params, ao, K, sm + L.ls, L.ldo);
        if (a.value) net_forward<TM, H>(sm, L, O, a.params, co, 1, sm + L.vs, 1);
        const int t = threadIdx.x;
        if (t < rows) {
            const int64_t r = r0 + t;
            if (a.value) a.value[r] = sm[L.vs + t];
            if (a.want_policy) {
                const float* x = sm + L.ls + t * L.ldo;
                const bool masked = a.mask != nullptr;
                const uint64_t mbits = masked ? pack_mask(a.mask + r * K, K) : ~0ull;
                int32_t acts[PPO_MAX_HEADS];
                const RowOut o = row_eval(x, a.d.heads, a.d.nvec, mbits, masked,
                                          a.actions_in ? a.actions_in + r * a.d.heads : nullptr,
                                          a.uniforms ? a.uniforms + r * a.d.heads : nullptr, a.seed, (uint32_t)r,
                                          a.counter, a.entropy_mode, acts);
                if (a.actions_out)
                    for (int h = 0; h < a.d.heads; h++) a.actions_out[r * a.d.heads + h] = acts[h];
                if (a.logprob) a.logprob[r] = o.logp;
                if (a.entropy) a.entropy[r] = o.entropy;
                if (a.logits_out)
                    for (int j = 0; j < K; j++) a.logits_out[r * K + j] = x[j];
            }
        }
    }
}

template <int TM, int H>
static int launch_fwd(const FwdArgs& a, cudaStream_t st) {
    const TileLayout L = tile_layout(TM, H, a.d.O, a.d.K);
    const size_t smem = (size_t)L.end * sizeof(float);
    static bool attr_done = false;
    if (!attr_done) {
        PPO_CUDA(cudaFuncSetAttribute(policy_forward_kernel<TM, H>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      200 * 1024));
        attr_done = true;
    }
    PPO_REQUIRE(smem <= 200 * 1024, "policy forward: tile needs %zu B of shared memory", smem);
    const int ntiles = (a.n + TM - 1) / TM;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int grid = ntiles < 2 * sms ? ntiles : 2 * sms;
    policy_forward_kernel<TM, H><<<grid, MLP_THREADS, smem, st>>>(a);
    PPO_LAUNCH_CHECK();
    return PPO_OK;
}

static int dispatch_fwd(const FwdArgs& a, cudaStream_t st) {
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const bool big = a.n >= 128 * sms;
    switch (a.d.H) {
        case 64: return big ? launch_fwd<128, 64>(a, st) : launch_fwd<32, 64>(a, st);
        case 128: return big ? launch_fwd<64, 128>(a, st) : launch_fwd<32, 128>(a, st);
        case 256: return launch_fwd<32, 256>(a, st);
        default: set_error("hidden=%d unsupported (64, 128 or 256)", a.d.H); return PPO_ERR_UNSUPPORTED;
    }
}

}  // namespace ppo

using namespace ppo;

extern "C" int ppo_policy_forward(const float* params, const ppo_dims* dims, int n, const float* obs,
                                  const uint8_t* mask, const int32_t* actions_in, const float* uniforms,
                                  uint64_t seed, uint64_t counter, int entropy_mode, int32_t* actions_out,
                                  float* logprob_out, float* entropy_out, float* value_out, float* logits_out,
                                  ppo_stream_t stream) {
    FwdArgs a{};
    if (int rc = make_dims(dims, &a.d)) return rc;
    PPO_REQUIRE(n >= 0 && params && obs, "ppo_policy_forward: null params/obs or negative n");
    if (n == 0) return PPO_OK;
    a.params = params; a.n = n; a.obs = obs; a.mask = mask; a.actions_in = actions_in; a.uniforms = uniforms;
    a.seed = seed; a.counter = counter; a.entropy_mode = entropy_mode; a.actions_out = actions_out;
    a.logprob = logprob_out; a.entropy = entropy_out; a.value = value_out; a.logits_out = logits_out;
    a.want_policy = 1;
    return dispatch_fwd(a, as_stream(stream));
}

extern "C" int ppo_value_forward(const float* params, const ppo_dims* dims, int n, const float* obs,
                                 float* value_out, ppo_stream_t stream) {
    FwdArgs a{};
    if (int rc = make_dims(dims, &a.d)) return rc;
    PPO_REQUIRE(n >= 0 && params && obs && value_out, "ppo_value_forward: null buffer or negative n");
    if (n == 0) return PPO_OK;
    a.params = params; a.n = n; a.obs = obs; a.value = value_out; a.want_policy = 0;
    return dispatch_fwd(a, as_stream(stream));
}
