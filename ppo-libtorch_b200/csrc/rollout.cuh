// rollout.cuh — argument block shared by the rollout kernels (rollout.cu: FMA tiles; rollout_tc.cu: tcgen05 tiles)
#pragma once
#include "env.cuh"
#include "policy.cuh"

namespace ppo {

struct RollArgs {
    const float* params;
    Dims d;
    ppo_rollout_cfg cfg;
    float* state;
    int32_t* ep_len;
    float* ep_ret;
    float* next_done;
    float* next_value;
    float* obs_buf;
    int32_t* actions_buf;
    float* logprobs_buf;
    float* rewards_buf;
    float* dones_buf;
    float* values_buf;
    uint8_t* mask_buf;
    double* ep_stats;
};

// rollout_tc.cu: H = 64, 128-env tiles with both 64 x 64 layers on the tensor core (split 16-bit operands, fp32-accurate)
template <class Env>
int launch_rollout_tc(const RollArgs& a, cudaStream_t st);
bool rollout_tc_supported(const Dims& d);

// rollout_h256.cu: H = 256 in the 16-bit arithmetic modes: actor-only rollout kernel with W2 resident + one batched critic
// pass for m_values / next_value (fmt: umma::FMT_F16 / FMT_BF16)
template <class Env>
int launch_rollout_h256(const RollArgs& a, int fmt, cudaStream_t st);
bool rollout_h256_supported(const Dims& d);
int launch_value_h256(const float* params, const Dims& d, const float* obs_a, float* out_a, int64_t rows_a,
                      const float* soa_b, float* out_b, int rows_b, int fmt, cudaStream_t st);

}  // namespace ppo
