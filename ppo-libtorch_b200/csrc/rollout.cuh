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

}  // namespace ppo
