// Agent.h — the reference's actor-critic module (PPO/Agent.h:22-54, PPO/Agent.cpp:14-175) over the C ABI.
// Same class surface: a torch::nn::Module with m_Critic / m_Actor Sequentials under the reference's registered names
// (the checkpoint layout), ppoLayerInit (orthogonal weights, constant bias), getValue, getActionAndValueDiscrete,
// getActionAndValueMasked, printAgent, m_actionSpace / m_actionSpaceSum / m_actionSpaceSize / m_device.
// The three forward methods do not run the Sequentials: the 12 parameter tensors are views into ONE flat fp32 slab in
// the reference's registration order (critic, then actor; [out, in] row-major) and a call is one fused sm_100a kernel —
// ppo_policy_forward (actor MLP + critic MLP + Categorical[Masked] sample / log-prob / entropy) or ppo_value_forward.
// Inputs must be CUDA tensors; there is no CPU fallback.
// Extensions (defaulted, so reference call sites compile unchanged): `hidden` (the reference hard-codes 64 and keeps 256
// as a commented alternative, Agent.cpp:25-37) and `actionSpace` (the MultiDiscrete nvec the reference leaves commented
// at Agent.cpp:16).
#pragma once
#include <torch/torch.h>

#include <cmath>
#include <memory>
#include <vector>

#include "Distributions.h"

// Define agent return structs for readability of code (PPO/Agent.h:22-27)
struct AgentOutput {
    torch::Tensor action;
    torch::Tensor logprob;
    torch::Tensor entropy;
    torch::Tensor value;
};

class Agent : public torch::nn::Module {
public:
    Agent(int64_t obsSize, int64_t actionSize, std::shared_ptr<torch::Device> device, int64_t hidden = 64,
          std::vector<int64_t> actionSpace = {});
    ~Agent();
    torch::nn::Linear ppoLayerInit(torch::nn::Linear layer, const double stdDev = sqrt(2), const double bias_const = 0.0);
    torch::Tensor getValue(const torch::Tensor& x);
    AgentOutput getActionAndValueDiscrete(const torch::Tensor& x, torch::Tensor action = torch::Tensor());
    AgentOutput getActionAndValueMasked(const torch::Tensor& x, const torch::Tensor& mask, torch::Tensor action = torch::Tensor());
    void printAgent();

    torch::nn::Sequential m_Critic{nullptr};
    torch::nn::Sequential m_Actor{nullptr};
    std::shared_ptr<torch::Tensor> m_actor_logstd;
    std::vector<int64_t> m_actionSpace;
    int64_t m_actionSpaceSum;
    int64_t m_actionSpaceSize;
    std::shared_ptr<torch::Device> m_device;

    // --- B200 additions ---
    // The flat parameter slab the kernels read.  Re-established (one copy) if the parameters were rebound since the last
    // call, e.g. by torch::load(agent, file) or agent->to(device).
    torch::Tensor flatParameters();
    // Make the parameters views of an EXTERNAL slab (the device trainer's arena): current values are copied into it.
    void aliasParameters(torch::Tensor flat);
    int64_t hidden() const { return m_hidden; }
    int64_t obsSize() const { return m_obsSize; }
    // entropy of the unmasked path: false (default) = the reference's clamp quirk (Categorical.cpp:114-117), true = exact
    bool m_exact_entropy = false;

private:
    int64_t m_obsSize, m_hidden;
    torch::Tensor m_flat;
    AgentOutput forward_fused(const torch::Tensor& x, const torch::Tensor* mask, torch::Tensor action, bool masked_api);
};
