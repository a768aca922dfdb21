// PPO_Discrete.h — class PPO_Discrete with the reference's full public surface (PPO/PPO_Discrete.h:24-108): the life
// cycle / hyper-parameter members / train() of PPOTrainerBase (PPO_Trainers.h) plus the tensor-level methods and members
//   computeActionLogic, calcAdvantage, getApproxKLAndClippedObj, initEnvs, stepEnvs,
//   m_device, m_agent, m_obs, m_actions, m_logprobs, m_rewards, m_dones, m_values, m_action_masks
// (PPOTensorTrainer below; PPO_MultiDiscrete.h adds the masked signatures).  The tensors are zero-copy torch views of the
// device trainer's arena (torch::from_blob on ppo_trainer_buffer pointers); m_agent's parameters alias the trainer's
// parameter slab, so a forward through m_agent always sees the weights train() is optimising.  Every method is one or
// two launches of libppo_b200 kernels; none of them is needed by train(), which runs the whole update as ONE
// ppo_trainer_iterate call — they exist so that code written against the reference's per-step API keeps working.
// Layout differences from the reference, stated: m_actions is int32 [T, N, heads] (the reference stores float [T, N, 1]).
#pragma once
#include <torch/torch.h>

#include <array>
#include <memory>
#include <tuple>

#include "Agent.h"
#include "PPO_Trainers.h"

class PPOTensorTrainer : public PPOTrainerBase {
public:
    // ALGO LOGIC (PPO/PPO_Discrete.h:35-38)
    AgentOutput computeActionLogic(const torch::Tensor& next_obs) const;
    AgentOutput computeActionLogic(const torch::Tensor& next_obs, const torch::Tensor& action_mask,
                                   const torch::Tensor& input_action = torch::Tensor()) const;   // PPO_MultiDiscrete.h:34
    std::array<torch::Tensor, 2> calcAdvantage(const torch::Tensor& next_obs, const torch::Tensor& next_done) const;
    torch::Tensor getApproxKLAndClippedObj(const torch::Tensor& ratio, const torch::Tensor& logratio);
    // Controlling Environments (:41-43).  The envs are the device trainer's SoA state, stepped by env_step_kernel.
    torch::Tensor initEnvs() const;
    torch::Tensor initEnvs(const torch::Tensor& action_mask);                                     // PPO_MultiDiscrete.h:42
    std::tuple<torch::Tensor, torch::Tensor, torch::Tensor> stepEnvs(const torch::Tensor& action);

    std::shared_ptr<torch::Device> m_device;
    std::shared_ptr<Agent> m_agent;
    torch::Tensor m_obs = torch::Tensor();
    torch::Tensor m_actions = torch::Tensor();
    torch::Tensor m_logprobs = torch::Tensor();
    torch::Tensor m_rewards = torch::Tensor();
    torch::Tensor m_dones = torch::Tensor();
    torch::Tensor m_values = torch::Tensor();
    torch::Tensor m_action_masks = torch::Tensor();

protected:
    PPOTensorTrainer(bool multidiscrete, const std::string& config_path) : PPOTrainerBase(multidiscrete, config_path) {}
    void bindTensors();   // after build(): arena views + m_agent aliasing the trainer's parameter slab

private:
    torch::Tensor arenaView(const char* name, c10::ScalarType dtype, std::vector<int64_t> shape) const;
    uint64_t m_step_calls = 0;   // stepEnvs counter: keys the auto-reset draws
};

class PPO_Discrete : public PPOTensorTrainer {
public:
    explicit PPO_Discrete(const std::string& config_path = "./PPOConfig.toml") : PPOTensorTrainer(false, config_path) {
        build();
        bindTensors();
    }
};
