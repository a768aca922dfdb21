// PPO_MultiDiscrete.h — class PPO_MultiDiscrete (PPO/PPO_MultiDiscrete.h:23-111): MountainCar by default, CategoricalMasked
// heads, action-mask buffer m_action_masks.  Same surface as PPO_Discrete plus the masked overloads (PPO_Discrete.h).
// `[b200] nvec = [3, 3, 3, 3, 3, 2, 2, 2]` in PPOConfig.toml builds the multi-head agent the reference leaves commented
// out at PPO/Agent.cpp:16.
#pragma once
#include "PPO_Discrete.h"

class PPO_MultiDiscrete : public PPOTensorTrainer {
public:
    explicit PPO_MultiDiscrete(const std::string& config_path = "./PPOConfig.toml") : PPOTensorTrainer(true, config_path) {
        build();
        bindTensors();
    }
};
