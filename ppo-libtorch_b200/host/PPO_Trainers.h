// PPO_Trainers.h — host facade with the reference's class surface over the C ABI (include/ppo_b200.h):
//   class PPO_Discrete       PPO/PPO_Discrete.h:24-108      (CartPole, one Categorical head)
//   class PPO_MultiDiscrete  PPO/PPO_MultiDiscrete.h:23-111  (MountainCar, CategoricalMasked head, mask buffer)
// Same life cycle: the default constructor reads ./PPOConfig.toml from the working directory (getArgs, :107-255),
// builds the agent (orthogonal init, Agent.cpp:91-99) and AdamW(lr).eps(1e-5) (:76-78), auto-resumes from the newest
// file in ./ModelCheckpoints + ./OptimizerCheckpoints (loadPolicyFromCheckpoint, :782-835), and train() runs
// (total_timesteps - global_step) / batch_size updates (:485-690), printing the SB3-style table (:700-774) and
// checkpointing every checkpoint_updates updates (:662-673) plus the final pair under ./Models (:678-685).
// The whole rollout -> GAE -> E x MB optimiser steps of one update is ONE ppo_trainer_iterate call (a CUDA graph of
// sm_100a kernels); the host only touches scalars.  Hyper-parameters are public members with the reference's names.
// Errors surface as std::runtime_error carrying ppo_last_error(), which the reference's main() already catches
// (driver.cpp:15-17).  There is no CPU fallback: use_cuda = false, or no CUDA device, throws.
#pragma once
#include <chrono>
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "../../include/ppo_b200.h"
#include "PPOConfig.h"
#include "PPOUtils.h"

namespace ppo_host {
struct CheckpointShell;
}

class PPOTrainerBase {
public:
    virtual ~PPOTrainerBase();
    void train();
    void getArgs();                       // re-reads ./PPOConfig.toml into the members below
    void loadPolicyFromCheckpoint();
    void saveCheckpoint(const std::string& agent_file, const std::string& optimizer_file);
    void printPPOResults(int64_t update, int64_t global_step, std::chrono::milliseconds fps,
                         std::chrono::milliseconds time_elapsed, const ppo_iter_result& r);
    // one PPO update (rollout + GAE + all optimiser steps) at learning rate lr; what train() loops over.  Virtual: the
    // host-environment bridge (PPO_HostEnv.h) runs the rollout on the host and only the update on the device.
    virtual ppo_iter_result iterate(float lr);
    // true: finished episodes are pushed into m_episode_stats by iterate() itself (host envs), not read from the device log
    virtual bool pushesEpisodeStats() const { return false; }
    std::vector<float> getParameters() const;          // flat slab, reference registration order
    ppo_trainer* handle() const { return m_b200; }
    bool multidiscrete() const { return m_multidiscrete; }
    void* stream() const { return m_stream; }
    int envKind() const { return m_env_kind; }

    // Hyperparameters (PPO/PPO_Discrete.h:51-76)
    int64_t m_obs_size, m_action_size;
    float m_learning_rate;
    int64_t m_seed, m_total_timesteps;
    bool m_use_cuda, m_torch_deterministic;
    int64_t m_num_envs, m_num_steps;
    bool m_anneal_lr, m_use_gae;
    float m_gamma, m_gae_lambda;
    int64_t m_num_minibatches, m_update_epochs;
    bool m_norm_adv;
    float m_clip_coef;
    bool m_clip_vloss;
    float m_ent_coef, m_vf_coef, m_max_grad_norm;
    int64_t m_checkpoint_updates, m_max_episode_steps;
    int64_t m_batch_size, m_minibatch_size;
    uint64_t m_global_step = 0;
    std::vector<float> m_clipfracs;                    // per-update mean clip fraction (table row clip_fraction)
    std::unique_ptr<ppo_host::CircularBuffer> m_episode_stats;
    ppo_host::Config m_config;

protected:
    PPOTrainerBase(bool multidiscrete, const std::string& config_path);
    void build();                                      // agent shell, device trainer, resume
    virtual int deviceEnvKind(const std::string& env_name) const;   // registry lookup; the host bridge returns PPO_ENV_HOST
    void* ensureStream();
    void noteLearningRate(float lr) { m_last_lr = lr; }   // what a checkpoint records as the optimizer's lr

private:
    bool m_multidiscrete;
    int m_env_kind = 0;
    std::string m_config_path;
    ppo_trainer* m_b200 = nullptr;
    ppo_host::CheckpointShell* m_shell = nullptr;
    int64_t m_opt_steps = 0;
    float m_last_lr = -1.0f;                           // learning rate of the last iterate() (what a checkpoint records)
    void* m_stream = nullptr;                          // cudaStream_t the iterations are enqueued on
    template <typename T> void printElement(T t, int width);
};

// class PPO_Discrete / class PPO_MultiDiscrete: PPO_Discrete.h / PPO_MultiDiscrete.h (they add the tensor-level surface)
