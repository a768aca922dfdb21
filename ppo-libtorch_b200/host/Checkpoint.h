// Checkpoint.h — the reference's checkpoint layout (PPO/PPO_Discrete.cpp:662-685 save, :782-835 resume),
// kept byte-compatible by going through LibTorch exactly as the reference does: a host-side shell of the Agent module
// (same registered names: m_Critic.{criticInputLayer,Tanh1,criticMiddleLayer,Tanh2,criticOutputLayer}, m_Actor.{actor...},
// PPO/Agent.cpp:38-66) and a torch::optim::AdamW over its parameters are torch::save()d / torch::load()ed; the device
// slabs are copied in and out at those boundaries only.  This header is torch-free; Checkpoint.cpp is the only
// translation unit that includes LibTorch.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace ppo_host {

struct CheckpointShell;   // Agent shell + AdamW (opaque: LibTorch types)

// obs/hidden/actions define the 12 parameter tensors; lr etc. only seed the optimiser's option block
CheckpointShell* checkpoint_shell_create(int64_t obs_size, int64_t hidden, int64_t action_size, double lr, int64_t seed);
void checkpoint_shell_destroy(CheckpointShell* s);
int64_t checkpoint_num_params(const CheckpointShell* s);
// flat parameter slab in the reference's registration order (critic then actor, weights [out,in] row-major)
void checkpoint_get_params(const CheckpointShell* s, float* flat);       // e.g. the orthogonal init (Agent.cpp:91-99)
void checkpoint_set_params(CheckpointShell* s, const float* flat);
// AdamW state <-> flat slabs (exp_avg, exp_avg_sq, step); lr goes into / comes from the param-group options
void checkpoint_set_opt_state(CheckpointShell* s, const float* exp_avg, const float* exp_avg_sq, int64_t step, double lr);
bool checkpoint_get_opt_state(const CheckpointShell* s, float* exp_avg, float* exp_avg_sq, int64_t* step, double* lr);
// torch::save / torch::load of the two archives; throw std::exception (c10::Error) like the reference's calls
void checkpoint_save_agent(const CheckpointShell* s, const std::string& file);
void checkpoint_save_optimizer(const CheckpointShell* s, const std::string& file);
void checkpoint_load_agent(CheckpointShell* s, const std::string& file);
void checkpoint_load_optimizer(CheckpointShell* s, const std::string& file);

}  // namespace ppo_host
