// host_capi.cpp — C entry points over the host-side (non-CUDA) pieces, so that the pytest suite can check them without a
// GPU: the PPOConfig.toml reader (PPOConfig.h), the checkpoint file-name helpers and the episode-statistics ring.
#include <cstring>

#include "PPOConfig.h"
#include "PPOUtils.h"

extern "C" {
struct ppo_host_config {
    int64_t obs_size, action_size, max_episode_steps, seed, total_timesteps, checkpoint_updates, num_envs, num_steps,
        num_minibatches, update_epochs, hidden, batch_size, minibatch_size;
    float learning_rate, gamma, gae_lambda, clip_coef, ent_coef, vf_coef, max_grad_norm, action_high, action_low;
    int32_t use_cuda, torch_deterministic, anneal_lr, use_gae, norm_adv, clip_vloss, use_cuda_graph, exact_entropy;
    char env[32];
};
// returns load_config's status (0 loaded, 1 missing, 2 parse error); `is_multidiscrete` selects that class's defaults
int ppo_host_load_config(const char* path, int is_multidiscrete, ppo_host_config* o) {
    ppo_host::Config c;
    if (is_multidiscrete) c.max_episode_steps = 200;   // PPO_MultiDiscrete.cpp:31
    std::ostringstream sink;
    const int rc = ppo_host::load_config(path, c, sink);
    if (rc == 2 && is_multidiscrete) c.max_episode_steps = 200;
    *o = ppo_host_config{c.obs_size, c.action_size, c.max_episode_steps, c.seed, c.total_timesteps, c.checkpoint_updates,
                         c.num_envs, c.num_steps, c.num_minibatches, c.update_epochs, c.hidden, c.batch_size,
                         c.minibatch_size, c.learning_rate, c.gamma, c.gae_lambda, c.clip_coef, c.ent_coef, c.vf_coef,
                         c.max_grad_norm, c.action_high, c.action_low, c.use_cuda, c.torch_deterministic, c.anneal_lr,
                         c.use_gae, c.norm_adv, c.clip_vloss, c.use_cuda_graph, c.exact_entropy, {0}};
    std::strncpy(o->env, c.env.c_str(), sizeof(o->env) - 1);
    return rc;
}
// Utils.cpp:38-48 + PPO_Discrete.cpp:809-811: step count after `prefix` in a checkpoint file name (0 if not a number);
// parsed as 64-bit (the reference's stoi throws above 2^31-1)
uint64_t ppo_host_step_from_filename(const char* file_name, const char* prefix) {
    return ppo_host::step_from_filename(file_name, prefix);
}
// CircularBuffer(capacity) fed with n (reward, length) pairs: means of what it holds (Utils.h:30-79)
void ppo_host_circular_means(int capacity, int n, const float* rewards, const int64_t* lengths, double* mean_len,
                             float* mean_rew) {
    ppo_host::CircularBuffer b((size_t)capacity);
    for (int i = 0; i < n; i++) b.add(rewards[i], lengths[i]);
    *mean_len = b.avgLength();
    *mean_rew = b.avgReward();
}
}
