// PPO_HostEnv.h — SURVEY.md §8 f4: the host-environment bridge.  Keeps the reference's LITERAL environment interface usable
// for environments that cannot be written as device functors (a game engine, a simulator behind a socket; README
// "Custom Environment Framework", PPO/PPO_Discrete.cpp:413-483):
//
//     struct MyEnv {                                        // duck-typed, exactly as the reference's CartPole / MountainCar
//         std::vector<float> reset();
//         std::tuple<std::vector<float>, float, bool, bool> step(const int64_t& action);   // obs, reward, terminated, info
//         int64_t episode_length;  float episode_reward;    // read by the trainer (:443, :455-456)
//     };
//     PPO_DiscreteHostEnv<MyEnv> algo;  algo.train();       // same life cycle, config file, table and checkpoints
//
// What stays on the host: the environments (m_envs, one job per env per step on a thread pool, truncation at
// max_episode_steps, auto-reset, episode statistics — stepEnvs as in the reference).  What runs on the GPU: the policy
// forward of every step (one fused kernel on the [N, O] observations) and the whole update (ppo_trainer_update: GAE,
// advantage statistics, E x MB fused forward / loss / backward / AdamW steps).  Per step the bridge moves N actions down and
// N observations + rewards + done flags up through pinned staging buffers.  This is NOT a CPU fallback for the kernels:
// only the user's environment code runs on the host.
#pragma once
#include <cuda_runtime_api.h>

#include "PPO_Discrete.h"
#include "ThreadPool.h"

template <class Env>
class PPO_DiscreteHostEnv : public PPOTensorTrainer {
public:
    explicit PPO_DiscreteHostEnv(const std::string& config_path = "./PPOConfig.toml") : PPOTensorTrainer(false, config_path) {
        build();
        bindTensors();
        m_threadPool = std::make_shared<ppo_host::ThreadPool>();
        for (int64_t i = 0; i < m_num_envs; i++) m_envs.push_back(std::make_shared<Env>());
        const size_t N = (size_t)m_num_envs, O = (size_t)m_obs_size;
        if (cudaMallocHost(reinterpret_cast<void**>(&m_h_obs), N * O * 4) != cudaSuccess ||
            cudaMallocHost(reinterpret_cast<void**>(&m_h_rew), N * 4) != cudaSuccess ||
            cudaMallocHost(reinterpret_cast<void**>(&m_h_done), N * 4) != cudaSuccess ||
            cudaMallocHost(reinterpret_cast<void**>(&m_h_act), N * 8) != cudaSuccess)
            throw std::runtime_error("PPO_DiscreteHostEnv: cudaMallocHost failed");
        auto f32 = torch::TensorOptions().dtype(torch::kFloat32).device(*m_device);
        m_next_obs = torch::zeros({m_num_envs, m_obs_size}, f32);
        m_next_done = torch::zeros({m_num_envs}, f32);
    }
    ~PPO_DiscreteHostEnv() override {
        cudaFreeHost(m_h_obs); cudaFreeHost(m_h_rew); cudaFreeHost(m_h_done); cudaFreeHost(m_h_act);
    }

    // PPO_Discrete.cpp:365-402: reset every env on the pool, observations [N, O] on the device
    torch::Tensor initEnvs() {
        const int64_t N = m_num_envs, O = m_obs_size;
        m_threadPool->parallelFor(N, [&](int64_t i) {
            const std::vector<float> o = m_envs[i]->reset();
            if ((int64_t)o.size() != O) throw std::runtime_error("environment observation size does not match obs_size");
            for (int64_t k = 0; k < O; k++) m_h_obs[i * O + k] = o[k];
        });
        return upload(m_h_obs, {N, O});
    }
    // PPO_Discrete.cpp:413-483: one step of every env with truncation and auto-reset; episode statistics in env order
    std::tuple<torch::Tensor, torch::Tensor, torch::Tensor> stepEnvs(const torch::Tensor& action) {
        const int64_t N = m_num_envs, O = m_obs_size;
        cudaStream_t st = static_cast<cudaStream_t>(ppo_host::stream_of(m_next_obs));
        torch::Tensor a64 = action.to(torch::kLong).reshape({N}).contiguous();
        cudaMemcpyAsync(m_h_act, a64.data_ptr<int64_t>(), (size_t)N * 8, cudaMemcpyDeviceToHost, st);
        cudaStreamSynchronize(st);
        std::vector<int64_t> fin_len((size_t)N, 0);
        std::vector<float> fin_ret((size_t)N, 0.0f);
        m_threadPool->parallelFor(N, [&](int64_t i) {
            Env& e = *m_envs[i];
            auto [obs, reward, terminated, info] = e.step(m_h_act[i]);
            (void)info;
            if (e.episode_length == m_max_episode_steps) terminated = true;          // :443
            if (terminated) {                                                         // :453-458
                fin_len[(size_t)i] = e.episode_length;
                fin_ret[(size_t)i] = e.episode_reward;
                obs = e.reset();
            }
            for (int64_t k = 0; k < O; k++) m_h_obs[i * O + k] = obs[k];
            m_h_rew[i] = reward;
            m_h_done[i] = terminated ? 1.0f : 0.0f;
        });
        for (int64_t i = 0; i < N; i++)                                               // :474-480
            if (fin_len[(size_t)i] > 0) m_episode_stats->add(fin_ret[(size_t)i], fin_len[(size_t)i]);
        torch::Tensor obs = upload(m_h_obs, {N, O}), rew = upload(m_h_rew, {N, 1}), done = upload(m_h_done, {N, 1});
        return {obs, rew, done.to(torch::kInt32)};
    }

    // one update = the reference's rollout loop (:524-554) against the host envs + the device update
    ppo_iter_result iterate(float lr) override {
        torch::NoGradGuard ng;
        const int64_t T = m_num_steps, N = m_num_envs;
        if (!m_started) { m_next_obs = initEnvs(); m_next_done.zero_(); m_started = true; }
        for (int64_t t = 0; t < T; t++) {
            m_obs[t].copy_(m_next_obs);                                               // :529-530
            m_dones[t].copy_(m_next_done);
            AgentOutput out = computeActionLogic(m_next_obs);                         // :533
            m_values[t].copy_(out.value.reshape({N}));
            m_actions[t].copy_(out.action.to(torch::kInt32).reshape({N, 1}));
            m_logprobs[t].copy_(out.logprob);
            auto [obs, reward, done] = stepEnvs(out.action);                          // :543
            m_rewards[t].copy_(reward.reshape({N}));
            m_next_obs = obs;
            m_next_done = done.to(torch::kFloat32).reshape({N});
        }
        // bootstrap value and the flag entering the next rollout (calcAdvantage :280), then everything behind the rollout
        int64_t bytes = 0;
        float* nv = static_cast<float*>(ppo_trainer_buffer(handle(), "next_value", &bytes));
        float* nd = static_cast<float*>(ppo_trainer_buffer(handle(), "next_done", &bytes));
        auto f32 = torch::TensorOptions().dtype(torch::kFloat32).device(*m_device);
        torch::from_blob(nv, {N}, f32).copy_(m_agent->getValue(m_next_obs).reshape({N}));
        torch::from_blob(nd, {N}, f32).copy_(m_next_done);
        cudaStream_t st = static_cast<cudaStream_t>(ppo_host::stream_of(m_next_obs));
        cudaStreamSynchronize(st);   // the rows above were written on torch's current stream
        ppo_iter_result r{};
        ppo_host::check_rc(ppo_trainer_update(handle(), lr, &r, ensureStream()), "ppo_trainer_update");
        noteLearningRate(lr);
        return r;
    }
    bool pushesEpisodeStats() const override { return true; }

    std::vector<std::shared_ptr<Env>> m_envs;                  // PPO_Discrete.h:106
    std::shared_ptr<ppo_host::ThreadPool> m_threadPool;        // :103

protected:
    int deviceEnvKind(const std::string&) const override { return PPO_ENV_HOST; }

private:
    torch::Tensor upload(const float* h, std::vector<int64_t> shape) {
        auto f32 = torch::TensorOptions().dtype(torch::kFloat32).device(*m_device);
        torch::Tensor d = torch::empty(shape, f32);
        cudaStream_t st = static_cast<cudaStream_t>(ppo_host::stream_of(d));
        cudaMemcpyAsync(d.data_ptr<float>(), h, (size_t)d.numel() * 4, cudaMemcpyHostToDevice, st);
        cudaStreamSynchronize(st);   // the staging buffer is reused by the next step
        return d;
    }
    float *m_h_obs = nullptr, *m_h_rew = nullptr, *m_h_done = nullptr;
    int64_t* m_h_act = nullptr;
    torch::Tensor m_next_obs, m_next_done;
    bool m_started = false;
};
