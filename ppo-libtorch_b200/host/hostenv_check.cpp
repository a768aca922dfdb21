// hostenv_check.cpp — the f4 bridge end to end: a CartPole written against the reference's duck-typed HOST interface
// (reset() / step(const int64_t&) / episode_length / episode_reward, Environments/CartPole.h:43-74; own code, std::mt19937
// per env) trained through PPO_DiscreteHostEnv<HostCartPole> from ./PPOConfig.toml.  Prints "ok ..." lines, exits 1 on failure.
#include <cmath>
#include <iostream>
#include <random>

#include "PPO_HostEnv.h"

struct HostCartPole {
    HostCartPole() : rng(next_seed()++) { reset(); }
    std::vector<float> reset() {
        std::uniform_real_distribution<float> d(-0.05f, 0.05f);
        for (float& v : s) v = d(rng);
        episode_length = 0;
        episode_reward = 0.0f;
        return {s[0], s[1], s[2], s[3]};
    }
    std::tuple<std::vector<float>, float, bool, bool> step(const int64_t& action) {
        const float force = action == 0 ? -10.0f : 10.0f, ct = std::cos(s[2]), sn = std::sin(s[2]);
        const float temp = (force + 0.05f * s[3] * s[3] * sn) / 1.1f;
        const float tacc = (9.8f * sn - ct * temp) / (0.5f * (4.0f / 3.0f - 0.1f * ct * ct / 1.1f));
        const float xacc = temp - 0.05f * tacc * ct / 1.1f;
        s[0] += 0.02f * s[1]; s[1] += 0.02f * xacc; s[2] += 0.02f * s[3]; s[3] += 0.02f * tacc;
        const bool term = std::fabs(s[0]) > 2.4f || std::fabs(s[2]) > 0.20943951f;
        const float r = term ? -1.0f : 1.0f;
        episode_length++;
        episode_reward += r;
        return {std::vector<float>{s[0], s[1], s[2], s[3]}, r, term, false};
    }
    int64_t episode_length = 0;
    float episode_reward = 0.0f;

private:
    static unsigned& next_seed() { static unsigned s = 1234; return s; }
    std::mt19937 rng;
    float s[4];
};

static void expect(bool c, const std::string& what) {
    if (!c) throw std::runtime_error("FAILED " + what);
    std::cout << "ok " << what << std::endl;
}

int main() {
    try {
        PPO_DiscreteHostEnv<HostCartPole> algo;
        const int64_t N = algo.m_num_envs;
        expect((int64_t)algo.m_envs.size() == N && algo.m_threadPool->size() >= 1, "m_envs / m_threadPool");
        torch::Tensor obs = algo.initEnvs();
        expect(obs.is_cuda() && obs.size(0) == N && obs.size(1) == 4 && obs.abs().max().item<float>() <= 0.05f, "initEnvs");
        AgentOutput out = algo.computeActionLogic(obs);
        auto [o2, r, d] = algo.stepEnvs(out.action);
        expect(o2.is_cuda() && r.size(0) == N && d.scalar_type() == torch::kInt32, "stepEnvs through the pinned staging buffers");
        expect(algo.m_envs[0]->episode_length == 1, "the host envs were stepped");
        algo.train();
        const double len = algo.m_episode_stats->avgLength();
        std::cout << "mean episode length after training: " << len << std::endl;
        expect(len > 60.0, "PPO_DiscreteHostEnv<HostCartPole>::train() learns (mean episode length of the last 100 episodes > 60)");
        std::cout << "hostenv_check: all ok" << std::endl;
    } catch (const std::exception& e) {
        std::cout << e.what() << std::endl;
        return 1;
    }
    return 0;
}
