// PPOUtils.h — host bookkeeping with the reference's semantics (Utils/Utils.h:13-79, Utils/Utils.cpp:5-60): the
// last-100-episodes ring behind ep_len_mean / ep_rew_mean, the mean of the per-minibatch clip fractions and the
// "step count inside a checkpoint file name" helper used by auto-resume.
#pragma once
#include <cctype>
#include <cstdint>
#include <string>
#include <vector>

namespace ppo_host {

// Ring of the most recent `capacity` finished episodes with running sums (Utils.h:30-79).
class CircularBuffer {
public:
    explicit CircularBuffer(size_t capacity) : rew_(capacity), len_(capacity), cap_(capacity) {}
    void add(float reward, int64_t length) {
        if (cap_ == 0) return;
        if (size_ == cap_) {
            rew_sum_ -= rew_[head_];
            len_sum_ -= (double)len_[head_];
        } else {
            size_++;
        }
        rew_[head_] = reward;
        len_[head_] = length;
        rew_sum_ += reward;
        len_sum_ += (double)length;
        head_ = (head_ + 1) % cap_;
    }
    bool empty() const { return size_ == 0; }
    size_t size() const { return size_; }
    float avgReward() const { return size_ ? (float)(rew_sum_ / (double)size_) : 0.0f; }
    double avgLength() const { return size_ ? len_sum_ / (double)size_ : 0.0; }

private:
    std::vector<float> rew_;
    std::vector<int64_t> len_;
    size_t cap_, size_ = 0, head_ = 0;
    double rew_sum_ = 0.0, len_sum_ = 0.0;
};

inline float vector_mean(const std::vector<float>& v) {   // Utils.cpp:5-14
    if (v.empty()) return 0.0f;
    double s = 0.0;
    for (float x : v) s += x;
    return (float)(s / (double)v.size());
}

// digits between `prefix` and the next '_' of the file name (".../PPO_Agent_<steps>_steps.pt"); 0 when absent
inline uint64_t step_from_filename(const std::string& file_name, const std::string& prefix) {
    const size_t p = file_name.rfind(prefix);
    if (p == std::string::npos) return 0;
    size_t i = p + prefix.size();
    uint64_t v = 0;
    bool any = false;
    while (i < file_name.size() && std::isdigit((unsigned char)file_name[i])) {
        v = v * 10 + (uint64_t)(file_name[i] - '0');
        any = true;
        i++;
    }
    if (!any || i >= file_name.size() || file_name[i] != '_') return 0;
    return v;
}

}  // namespace ppo_host
