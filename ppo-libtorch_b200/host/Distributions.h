// Distributions.h — the reference's distribution classes over the C ABI:
//   class Categorical        Distributions/Categorical.h:8-29       (Categorical.cpp:28-119)
//   class CategoricalMasked  Distributions/CategoricalMasked.h:9-30 (CategoricalMasked.cpp:31-144)
// Same constructors, methods and public members (m_logits, m_probs, m_masks, m_num_events, m_device).  Every number
// comes from a libppo_b200 kernel: ppo_categorical_normalize (constructor: m_logits = x - logsumexp, m_probs = softmax),
// ppo_categorical (sample / log_prob / entropy).  Inputs must be CUDA tensors — there is no CPU fallback.
// sample() draws from the library's counter-based Philox stream keyed by torch's default-generator seed
// (torch::manual_seed) and a process-wide call counter; LibTorch's multinomial stream (Categorical.cpp:77) is an
// exponential race on its own generator and cannot be reproduced by a counter-based one.
#pragma once
#include <torch/torch.h>

#include <memory>

class Categorical {
public:
    Categorical();
    Categorical(const torch::Tensor& logits, std::shared_ptr<torch::Device> device);
    ~Categorical();
    torch::Tensor logits_to_probs(torch::Tensor logits, bool is_binary = false);
    torch::Tensor sample();
    torch::Tensor log_prob(torch::Tensor value);
    torch::Tensor entropy();
    torch::Tensor mean();
    torch::Tensor mode();
    torch::Tensor variance();
    torch::Tensor enumerate_support();

    torch::Tensor m_logits, m_probs;
    int64_t m_num_events;
    std::shared_ptr<torch::Device> m_device;

private:
    torch::Tensor m_raw;   // [n, K] fp32 contiguous: what the kernels read
    std::vector<int64_t> m_batch_shape;
};

class CategoricalMasked {
public:
    CategoricalMasked();
    CategoricalMasked(const torch::Tensor& logits, const torch::Tensor& masks, std::shared_ptr<torch::Device> device);
    ~CategoricalMasked();
    torch::Tensor logits_to_probs(torch::Tensor logits, bool is_binary = false);
    torch::Tensor sample();
    torch::Tensor log_prob(torch::Tensor value);
    torch::Tensor entropy();
    torch::Tensor mean();
    torch::Tensor mode();
    torch::Tensor variance();
    torch::Tensor enumerate_support();

    torch::Tensor m_logits, m_probs, m_masks;
    int64_t m_num_events;
    std::shared_ptr<torch::Device> m_device;

private:
    torch::Tensor m_raw, m_mask8;   // [n, K] fp32 / u8 contiguous
    std::vector<int64_t> m_batch_shape;
};

namespace ppo_host {
// (seed, counter) of the next sampling call: seed = torch's default CPU generator seed, counter = process-wide
void next_sampling_key(uint64_t* seed, uint64_t* counter);
// throws std::runtime_error carrying ppo_last_error() when rc != 0
void check_rc(int rc, const char* what);
// current CUDA stream of the tensor's device as the ABI's stream handle
void* stream_of(const torch::Tensor& t);
}  // namespace ppo_host
