// ThreadPool.h — the host-environment bridge's worker pool (the role of the reference's Utils/ThreadPool.h:14-75: one job
// per environment per step, a barrier before the results are used).  Written for the bridge: parallelFor(n, f) runs
// f(0) .. f(n - 1) on the workers in contiguous chunks and returns when all are done; an exception thrown by a job is
// rethrown on the caller's thread (the reference's pool swallows them).
#pragma once
#include <condition_variable>
#include <exception>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace ppo_host {

class ThreadPool {
public:
    explicit ThreadPool(unsigned threads = 0) {
        unsigned n = threads ? threads : std::thread::hardware_concurrency();
        if (n == 0) n = 1;
        for (unsigned i = 0; i < n; i++) workers_.emplace_back([this, i] { loop(i); });
    }
    ~ThreadPool() {
        {
            std::lock_guard<std::mutex> g(m_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto& w : workers_) w.join();
    }
    size_t size() const { return workers_.size(); }

    void parallelFor(int64_t n, const std::function<void(int64_t)>& f) {
        if (n <= 0) return;
        std::unique_lock<std::mutex> g(m_);
        job_ = &f; n_ = n; pending_ = (int)workers_.size(); error_ = nullptr; epoch_++;
        cv_.notify_all();
        done_.wait(g, [this] { return pending_ == 0; });
        job_ = nullptr;
        if (error_) std::rethrow_exception(error_);
    }

private:
    void loop(unsigned id) {
        uint64_t seen = 0;
        for (;;) {
            const std::function<void(int64_t)>* job;
            int64_t n;
            {
                std::unique_lock<std::mutex> g(m_);
                cv_.wait(g, [&] { return stop_ || epoch_ != seen; });
                if (stop_) return;
                seen = epoch_;
                job = job_;
                n = n_;
            }
            const int64_t w = (int64_t)workers_.size(), per = (n + w - 1) / w;
            const int64_t i0 = id * per, i1 = i0 + per < n ? i0 + per : n;
            std::exception_ptr err;
            try {
                for (int64_t i = i0; i < i1; i++) (*job)(i);
            } catch (...) {
                err = std::current_exception();
            }
            std::lock_guard<std::mutex> g(m_);
            if (err && !error_) error_ = err;
            if (--pending_ == 0) done_.notify_one();
        }
    }
    std::vector<std::thread> workers_;
    std::mutex m_;
    std::condition_variable cv_, done_;
    const std::function<void(int64_t)>* job_ = nullptr;
    int64_t n_ = 0;
    int pending_ = 0;
    uint64_t epoch_ = 0;
    bool stop_ = false;
    std::exception_ptr error_;
};

}  // namespace ppo_host
