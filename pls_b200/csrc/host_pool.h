// pls_b200/csrc/host_pool.h — a tiny persistent parallel-for pool of host threads (packing pageable host matrices
// into pinned staging buffers for the H2D copy engine; dataset.cu).
#pragma once
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace plsb {

class HostPool {
  public:
    explicit HostPool(int nthreads) : n_(nthreads < 1 ? 1 : nthreads) {
        for (int t = 1; t < n_; ++t) th_.emplace_back([this, t] { loop(t); });
    }
    ~HostPool() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
            ++epoch_;
        }
        cv_.notify_all();
        for (auto &t : th_) t.join();
    }
    int size() const { return n_; }
    // run fn(t, n) for t = 0..n-1 (the caller is thread 0) and wait for all of them
    void run(const std::function<void(int, int)> &fn) {
        {
            std::lock_guard<std::mutex> lk(mu_);
            fn_ = &fn;
            pending_ = n_ - 1;
            ++epoch_;
        }
        cv_.notify_all();
        fn(0, n_);
        std::unique_lock<std::mutex> lk(mu_);
        done_.wait(lk, [this] { return pending_ == 0; });
        fn_ = nullptr;
    }

  private:
    void loop(int t) {
        unsigned long seen = 0;
        for (;;) {
            const std::function<void(int, int)> *fn;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return epoch_ != seen; });
                seen = epoch_;
                if (stop_) return;
                fn = fn_;
            }
            (*fn)(t, n_);
            {
                std::lock_guard<std::mutex> lk(mu_);
                --pending_;
            }
            done_.notify_one();
        }
    }
    int n_;
    std::vector<std::thread> th_;
    std::mutex mu_;
    std::condition_variable cv_, done_;
    const std::function<void(int, int)> *fn_ = nullptr;
    int pending_ = 0;
    unsigned long epoch_ = 0;
    bool stop_ = false;
};

}  // namespace plsb
