// A small persistent pool of host threads for the staging layer: page-cache reads, gathers of trimmed
// records and bounce-buffer copies are all "n independent chunks" jobs that run once per 64 MiB tile;
// creating 16 threads per tile costs as much as the tile's PCIe time.
#pragma once
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace pdx {

class HostPool {
 public:
  static HostPool& Get() {
    static HostPool pool;
    return pool;
  }
  uint32_t threads() const { return static_cast<uint32_t>(workers_.size()) + 1; }

  // Runs fn(i) for every i in [0, n); the caller works too.  Jobs are serialised (one at a time).
  void ParallelFor(uint64_t n, const std::function<void(uint64_t)>& fn) {
    if (n == 0) return;
    if (n == 1 || workers_.empty()) {
      for (uint64_t i = 0; i < n; ++i) fn(i);
      return;
    }
    std::lock_guard<std::mutex> job_lock(job_mu_);
    {
      std::lock_guard<std::mutex> lk(mu_);
      fn_ = &fn;
      n_ = n;
      next_.store(0, std::memory_order_relaxed);
      pending_ = static_cast<uint32_t>(workers_.size());
      ++generation_;
    }
    cv_.notify_all();
    Work();
    std::unique_lock<std::mutex> lk(mu_);
    done_cv_.wait(lk, [&] { return pending_ == 0; });
    fn_ = nullptr;
  }

 private:
  HostPool() {
    const uint32_t hw = std::max(2u, std::min(64u, std::thread::hardware_concurrency()));
    for (uint32_t t = 1; t < hw; ++t) workers_.emplace_back([this] { Loop(); });
  }
  ~HostPool() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      stop_ = true;
      ++generation_;
    }
    cv_.notify_all();
    for (auto& w : workers_) w.join();
  }
  void Work() {
    for (;;) {
      const uint64_t i = next_.fetch_add(1, std::memory_order_relaxed);
      if (i >= n_) break;
      (*fn_)(i);
    }
  }
  void Loop() {
    uint64_t seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return generation_ != seen; });
        seen = generation_;
        if (stop_) return;
      }
      Work();
      {
        std::lock_guard<std::mutex> lk(mu_);
        if (--pending_ == 0) done_cv_.notify_one();
      }
    }
  }

  std::vector<std::thread> workers_;
  std::mutex job_mu_, mu_;
  std::condition_variable cv_, done_cv_;
  const std::function<void(uint64_t)>* fn_ = nullptr;
  uint64_t n_ = 0;
  std::atomic<uint64_t> next_{0};
  uint32_t pending_ = 0;
  uint64_t generation_ = 0;
  bool stop_ = false;
};

}  // namespace pdx
