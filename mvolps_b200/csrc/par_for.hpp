// par_for.hpp -- a few sleeping worker threads for the per-LP host loops of a B&B round (reading a
// launch's results back into the LP objects: every LP is independent).  Not OpenMP: its idle
// workers spin, and with one process per GPU eight ranks share the host cores.  The workers here
// wait on a condition variable and cost nothing between two calls.  One pool per process, shared by
// the engine and the driver (both are in the same library); created on first use (after the CLI's
// fork), sized by set_threads() (the B&B driver: min(4, cores / (2 x ranks))) or MVOLPS_HOST_THREADS, default min(4, cores / 2).
#pragma once

#include <algorithm>
#include <condition_variable>
#include <cstdlib>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

class ParPool {
  std::vector<std::thread> workers_;
  std::mutex mu_;
  std::condition_variable go_, done_;
  const std::function<void(size_t, size_t)> *fn_ = nullptr;
  size_t n_ = 0, chunk_ = 0, next_ = 0, pending_ = 0;
  unsigned long long epoch_ = 0;
  bool quit_ = false;
  int want_ = 0;

  void loop() {
    unsigned long long seen = 0;
    std::unique_lock<std::mutex> lk(mu_);
    while (true) {
      go_.wait(lk, [&] { return quit_ || epoch_ != seen; });
      if (quit_) return;
      seen = epoch_;
      while (next_ < n_) {
        const size_t a = next_, b = std::min(n_, a + chunk_);
        next_ = b;
        lk.unlock();
        (*fn_)(a, b);
        lk.lock();
      }
      if (--pending_ == 0) done_.notify_one();
    }
  }

 public:
  static ParPool &get() {
    static ParPool p;
    return p;
  }
  ~ParPool() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      quit_ = true;
    }
    go_.notify_all();
    for (auto &t : workers_) t.join();
  }
  // threads the next calls use (the caller's included); 0 = the default
  void set_threads(int t) { want_ = t; }
  int threads() {
    if (const char *e = getenv("MVOLPS_HOST_THREADS")) return std::max(1, atoi(e));
    if (want_ > 0) return want_;
    const int hw = (int)std::thread::hardware_concurrency();
    return std::max(1, std::min(4, hw / 2));
  }
  // fn(a, b) over disjoint ranges that cover [0, n); returns when all of them are done
  void run(size_t n, const std::function<void(size_t, size_t)> &fn, size_t min_chunk = 16) {
    const int T = threads();
    if (T <= 1 || n < 2 * min_chunk) { fn(0, n); return; }
    const int helpers = std::min<int>(T - 1, (int)(n / min_chunk) - 1);
    if (helpers < 1) { fn(0, n); return; }
    while ((int)workers_.size() < helpers) workers_.emplace_back([this] { loop(); });
    std::unique_lock<std::mutex> lk(mu_);
    fn_ = &fn;
    n_ = n;
    chunk_ = std::max(min_chunk, (n + 4 * (size_t)(helpers + 1) - 1) / (4 * (size_t)(helpers + 1)));
    next_ = 0;
    pending_ = workers_.size();
    epoch_++;
    go_.notify_all();
    while (next_ < n_) {  // the caller takes chunks as well
      const size_t a = next_, b = std::min(n_, a + chunk_);
      next_ = b;
      lk.unlock();
      fn(a, b);
      lk.lock();
    }
    done_.wait(lk, [&] { return pending_ == 0; });
    fn_ = nullptr;
  }
};

inline void par_for(size_t n, const std::function<void(size_t, size_t)> &fn, size_t min_chunk = 16) {
  ParPool::get().run(n, fn, min_chunk);
}
