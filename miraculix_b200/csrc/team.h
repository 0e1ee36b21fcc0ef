// team.h -- one host thread per device of a single-process multi-GPU call (multi.cu,
// crossprod_host.cu).  Member 0 is the calling thread; members 1..n-1 are persistent threads that
// spin briefly for the next job (iterative solvers call back to back) before they sleep.
#pragma once
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>
#include "common.cuh"

namespace mb200 {

class Team {
 public:
  explicit Team(int n) : n_(n), rc_(n, 0), st_(n) {
    for (int g = 1; g < n_; g++) threads_.emplace_back([this, g] { loop(g); });
  }
  ~Team() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      quit_.store(true, std::memory_order_release);
      gen_.fetch_add(1, std::memory_order_release);
    }
    cv_.notify_all();
    for (auto &t : threads_) t.join();
  }
  int size() const { return n_; }

  // Runs job(g) on every member and returns the first non-zero result; that member's error text
  // becomes the calling thread's status.  A member that fails releases everybody else from the
  // barriers of this job (they return non-zero too).
  int run(const std::function<int(int)> &job) {
    failed_.store(0, std::memory_order_relaxed);
    bar_count_.store(0, std::memory_order_relaxed);
    job_ = &job;
    done_.store(0, std::memory_order_relaxed);
    {
      std::lock_guard<std::mutex> lk(mu_);
      gen_.fetch_add(1, std::memory_order_release);
    }
    if (sleepers_.load(std::memory_order_acquire) > 0) cv_.notify_all();
    execute(0);
    int spins = 0;
    while (done_.load(std::memory_order_acquire) != n_ - 1)
      if (++spins > 2000) std::this_thread::yield();
    job_ = nullptr;
    for (int g = 0; g < n_; g++)
      if (rc_[g] != 0) {
        if (g != 0) status() = st_[g];
        return rc_[g];
      }
    return 0;
  }

  // All members of the running job meet here.  Returns non-zero when some member has failed (the
  // caller should unwind: nothing after a failed barrier may depend on the peers).
  int barrier() {
    if (n_ == 1) return 0;
    const int gen = bar_gen_.load(std::memory_order_acquire);
    if (bar_count_.fetch_add(1, std::memory_order_acq_rel) + 1 == n_) {
      bar_count_.store(0, std::memory_order_relaxed);
      bar_gen_.fetch_add(1, std::memory_order_release);
    } else {
      int spins = 0;
      while (bar_gen_.load(std::memory_order_acquire) == gen) {
        if (failed_.load(std::memory_order_acquire)) return -40;
        if (++spins > 4000) std::this_thread::yield();
      }
    }
    return failed_.load(std::memory_order_acquire) ? -40 : 0;
  }
  void mark_failed() { failed_.store(1, std::memory_order_release); }

 private:
  void execute(int g) {
    clear_status();
    int rc = 0;
    try {
      rc = (*job_)(g);
    } catch (...) {
      rc = fail(-41, __func__, __LINE__, "exception in device worker %d", g);
    }
    if (rc != 0) {
      mark_failed();
      if (rc == -40 && status().code == 0) fail(-40, __func__, __LINE__, "aborted: a peer device worker failed");
      st_[g] = status();
    }
    rc_[g] = rc;
  }
  void loop(int g) {
    unsigned long long seen = 0;
    for (;;) {
      // spin for up to ~2 ms for the next job, then sleep on the condition variable
      const auto t0 = std::chrono::steady_clock::now();
      int spins = 0;
      while (gen_.load(std::memory_order_acquire) == seen) {
        if (++spins % 256 == 0) {
          if (std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(2)) {
            std::unique_lock<std::mutex> lk(mu_);
            sleepers_.fetch_add(1, std::memory_order_release);
            cv_.wait(lk, [&] { return gen_.load(std::memory_order_acquire) != seen; });
            sleepers_.fetch_sub(1, std::memory_order_release);
            break;
          }
        }
      }
      seen = gen_.load(std::memory_order_acquire);
      if (quit_.load(std::memory_order_acquire)) return;
      execute(g);
      done_.fetch_add(1, std::memory_order_release);
    }
  }

  const int n_;
  std::vector<std::thread> threads_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::atomic<unsigned long long> gen_{0};
  std::atomic<int> sleepers_{0}, done_{0}, failed_{0}, bar_count_{0}, bar_gen_{0};
  const std::function<int(int)> *job_ = nullptr;
  std::atomic<bool> quit_{false};
  std::vector<int> rc_;
  std::vector<Status> st_;
};

}  // namespace mb200
