// Host worker pool for the Fiat-Shamir tail.  The sum-check and bullet rounds are strictly
// sequential (every challenge depends on the previous commitments), so what is left to
// parallelise on the host are the independent fixed-base scalar multiplications and point
// compressions INSIDE one round (sumcheck.rs:659-706, nizk/mod.rs:333-345), and the
// commitments whose scalars come from the RandomTape alone, which can be computed ahead of
// the rounds while the GPU is busy with the witness commitment.
//
// One mailbox per worker, tasks assigned explicitly by the (single) posting thread: no shared
// queue, no stragglers.  Workers spin for a short while after their last task and then sleep
// on a condition variable, so an idle library costs nothing between proofs.
#pragma once
#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>
#if defined(__linux__)
#include <pthread.h>
#include <sched.h>
#endif

#if defined(__x86_64__)
#include <immintrin.h>
#define SB_CPU_RELAX() _mm_pause()
#else
#define SB_CPU_RELAX() std::this_thread::yield()
#endif

namespace sbh {

class HostPool {
  struct alignas(64) Slot {
    std::function<void()> fn;
    std::atomic<uint64_t> posted{0}, done{0};
    std::atomic<bool> sleeping{false};
    std::mutex mu;
    std::condition_variable cv;
    std::thread th;
  };
  std::vector<std::unique_ptr<Slot>> slots_;
  std::atomic<bool> stop_{false};

  void worker(Slot* s) {
    uint64_t seen = 0;
    for (;;) {
      uint64_t p = 0;
      bool got = false;
      for (int spin = 0; spin < 200000; spin++) {  // ~ a few ms
        p = s->posted.load(std::memory_order_acquire);
        if (p != seen || stop_.load(std::memory_order_relaxed)) { got = true; break; }
        SB_CPU_RELAX();
      }
      if (!got) {
        std::unique_lock<std::mutex> lk(s->mu);
        s->sleeping.store(true, std::memory_order_seq_cst);
        s->cv.wait(lk, [&] { return s->posted.load(std::memory_order_seq_cst) != seen || stop_.load(); });
        s->sleeping.store(false, std::memory_order_seq_cst);
        p = s->posted.load(std::memory_order_acquire);
      }
      if (stop_.load()) return;
      if (p == seen) continue;
      s->fn();
      seen = p;
      s->done.store(p, std::memory_order_release);
    }
  }

 public:
  // SB200_PIN_WORKERS=1: each worker on a core of its own (the k-th CPU of the set this process may use, the first left
  // to the posting thread).  Off by default: a 2^16 proof comes out at 2.06 - 2.30 ms from process to process either way
  // (three alternating runs each on a B200 box), so the spread is not the scheduler moving the workers.
  explicit HostPool(int n) {
    std::vector<int> cpus;
#if defined(__linux__)
    const char* pe = getenv("SB200_PIN_WORKERS");
    if (pe && atoi(pe) != 0) {
      cpu_set_t cur;
      CPU_ZERO(&cur);
      if (sched_getaffinity(0, sizeof cur, &cur) == 0)
        for (int c = 0; c < CPU_SETSIZE; c++)
          if (CPU_ISSET(c, &cur)) cpus.push_back(c);
      if ((int)cpus.size() < n + 1) cpus.clear();  // fewer cores than threads: pinning would stack workers on the caller's core
    }
#endif
    for (int i = 0; i < n; i++) {
      slots_.emplace_back(new Slot());
      Slot* s = slots_.back().get();
      s->th = std::thread([this, s] { worker(s); });
#if defined(__linux__)
      if (!cpus.empty()) {
        cpu_set_t one;
        CPU_ZERO(&one);
        CPU_SET(cpus[(size_t)i + 1], &one);
        pthread_setaffinity_np(s->th.native_handle(), sizeof one, &one);
      }
#endif
    }
  }
  ~HostPool() {
    stop_.store(true);
    for (auto& s : slots_) {
      { std::lock_guard<std::mutex> lk(s->mu); }
      s->cv.notify_all();
      if (s->th.joinable()) s->th.join();
    }
  }
  // SB200_HOST_THREADS overrides the worker count (0 = everything on the calling thread).
  static HostPool& get() {
    static HostPool* pool = [] {
      int hw = (int)std::thread::hardware_concurrency();
      if (const char* lw = getenv("LOCAL_WORLD_SIZE")) {  // one process per GPU (torchrun): share the host cores
        int ranks = atoi(lw);
        if (ranks > 1) hw /= ranks;
      }
      int n = hw > 1 ? hw - 1 : 0;  // this thread + n workers = one thread per core (the last worker draws the tape first)
      if (n > 10) n = 10;
      if (const char* e = getenv("SB200_HOST_THREADS")) n = atoi(e);
      if (n < 0) n = 0;
      if (n > 32) n = 32;
      return new HostPool(n);  // lives for the process: worker threads must not be joined from a static destructor
    }();
    return *pool;
  }
  int workers() const { return (int)slots_.size(); }

  // Hand `fn` to worker w (which must be idle: every earlier post to it has been wait()ed for).
  void post(int w, std::function<void()> fn) {
    Slot* s = slots_[w].get();
    s->fn = std::move(fn);
    s->posted.store(s->posted.load(std::memory_order_relaxed) + 1, std::memory_order_seq_cst);
    if (s->sleeping.load(std::memory_order_seq_cst)) {
      { std::lock_guard<std::mutex> lk(s->mu); }
      s->cv.notify_one();
    }
  }
  void wait(int w) {
    Slot* s = slots_[w].get();
    uint64_t p = s->posted.load(std::memory_order_relaxed);
    while (s->done.load(std::memory_order_acquire) != p) SB_CPU_RELAX();
  }
  bool idle(int w) const {
    const Slot* s = slots_[w].get();
    return s->done.load(std::memory_order_acquire) == s->posted.load(std::memory_order_relaxed);
  }
  // f(0..n) spread over the caller and workers [0, use); returns when all are done.
  template <typename F>
  void parallel(size_t n, int use, F&& f) {
    if (use > workers()) use = workers();
    if (n <= 1 || use <= 0) {
      for (size_t i = 0; i < n; i++) f(i);
      return;
    }
    const size_t parts = (size_t)use + 1;
    int posted = 0;
    for (int w = 0; w < use; w++) {
      size_t first = (size_t)w + 1;  // the caller takes i = 0, parts, 2*parts, ...
      if (first >= n) break;
      post(w, [&f, first, n, parts] { for (size_t i = first; i < n; i += parts) f(i); });
      posted++;
    }
    for (size_t i = 0; i < n; i += parts) f(i);
    for (int w = 0; w < posted; w++) wait(w);
  }
};

}  // namespace sbh
