// thread_barrier.h (facade) -- ThreadBarrier(n) / bool wait() of dryman/HashMergeJoin
// thread_barrier.h:23-35.  On the GPU path the barrier's job is done by kernel boundaries, the
// decoupled look-back and stream order; this host shim only keeps callers of the class compiling.
// wait() returns true for exactly one of the n arriving threads (the last one).
#ifndef THREAD_BARRIER_H
#define THREAD_BARRIER_H 1

#include <condition_variable>
#include <mutex>

class ThreadBarrier {
 public:
  explicit ThreadBarrier(unsigned int n) : expected_(n), arrived_(0), phase_(0) {}
  ThreadBarrier(const ThreadBarrier&) = delete;
  ThreadBarrier& operator=(const ThreadBarrier&) = delete;

  bool wait() {
    std::unique_lock<std::mutex> lock(m_);
    const unsigned long my_phase = phase_;
    if (++arrived_ == expected_) {
      arrived_ = 0;
      ++phase_;
      cv_.notify_all();
      return true;
    }
    cv_.wait(lock, [&] { return phase_ != my_phase; });
    return false;
  }

 private:
  std::mutex m_;
  std::condition_variable cv_;
  unsigned int expected_, arrived_;
  unsigned long phase_;
};

#endif  // THREAD_BARRIER_H
