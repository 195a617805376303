// ORACLE — TEST INFRASTRUCTURE ONLY.  See common.h.
#include "common.h"

#include <cstdlib>

namespace oracle {

ThreadPool& ThreadPool::instance() {
  static ThreadPool pool;
  return pool;
}

// Default thread count follows src/threading.rs:41-63: RTEN_NUM_THREADS if set
// (clamped to the logical CPU count), else the machine's core count.
ThreadPool::ThreadPool() {
  int hw = (int)std::thread::hardware_concurrency();
  if (hw < 1) hw = 1;
  int n = hw;
  if (const char* env = std::getenv("RTEN_NUM_THREADS")) {
    int req = std::atoi(env);
    if (req >= 1) n = req > hw ? hw : req;
  }
  start(n);
}

ThreadPool::~ThreadPool() { stop(); }

void ThreadPool::start(int n) {
  n_threads_ = n < 1 ? 1 : n;
  quit_ = false;
  for (int i = 1; i < n_threads_; i++) threads_.emplace_back([this] { worker(); });
}

void ThreadPool::stop() {
  {
    std::lock_guard<std::mutex> lk(mu_);
    quit_ = true;
  }
  cv_.notify_all();
  for (auto& t : threads_) t.join();
  threads_.clear();
}

void ThreadPool::set_num_threads(int n) {
  stop();
  start(n);
}

bool ThreadPool::run_one(Job* job) {
  size_t i = job->next.fetch_add(1, std::memory_order_relaxed);
  if (i >= job->n) return false;
  (*job->fn)(i);
  job->done.fetch_add(1, std::memory_order_release);
  return true;
}

void ThreadPool::worker() {
  for (;;) {
    Job* job = nullptr;
    {
      std::unique_lock<std::mutex> lk(mu_);
      for (;;) {
        if (quit_) return;
        for (Job* j : jobs_) {
          if (j->next.load(std::memory_order_relaxed) < j->n) {
            job = j;
            break;
          }
        }
        if (job) break;
        cv_.wait(lk);
      }
      // Claim an index while still holding the lock so the job cannot be retired
      // (removed from jobs_ and destroyed) between selection and claim.
      size_t i = job->next.fetch_add(1, std::memory_order_relaxed);
      if (i >= job->n) continue;
      lk.unlock();
      (*job->fn)(i);
      job->done.fetch_add(1, std::memory_order_release);
    }
    // Keep draining this job without the lock: `done < n` until we finish the
    // indices we claim, so the owner cannot retire it under us.
    for (;;) {
      // Peek before claiming: claiming past n is harmless, but touching the job
      // after our last completed index is not, so re-validate through the lock.
      std::unique_lock<std::mutex> lk(mu_);
      bool alive = false;
      for (Job* j : jobs_) alive |= (j == job);
      if (!alive) break;
      size_t i = job->next.fetch_add(1, std::memory_order_relaxed);
      if (i >= job->n) break;
      lk.unlock();
      (*job->fn)(i);
      job->done.fetch_add(1, std::memory_order_release);
    }
  }
}

void ThreadPool::parallel_for(size_t n, const std::function<void(size_t)>& fn) {
  if (n == 0) return;
  if (n == 1 || n_threads_ == 1) {
    for (size_t i = 0; i < n; i++) fn(i);
    return;
  }
  Job job;
  job.fn = &fn;
  job.n = n;
  {
    std::lock_guard<std::mutex> lk(mu_);
    jobs_.push_back(&job);
  }
  cv_.notify_all();
  while (run_one(&job)) {
  }
  // Wait for helpers to finish the indices they claimed.
  while (job.done.load(std::memory_order_acquire) < n) std::this_thread::yield();
  {
    std::lock_guard<std::mutex> lk(mu_);
    for (size_t i = 0; i < jobs_.size(); i++) {
      if (jobs_[i] == &job) {
        jobs_.erase(jobs_.begin() + i);
        break;
      }
    }
  }
}

}  // namespace oracle
