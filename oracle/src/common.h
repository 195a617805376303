// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// CPU restatement of the RTen 0.12.0 dense hot path (reference: hsfzxjy/rten-fork).
// The reference is Rust and cannot be compiled in this image (no rustc/cargo), so
// this is a function-by-function C++ restatement; every function cites the
// reference file:line it follows.  Parity is pinned by the reference's own golden
// vectors and known-answer tests (see tests/test_oracle_*.py, tests/golden/).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load this library.  The product path (rten_fork_b200/) never does.
#pragma once
#include <atomic>
#include <condition_variable>
#include <cstddef>
#include <cstdint>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace oracle {

// ---------------------------------------------------------------------------
// Strided 2-D view.  Mirrors rten-tensor `Matrix` (ptr + [rows, cols] + element
// strides) as used by src/gemm.rs.
// ---------------------------------------------------------------------------
struct Matrix {
  const float* data;
  size_t rows, cols;
  ptrdiff_t row_stride, col_stride;
  float at(size_t r, size_t c) const { return data[(ptrdiff_t)r * row_stride + (ptrdiff_t)c * col_stride]; }
};

// ---------------------------------------------------------------------------
// Work-sharing pool standing in for Rayon (src/threading.rs:18-63).  Supports
// nested parallel_for the way the reference nests col-block / row-block /
// per-image parallelism: the caller always executes indices of its own job, idle
// workers help with any open job.
// ---------------------------------------------------------------------------
class ThreadPool {
 public:
  static ThreadPool& instance();
  // (Re)create the pool with `n` threads in total (caller thread included).
  void set_num_threads(int n);
  int num_threads() const { return n_threads_; }
  void parallel_for(size_t n, const std::function<void(size_t)>& fn);

 private:
  struct Job {
    const std::function<void(size_t)>* fn;
    size_t n;
    std::atomic<size_t> next{0};
    std::atomic<size_t> done{0};
  };
  ThreadPool();
  ~ThreadPool();
  void start(int n);
  void stop();
  void worker();
  bool run_one(Job* job);

  int n_threads_ = 1;
  std::vector<std::thread> threads_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::vector<Job*> jobs_;
  bool quit_ = false;
};

inline size_t div_ceil(size_t a, size_t b) { return (a + b - 1) / b; }
inline size_t round_up(size_t a, size_t b) { return div_ceil(a, b) * b; }

}  // namespace oracle
