// Host side of the input pipeline: fp32 -> bf16 conversion of a batch on CPU threads, into pinned memory.
//
// The native engine computes in bf16, so the first thing that happens to an fp32 input batch on the device is a cast.
// Doing that cast on the host - same round-to-nearest-even, bit-identical result - halves the bytes every step sends
// over PCIe. With eight GPUs streaming 67 MB batches concurrently the host links, not the GPUs, bound the end-to-end
// step (profiles/e2e_trace_8gpu_r1.log), so this is where the time is. A small persistent thread pool does the work;
// the call blocks until the batch is converted (the Python caller runs it on a helper thread, ctypes drops the GIL).
//
// No counterpart in the reference: it ships pickled NumPy arrays through Spark to the executors
// (elephas/utils/rdd_utils.py:12-27, elephas/worker.py:36-37).
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#include "../include/elephas_b200.h"
#include "host.h"

namespace eb {

// Row conversion with explicit AVX-512 / AVX2 intrinsics and non-temporal stores: host_cast.cpp
void convert_rows(const float* src, int64_t ld_src, uint16_t* dst, int64_t ld_dst, int64_t r0, int64_t r1, int32_t cols);
const char* host_cast_isa();

// Persistent worker pool: run(fn, n) executes fn(i) for i in [0, n) on the workers plus the calling thread.
// All job state (fn_, total_, next_) is written under m_ while no worker is inside drain(): a worker enters drain() only
// after it has seen a live job under m_ (active_ counts the workers inside), and run() returns only when every item is
// done AND active_ is back to zero. (Checked with ThreadSanitizer: tests/native/tsan_host_pool.cpp.)
class HostPool {
 public:
  explicit HostPool(int workers) {
    for (int i = 0; i < workers; ++i) threads_.emplace_back([this] { loop(); });
  }
  ~HostPool() {
    {
      std::lock_guard<std::mutex> g(m_);
      stop_ = true;
    }
    cv_.notify_all();
    for (auto& t : threads_) t.join();
  }
  int workers() const { return static_cast<int>(threads_.size()); }

  void run(const std::function<void(int)>& fn, int n) {
    std::lock_guard<std::mutex> one_job(job_mutex_);  // one batch at a time
    {
      std::lock_guard<std::mutex> g(m_);
      fn_ = &fn;
      total_ = n;
      pending_ = n;
      next_.store(0);
      ++generation_;
    }
    cv_.notify_all();
    drain();
    std::unique_lock<std::mutex> g(m_);
    done_cv_.wait(g, [this] { return pending_ == 0 && active_ == 0; });
    fn_ = nullptr;  // a worker that wakes up late finds no job
  }

 private:
  void drain() {
    for (;;) {
      const int i = next_.fetch_add(1);
      if (i >= total_) return;
      (*fn_)(i);
      std::lock_guard<std::mutex> g(m_);
      if (--pending_ == 0 && active_ == 0) done_cv_.notify_all();
    }
  }
  void loop() {
    uint64_t seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> g(m_);
        cv_.wait(g, [&] { return stop_ || generation_ != seen; });
        if (stop_) return;
        seen = generation_;
        if (fn_ == nullptr) continue;  // that job is over already
        ++active_;
      }
      drain();
      std::lock_guard<std::mutex> g(m_);
      if (--active_ == 0 && pending_ == 0) done_cv_.notify_all();
    }
  }
  std::vector<std::thread> threads_;
  std::mutex m_, job_mutex_;
  std::condition_variable cv_, done_cv_;
  const std::function<void(int)>* fn_ = nullptr;
  std::atomic<int> next_{0};
  int total_ = 0, pending_ = 0, active_ = 0;
  uint64_t generation_ = 0;
  bool stop_ = false;
};

static HostPool* g_pool = nullptr;
static std::mutex g_pool_mutex;

// One pool per process, grown when a caller asks for more threads than it has. Jobs are serialised by g_pool_mutex
// (one batch at a time anyway), so a pool is never replaced while a job runs on it.
static HostPool* pool_for_locked(int threads) {  // caller holds g_pool_mutex (also while it runs its job on the pool)
  const int workers = threads > 1 ? threads - 1 : 0;
  if (g_pool == nullptr) {
    g_pool = new HostPool(workers);
  } else if (g_pool->workers() < workers) {
    HostPool* old = g_pool;
    g_pool = new HostPool(workers);
    delete old;  // idle: joins its workers
  }
  return g_pool;
}

}  // namespace eb

extern "C" int eb_host_f32_to_bf16(const float* src, int64_t ld_src, void* dst, int64_t ld_dst, int64_t rows,
                                   int32_t cols, int32_t threads) {
  using namespace eb;
  if (rows <= 0 || cols <= 0) return 0;
  if (src == nullptr || dst == nullptr) return fail("eb_host_f32_to_bf16: null buffer");
  if (ld_src < cols || ld_dst < cols) return fail("eb_host_f32_to_bf16: leading dimensions smaller than cols");
  if (threads < 1) threads = 1;
  if (threads > 64) threads = 64;
  uint16_t* out = static_cast<uint16_t*>(dst);
  const int64_t chunk = 64;  // rows per work item
  const int items = static_cast<int>((rows + chunk - 1) / chunk);
  if (threads == 1 || items == 1) {
    convert_rows(src, ld_src, out, ld_dst, 0, rows, cols);
    return 0;
  }
  std::function<void(int)> fn = [&](int i) {
    const int64_t r0 = i * chunk, r1 = r0 + chunk < rows ? r0 + chunk : rows;
    convert_rows(src, ld_src, out, ld_dst, r0, r1, cols);
  };
  std::lock_guard<std::mutex> g(g_pool_mutex);
  pool_for_locked(threads)->run(fn, items);
  return 0;
}

extern "C" const char* eb_host_cast_isa() { return eb::host_cast_isa(); }
