// engine_util.h -- small helpers shared by the native engines (engine.cu, engine_coll.cu).
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <condition_variable>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../include/gmt_b200.h"
#include "check.h"

namespace gmt {

// the engines hop between devices; callers (e.g. a torch process) get their current device back
struct DeviceRestore {
  int dev = -1;
  DeviceRestore()
  {
    if (cudaGetDevice(&dev) != cudaSuccess) dev = -1;
  }
  ~DeviceRestore()
  {
    if (dev >= 0) cudaSetDevice(dev);
  }
};

// bump allocator over one zero-initialised block, 256-byte granules
inline size_t take(size_t& cursor, size_t bytes)
{
  const size_t at = cursor;
  cursor += (bytes + 255) & ~size_t(255);
  return at;
}

// ranks grouped by the device that hosts them, in first-seen order
struct DeviceRanks {
  int device = 0;
  std::vector<int> ranks;
};

inline std::vector<DeviceRanks> group_by_device(const std::vector<int>& device_of_rank)
{
  std::vector<DeviceRanks> groups;
  for (int r = 0; r < (int)device_of_rank.size(); ++r) {
    size_t k = 0;
    while (k < groups.size() && groups[k].device != device_of_rank[r]) ++k;
    if (k == groups.size()) {
      groups.emplace_back();
      groups.back().device = device_of_rank[r];
    }
    groups[k].ranks.push_back(r);
  }
  return groups;
}

// One worker thread per GPU: a collective (or a step) is one launch per GPU, and launches issued
// one after the other from a single host thread skew the GPUs by ~4-5 us each -- for the
// latency-bound collectives that skew IS the cost (daxpy + global sum on 4 GPUs: 16.9 us per call
// from one thread vs 10.7 us with one process per GPU).  Workers bind their device once, spin
// briefly for the next job (back-to-back calls hand over in about a microsecond) and fall asleep
// on a condition variable when the engine goes idle.  A single-GPU engine runs inline, no thread.
// GMT_ENGINE_THREADS=0 forces the sequential loop.
class DeviceExecutor {
 public:
  explicit DeviceExecutor(const std::vector<int>& devices) : devices_(devices), rc_(devices.size(), 0),
                                                            msg_(devices.size())
  {
    const char* env = getenv("GMT_ENGINE_THREADS");
    threaded_ = devices.size() > 1 && !(env && env[0] == '0');
    if (!threaded_) return;
    for (size_t i = 1; i < devices.size(); ++i) workers_.emplace_back([this, i] { loop(i); });
  }
  ~DeviceExecutor()
  {
    if (!threaded_) return;
    {
      std::lock_guard<std::mutex> lock(mu_);
      stop_ = true;
      ticket_.fetch_add(1, std::memory_order_release);
    }
    cv_.notify_all();
    for (auto& t : workers_) t.join();
  }
  DeviceExecutor(const DeviceExecutor&) = delete;
  DeviceExecutor& operator=(const DeviceExecutor&) = delete;

  // fn(i) runs with device i of the list current, on all devices at once; the calling thread
  // takes index 0.  Returns the first non-zero code (its message copied into the caller's buffer).
  int run(const std::function<int(int)>& fn)
  {
    const int n = (int)devices_.size();
    if (!threaded_) {
      for (int i = 0; i < n; ++i) {
        if (cudaSetDevice(devices_[i]) != cudaSuccess) return fail(GMT_EINVAL, "cudaSetDevice(%d) failed", devices_[i]);
        const int rc = fn(i);
        if (rc) return rc;
      }
      return 0;
    }
    fn_ = &fn;
    pending_.store(n - 1, std::memory_order_relaxed);
    {
      std::lock_guard<std::mutex> lock(mu_);   // pairs with a worker that is about to sleep
      ticket_.fetch_add(1, std::memory_order_release);
    }
    if (sleepers_.load(std::memory_order_acquire) > 0) cv_.notify_all();
    int rc0 = cudaSetDevice(devices_[0]) == cudaSuccess ? fn(0) : fail(GMT_EINVAL, "cudaSetDevice failed");
    while (pending_.load(std::memory_order_acquire) > 0) std::this_thread::yield();
    if (rc0) return rc0;
    for (int i = 1; i < n; ++i)
      if (rc_[i]) {
        strncpy(err_buf(), msg_[i].c_str(), 511);
        return rc_[i];
      }
    return 0;
  }

 private:
  void loop(size_t i)
  {
    (void)cudaSetDevice(devices_[i]);
    uint64_t seen = 0;
    for (;;) {
      int spins = 0;
      while (ticket_.load(std::memory_order_acquire) == seen) {
        if (++spins < 20000) {
          std::this_thread::yield();
          continue;
        }
        std::unique_lock<std::mutex> lock(mu_);
        sleepers_.fetch_add(1, std::memory_order_release);
        cv_.wait(lock, [&] { return ticket_.load(std::memory_order_acquire) != seen; });
        sleepers_.fetch_sub(1, std::memory_order_release);
      }
      seen = ticket_.load(std::memory_order_acquire);
      if (stop_) return;
      rc_[i] = (*fn_)((int)i);
      if (rc_[i]) msg_[i] = err_buf();
      pending_.fetch_sub(1, std::memory_order_release);
    }
  }

  std::vector<int> devices_;
  std::vector<int> rc_;
  std::vector<std::string> msg_;
  std::vector<std::thread> workers_;
  const std::function<int(int)>* fn_ = nullptr;
  std::atomic<uint64_t> ticket_{0};
  std::atomic<int> pending_{0}, sleepers_{0};
  std::mutex mu_;
  std::condition_variable cv_;
  bool threaded_ = false, stop_ = false;
};

}  // namespace gmt
