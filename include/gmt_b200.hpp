// gmt_b200.hpp -- thin C++ conveniences over the C ABI (gmt_b200.h); header-only.
//
//   gmt::check(rc, "what")     throws gmt::Error carrying the code and gmt_last_error()
//   gmt::NvtxRange r("name")   RAII push/pop range (the reference pushes and pops by hand around
//                              every phase, mpi_daxpy_nvtx.cc:177-313; a guard cannot leak a range
//                              on an early return)
//   gmt::ProfilerWindow w      RAII cudaProfilerStart/Stop (daxpy_nvtx.cu:65,105)
//   gmt::DeviceBuffer<T>       owning device / managed / pinned allocation (gmt_alloc / gmt_free)
//   gmt::Engine2D, gmt::CollEngine   owning handles on the native engines
#ifndef GMT_B200_HPP
#define GMT_B200_HPP

#include <cstddef>
#include <stdexcept>
#include <string>
#include <utility>

#include "gmt_b200.h"

namespace gmt {

class Error : public std::runtime_error {
 public:
  Error(int code, const std::string& what) : std::runtime_error(what), code_(code) {}
  int code() const { return code_; }

 private:
  int code_;
};

inline void check(int rc, const char* what)
{
  if (rc != 0) throw Error(rc, std::string(what) + ": error " + std::to_string(rc) + ": " + gmt_last_error());
}

class NvtxRange {
 public:
  explicit NvtxRange(const char* name) { gmt_nvtx_push(name); }
  ~NvtxRange() { gmt_nvtx_pop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

class ProfilerWindow {
 public:
  ProfilerWindow() { gmt_profiler_start(); }
  ~ProfilerWindow() { gmt_profiler_stop(); }
  ProfilerWindow(const ProfilerWindow&) = delete;
  ProfilerWindow& operator=(const ProfilerWindow&) = delete;
};

template <class T>
class DeviceBuffer {
 public:
  DeviceBuffer() = default;
  DeviceBuffer(std::size_t count, int space = GMT_SPACE_DEVICE) : space_(space), count_(count)
  {
    check(gmt_alloc(space, count * sizeof(T), reinterpret_cast<void**>(&ptr_)), "gmt_alloc");
  }
  ~DeviceBuffer() { reset(); }
  DeviceBuffer(DeviceBuffer&& o) noexcept : ptr_(o.ptr_), space_(o.space_), count_(o.count_) { o.ptr_ = nullptr; }
  DeviceBuffer& operator=(DeviceBuffer&& o) noexcept
  {
    if (this != &o) {
      reset();
      ptr_ = o.ptr_, space_ = o.space_, count_ = o.count_;
      o.ptr_ = nullptr;
    }
    return *this;
  }
  DeviceBuffer(const DeviceBuffer&) = delete;
  DeviceBuffer& operator=(const DeviceBuffer&) = delete;
  void reset()
  {
    if (ptr_) gmt_free(space_, ptr_);
    ptr_ = nullptr;
  }
  T* get() const { return ptr_; }
  std::size_t size() const { return count_; }
  std::size_t bytes() const { return count_ * sizeof(T); }
  T& operator[](std::size_t i) const { return ptr_[i]; }   // host-addressable spaces only

 private:
  T* ptr_ = nullptr;
  int space_ = GMT_SPACE_DEVICE;
  std::size_t count_ = 0;
};

class Engine2D {
 public:
  Engine2D(int px, int py, int64_t nx, int64_t ny, int gx = 2, int gy = 2, bool with_err = true,
           const int* devices = nullptr, int64_t spin_limit = 0)
  {
    check(gmt_engine2d_create(px, py, devices, nx, ny, gx, gy, with_err, spin_limit, &e_), "gmt_engine2d_create");
  }
  ~Engine2D() { gmt_engine2d_destroy(e_); }
  Engine2D(const Engine2D&) = delete;
  Engine2D& operator=(const Engine2D&) = delete;
  gmt_engine2d* get() const { return e_; }
  int world() const { return gmt_engine2d_world(e_); }
  int device_groups() const { return gmt_engine2d_device_groups(e_); }
  void init_analytic(double poison = 0.0) { check(gmt_engine2d_init_analytic(e_, poison), "init_analytic"); }
  void step() { check(gmt_engine2d_step(e_), "step"); }
  void exchange() { check(gmt_engine2d_exchange(e_), "exchange"); }
  void synchronize() { check(gmt_engine2d_synchronize(e_), "synchronize"); }
  void status() { check(gmt_engine2d_status(e_, nullptr), "status"); }
  double time_steps(int steps, int warmup)
  {
    double ms = 0.0;
    check(gmt_engine2d_time_steps(e_, steps, warmup, &ms), "time_steps");
    return ms;
  }
  double time_exchanges(int steps, int warmup)
  {
    double ms = 0.0;
    check(gmt_engine2d_time_exchanges(e_, steps, warmup, &ms), "time_exchanges");
    return ms;
  }
  double err_norm(int rank)
  {
    double v = 0.0;
    check(gmt_engine2d_err_norm(e_, rank, &v), "err_norm");
    return v;
  }

 private:
  gmt_engine2d* e_ = nullptr;
};

class CollEngine {
 public:
  CollEngine(int world, int64_t buf_doubles, int64_t stage_cap = 4096, const int* devices = nullptr,
             int64_t spin_limit = 0)
  {
    check(gmt_coll_engine_create(world, devices, buf_doubles, stage_cap, spin_limit, &e_), "gmt_coll_engine_create");
  }
  ~CollEngine() { gmt_coll_engine_destroy(e_); }
  CollEngine(const CollEngine&) = delete;
  CollEngine& operator=(const CollEngine&) = delete;
  gmt_coll_engine* get() const { return e_; }
  int world() const { return gmt_coll_engine_world(e_); }
  double* buffer(int rank) const
  {
    double* p = nullptr;
    check(gmt_coll_engine_buffer(e_, rank, &p, nullptr), "buffer");
    return p;
  }
  gmt_stream_t stream(int rank) const
  {
    gmt_stream_t s = nullptr;
    check(gmt_coll_engine_stream(e_, rank, &s), "stream");
    return s;
  }
  int rank_device(int rank) const
  {
    int d = 0;
    check(gmt_coll_engine_rank_device(e_, rank, &d), "rank_device");
    return d;
  }
  void bind(int rank, const double* in, double* out = nullptr, double* y = nullptr)
  {
    check(gmt_coll_engine_bind(e_, rank, in, out, y), "bind");
  }
  void allgather(int64_t n, bool from_in = false) { check(gmt_coll_engine_allgather(e_, n, from_in), "allgather"); }
  void gather(int64_t n, int root) { check(gmt_coll_engine_gather(e_, n, root), "gather"); }
  void allreduce(int64_t n, int algo = 0) { check(gmt_coll_engine_allreduce(e_, n, algo), "allreduce"); }
  void daxpy_allreduce(int64_t n, double a) { check(gmt_coll_engine_daxpy_allreduce(e_, n, a), "daxpy_allreduce"); }
  void barrier() { check(gmt_coll_engine_barrier(e_), "barrier"); }
  void synchronize() { check(gmt_coll_engine_synchronize(e_), "synchronize"); }
  void status() { check(gmt_coll_engine_status(e_, nullptr), "status"); }
  /* device milliseconds of fn(), max over the GPUs */
  template <class F>
  double timed(F&& fn)
  {
    double ms = 0.0;
    check(gmt_coll_engine_mark(e_, 0), "mark");
    fn();
    check(gmt_coll_engine_mark(e_, 1), "mark");
    check(gmt_coll_engine_elapsed_ms(e_, &ms), "elapsed_ms");
    return ms;
  }

 private:
  gmt_coll_engine* e_ = nullptr;
};

}  // namespace gmt
#endif  // GMT_B200_HPP
