// p2p.cuh -- cross-GPU memory-ordering primitives for kernels that talk over
// NVLink 5 / NVSwitch peer mappings (P2P or NVLS multicast).
//
// The reference has nothing like this: it serialises every exchange with
// device-wide syncs around MPI calls (mpi_stencil2d_gt.cc:202,254).  Here a
// producer CTA writes a peer's inbox with ordinary stores, then publishes an
// epoch number with a system-scope release; the consumer polls with a
// system-scope acquire.  Epochs only ever grow, so flags are never reset and
// there is no reset race; inboxes are double-buffered by epoch parity.
#pragma once
#include <stdint.h>

namespace gmt {

__device__ __forceinline__ void fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }

__device__ __forceinline__ void st_release_sys(uint64_t* p, uint64_t v)
{
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p)
{
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t ld_relaxed_sys(const uint64_t* p)
{
  uint64_t v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys(uint64_t* p, uint64_t v)
{
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void red_add_release_sys(uint64_t* p, uint64_t v)
{
  asm volatile("red.release.sys.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// data loads of memory a peer GPU wrote: must not be served from a stale L1 line
__device__ __forceinline__ double ld_peer_f64(const double* p)
{
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double2 ld_peer_f64x2(const double* p)
{
  double2 v;
  asm volatile("ld.relaxed.sys.global.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
  return v;
}

// NVLS multicast (scalar fp64 forms are the only ones the ISA has)
__device__ __forceinline__ double multimem_ld_reduce_add_f64(const double* mc)
{
  double v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f64 %0, [%1];" : "=d"(v) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f64(double* mc, double v)
{
  asm volatile("multimem.st.relaxed.sys.global.f64 [%0], %1;" ::"l"(mc), "d"(v) : "memory");
}
__device__ __forceinline__ void multimem_red_add_f64(double* mc, double v)
{
  asm volatile("multimem.red.relaxed.sys.global.add.f64 [%0], %1;" ::"l"(mc), "d"(v) : "memory");
}

// "LL" cells: 16 bytes carry one double as two 8-byte words, each = 32 payload bits + the 32-bit
// epoch.  8-byte stores are single-copy atomic, so a reader that sees the epoch in BOTH words has
// the whole value: data and flag travel together, no fence, no separate flag store.
__device__ __forceinline__ void st_ll_cell(uint64_t* cell, double v, uint64_t e32)
{
  const uint64_t bits = (uint64_t)__double_as_longlong(v);
  const uint64_t w0 = (bits & 0xffffffffull) | (e32 << 32), w1 = (bits >> 32) | (e32 << 32);
  asm volatile("st.volatile.global.v2.u64 [%0], {%1,%2};" ::"l"(cell), "l"(w0), "l"(w1) : "memory");
}
__device__ __forceinline__ bool ld_ll_cell(const uint64_t* cell, uint64_t e32, double& v)
{
  uint64_t w0, w1;
  asm volatile("ld.volatile.global.v2.u64 {%0,%1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(cell) : "memory");
  if ((w0 >> 32) != e32 || (w1 >> 32) != e32) return false;
  v = __longlong_as_double((long long)((w0 & 0xffffffffull) | (w1 << 32)));
  return true;
}

// the two halves of that: a raw load (several can be in flight) and the check
__device__ __forceinline__ void ld_ll_raw(const uint64_t* cell, uint64_t& w0, uint64_t& w1)
{
  asm volatile("ld.volatile.global.v2.u64 {%0,%1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(cell) : "memory");
}
__device__ __forceinline__ bool ll_unpack(uint64_t w0, uint64_t w1, uint64_t e32, double& v)
{
  if ((w0 >> 32) != e32 || (w1 >> 32) != e32) return false;
  v = __longlong_as_double((long long)((w0 & 0xffffffffull) | (w1 << 32)));
  return true;
}

constexpr int64_t kDefaultSpinLimit = 1LL << 28;  // ~ tens of seconds of polling

// Bounded spin until *flag >= want.  Returns false on expiry (dead / never-launched
// peer); the caller records that in its status word instead of hanging the GPU.
__device__ __forceinline__ bool wait_flag_ge(const uint64_t* flag, uint64_t want, int64_t limit)
{
  if (limit <= 0) limit = kDefaultSpinLimit;
  for (int64_t it = 0; it < limit; ++it) {
    if (ld_acquire_sys(flag) >= want) return true;
    if (it > 64) __nanosleep(it > 4096 ? 256 : 32);
  }
  return false;
}

}  // namespace gmt
