// common.cuh -- device-side building blocks shared by every kernel file.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../include/gmt_b200.h"
#include "check.h"

namespace gmt {

// B200: 148 SMs.  Persistent grids are sized as a multiple of the SM count
// queried at run time; this constant only bounds scratch sizes.
constexpr int kMaxCtas = 2048;
constexpr int kScratchHeaderBytes = 256;  // [0]: u32 ticket counter, rest pad
constexpr size_t kReduceScratchBytes = kScratchHeaderBytes + sizeof(double) * kMaxCtas * 2;

// 4th-order central first-derivative weights (mpi_stencil_gt.cc:46-47); the
// centre weight is exactly 0 and is not evaluated.
__host__ __device__ constexpr double c0() { return 1.0 / 12.0; }
__host__ __device__ constexpr double c1() { return -2.0 / 3.0; }
__host__ __device__ constexpr double c3() { return 2.0 / 3.0; }
__host__ __device__ constexpr double c4() { return -1.0 / 12.0; }

__device__ __forceinline__ double sten5(double a, double b, double d, double e)
{
  // same association order as the reference expression (left to right)
  return ((c0() * a + c1() * b) + c3() * d) + c4() * e;
}

// ---- 256-bit global accesses (sm_100a: LDG.E.256 / STG.E.256) --------------
struct __align__(32) double4_t {
  double x, y, z, w;
};

__device__ __forceinline__ double4_t ldg256(const double* p)
{
  double4_t r;
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w)
               : "l"(p)
               : "memory");
  return r;
}
// streaming read: read-only path, do not keep in L1
__device__ __forceinline__ double4_t ldg256_stream(const double* p)
{
  double4_t r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg256(double* p, const double4_t& v)
{
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.x), "d"(v.y), "d"(v.z),
               "d"(v.w)
               : "memory");
}
// streaming store: evict-first in L2, no L1 allocation
__device__ __forceinline__ void stg256_stream(double* p, const double4_t& v)
{
  asm volatile("st.global.cs.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.x), "d"(v.y), "d"(v.z),
               "d"(v.w)
               : "memory");
}
__device__ __forceinline__ double2 ldg128(const double* p)
{
  double2 r;
  asm volatile("ld.global.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void stg128(double* p, double2 v)
{
  asm volatile("st.global.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
}

// ---- reductions ------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sum over the CTA; result valid in thread 0.  smem: >= 32 doubles.
__device__ __forceinline__ double block_sum(double v, double* smem)
{
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  double r = 0.0;
  if (warp == 0) {
    r = lane < nw ? smem[lane] : 0.0;
    r = warp_sum(r);
  }
  __syncthreads();
  return r;
}

// Deterministic grid-wide sum without a second launch: every CTA parks its
// partial in scratch, takes a ticket, and the CTA that draws the last ticket
// adds the partials in index order.  Returns true (in every thread of the last
// CTA) with the total in `total` (valid in thread 0).  The ticket counter is
// reset by the last CTA, so the scratch is reusable by the next launch on the
// same stream without a memset.
__device__ __forceinline__ bool grid_sum_last_block(double partial_in_t0, void* scratch,
                                                    double* smem, double& total,
                                                    unsigned int nblk, unsigned int bid)
{
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  double* partials =
    reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + kScratchHeaderBytes);
  __shared__ bool is_last;
  if (threadIdx.x == 0) {
    partials[bid] = partial_in_t0;
    __threadfence();
    unsigned int t = atomicAdd(ticket, 1u);
    is_last = (t == nblk - 1);
  }
  __syncthreads();
  if (!is_last) return false;
  __threadfence();
  double v = 0.0;
  for (unsigned int i = threadIdx.x; i < nblk; i += blockDim.x) v += __ldcg(&partials[i]);
  total = block_sum(v, smem);
  if (threadIdx.x == 0) *ticket = 0u;
  return true;
}

inline int sm_count()
{
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev);
    cached_dev = dev;
    if (cached <= 0) cached = 148;
  }
  return cached;
}

// Zero-initialised, self-resetting work counters (4 x u32) private to (device, stream), for
// persistent kernels whose entry point has no scratch parameter.  Allocated on first use and
// kept for the life of the process (defined in runtime.cu).
int stream_counters(cudaStream_t stream, unsigned int** out);

inline cudaStream_t as_stream(gmt_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

template <typename T>
__host__ __device__ inline bool aligned(const T* p, size_t a)
{
  return (reinterpret_cast<uintptr_t>(p) % a) == 0;
}

}  // namespace gmt
