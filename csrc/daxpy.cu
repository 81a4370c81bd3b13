// daxpy.cu -- fp64 axpy kernels.
//
// gmt_daxpy_f64 replaces the cublasDaxpy call sites (daxpy.cu:73,
// mpi_daxpy_nvtx.cc:245, gt::blas::axpy mpi_daxpy_gt.cc:81).
// gmt_daxpy_sum_f64 additionally produces sum(y_new) in the same pass, which in
// the reference is a D2H copy plus a host loop (mpi_daxpy_nvtx.cc:251-268).
//
// 24 B/element, no reuse: a pure HBM stream.  Measured on B200 (tools/bw_sweep.cu):
// what matters is that the chip-wide access front stays tight in address order --
// one chunk per CTA in blockIdx order (6.9 TB/s) beats a grid-stride persistent loop
// (6.1-6.7 TB/s) -- while vector width, unroll and cache hints are second order.  The
// plain kernel therefore uses a one-shot grid; the kernel with the fused sum, which
// wants few partials, is persistent but draws its chunks from an atomic counter,
// which keeps the same in-order front.
#include "common.cuh"

namespace gmt {

constexpr int kAxpyThreads = 512;
constexpr int kAxpyUnroll = 2;
constexpr int64_t kAxpyChunk = (int64_t)kAxpyThreads * kAxpyUnroll;  // 256-bit vectors per chunk

// one chunk of kAxpyChunk vectors starting at vector index v0
template <bool kSum>
__device__ __forceinline__ void axpy_chunk(int64_t v0, int64_t n4, double a, const double* __restrict__ x,
                                           double* __restrict__ y, double& acc)
{
  double4_t xv[kAxpyUnroll], yv[kAxpyUnroll];
#pragma unroll
  for (int u = 0; u < kAxpyUnroll; ++u) {
    const int64_t i = v0 + threadIdx.x + (int64_t)u * kAxpyThreads;
    if (i < n4) {
      xv[u] = ldg256_stream(x + 4 * i);
      yv[u] = ldg256(y + 4 * i);
    }
  }
#pragma unroll
  for (int u = 0; u < kAxpyUnroll; ++u) {
    const int64_t i = v0 + threadIdx.x + (int64_t)u * kAxpyThreads;
    if (i < n4) {
      yv[u].x = a * xv[u].x + yv[u].x;
      yv[u].y = a * xv[u].y + yv[u].y;
      yv[u].z = a * xv[u].z + yv[u].z;
      yv[u].w = a * xv[u].w + yv[u].w;
      stg256(y + 4 * i, yv[u]);
      if (kSum) acc += (yv[u].x + yv[u].y) + (yv[u].z + yv[u].w);
    }
  }
}

__global__ void __launch_bounds__(kAxpyThreads, 2)
  daxpy_vec4_kernel(int64_t n4, int64_t n, double a, const double* __restrict__ x, double* __restrict__ y)
{
  double acc = 0.0;
  axpy_chunk<false>((int64_t)blockIdx.x * kAxpyChunk, n4, a, x, y, acc);
  // scalar tail (n not a multiple of 4): last CTA
  if (blockIdx.x == gridDim.x - 1) {
    for (int64_t k = 4 * n4 + threadIdx.x; k < n; k += kAxpyThreads) y[k] = a * x[k] + y[k];
  }
}

// persistent, chunks drawn in order from the counter in scratch word 1
__global__ void __launch_bounds__(kAxpyThreads, 2)
  daxpy_sum_vec4_kernel(int64_t n4, int64_t n, double a, const double* __restrict__ x,
                        double* __restrict__ y, double* out_sum, void* scratch)
{
  __shared__ double red[32];
  __shared__ int64_t next[2];
  unsigned int* counter = reinterpret_cast<unsigned int*>(scratch) + 1;
  const int64_t nchunks = (n4 + kAxpyChunk - 1) / kAxpyChunk;
  double acc = 0.0;
  // the draw for the next chunk is issued before the current chunk is processed, so the
  // atomic's round trip hides behind the chunk; two smem slots => one barrier per chunk
  if (threadIdx.x == 0) next[0] = (int64_t)atomicAdd(counter, 1u);
  __syncthreads();
  int64_t c = next[0];
  int slot = 0;
  while (c < nchunks) {
    unsigned int prefetched = 0;
    if (threadIdx.x == 0) prefetched = atomicAdd(counter, 1u);
    axpy_chunk<true>(c * kAxpyChunk, n4, a, x, y, acc);
    if (threadIdx.x == 0) next[slot ^ 1] = (int64_t)prefetched;
    __syncthreads();
    slot ^= 1;
    c = next[slot];
  }
  if (blockIdx.x == 0) {
    for (int64_t k = 4 * n4 + threadIdx.x; k < n; k += kAxpyThreads) {
      const double v = a * x[k] + y[k];
      y[k] = v;
      acc += v;
    }
  }
  const double part = block_sum(acc, red);
  double total;
  if (grid_sum_last_block(part, scratch, red, total, gridDim.x, blockIdx.x)) {
    if (threadIdx.x == 0) {
      *out_sum = total;
      *counter = 0u;
    }
  }
}

template <bool kSum>
__global__ void __launch_bounds__(kAxpyThreads, 2)
  daxpy_scalar_kernel(int64_t n, double a, const double* __restrict__ x, double* __restrict__ y,
                      double* out_sum, void* scratch)
{
  __shared__ double red[32];
  const int64_t stride = (int64_t)gridDim.x * kAxpyThreads;
  double acc = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * kAxpyThreads + threadIdx.x; i < n; i += stride) {
    const double v = a * x[i] + y[i];
    y[i] = v;
    if (kSum) acc += v;
  }
  if (kSum) {
    const double part = block_sum(acc, red);
    double total;
    if (grid_sum_last_block(part, scratch, red, total, gridDim.x, blockIdx.x)) {
      if (threadIdx.x == 0) *out_sum = total;
    }
  }
}

template <bool kSum>
static int launch_daxpy(int64_t n, double a, const double* x, double* y, double* out_sum,
                        void* scratch, cudaStream_t st)
{
  GMT_REQUIRE(n >= 0, GMT_EINVAL, "daxpy: negative n");
  GMT_REQUIRE(n == 0 || (x && y), GMT_EINVAL, "daxpy: NULL array");
  if (kSum) GMT_REQUIRE(out_sum && scratch, GMT_EINVAL, "daxpy_sum: NULL out/scratch");
  const int sms = sm_count();
  if (n == 0) {
    if (kSum) GMT_CUDA(cudaMemsetAsync(out_sum, 0, sizeof(double), st));
    return 0;
  }
  const bool vec = aligned(x, 32) && aligned(y, 32);
  if (vec) {
    const int64_t n4 = n / 4;
    const int64_t nchunks = std::max<int64_t>(1, (n4 + kAxpyChunk - 1) / kAxpyChunk);
    if (kSum) {
      const int grid = (int)std::min<int64_t>(nchunks, (int64_t)sms * 2);
      daxpy_sum_vec4_kernel<<<grid, kAxpyThreads, 0, st>>>(n4, n, a, x, y, out_sum, scratch);
    } else {
      GMT_REQUIRE(nchunks < (1LL << 31), GMT_ETOOBIG, "daxpy: n too large for a one-shot grid");
      daxpy_vec4_kernel<<<(int)nchunks, kAxpyThreads, 0, st>>>(n4, n, a, x, y);
    }
  } else {
    const int64_t want = (n + kAxpyThreads - 1) / kAxpyThreads;
    const int grid = (int)std::min<int64_t>(std::max<int64_t>(want, 1), (int64_t)sms * 2);
    daxpy_scalar_kernel<kSum><<<grid, kAxpyThreads, 0, st>>>(n, a, x, y, out_sum, scratch);
  }
  GMT_LAUNCH_CHECK();
  return 0;
}

}  // namespace gmt

extern "C" {

int gmt_daxpy_f64(int64_t n, double a, const double* x, double* y, gmt_stream_t stream)
{
  return gmt::launch_daxpy<false>(n, a, x, y, nullptr, nullptr, gmt::as_stream(stream));
}

int gmt_daxpy_sum_f64(int64_t n, double a, const double* x, double* y, double* out_sum,
                      void* scratch, gmt_stream_t stream)
{
  return gmt::launch_daxpy<true>(n, a, x, y, out_sum, scratch, gmt::as_stream(stream));
}

}  // extern "C"
