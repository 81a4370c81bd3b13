// noalloc.cu -- gmt_daxpy_sum_noalloc_f64: daxpy + sum for callers that must not allocate.
//
// No counterpart in the reference (its daxpy programs allocate everything they need with
// cudaMalloc / cudaMallocManaged up front, mpi_daxpy_nvtx.cc:177-199).
//
// Who this is for.  gmt_daxpy_sum_f64 wants a scratch buffer from its caller, and the Python
// layer gets that buffer from an allocator.  Some users of this library cannot do either:
//   * a solver that sized its arrays to fill the GPU's memory at start-up, or that lives
//     inside an application whose memory pool already owns the whole device, so that a late
//     cudaMalloc may simply fail;
//   * a time-stepping loop in which an allocator call is not acceptable because
//     cudaMalloc / cudaFree may synchronise the whole device and stall unrelated streams.
// For them the scratch lives in the library image itself: a file-scope __device__ array that
// every GPU loading the library holds until the process exits (kNoAllocScratchBytes, 64.25 KiB).
// The caller passes only inputs and outputs.  This entry point calls no cudaMalloc, cudaFree
// or any allocating library routine.
//
// What the fixed scratch costs: one partial sum per 4096-element chunk is parked in it, so
// n is limited to GMT_NOALLOC_MAX_N = 8192 * 4096 = 33,554,432 elements; larger n is refused
// with GMT_ETOOBIG.  (In exchange the chunks map 1:1 to CTAs in address order -- the fastest
// streaming pattern on B200 -- and the partials are added in index order, so the sum is
// bit-reproducible run to run, which the counter-scheduled gmt_daxpy_sum_f64 is not.)
//
// Rules of use
//   * The first call on each device does the set-up for that device (it asks the runtime for
//     the device address of the array).  Make that first call before any stream capture.
//   * All calls on a device share the one array, and the array keeps whatever the last call
//     left in it.  The words that must start at zero (the arrival counter) are cleared on the
//     stream at the start of every call.  Two calls that overlap on the GPU (different streams
//     without an ordering between them, or two host threads) race on the array and produce
//     wrong sums: order them.
#include <mutex>

#include "common.cuh"

namespace gmt {

constexpr int kNoAllocThreads = 512;
constexpr int kNoAllocUnroll = 2;
constexpr int64_t kNoAllocChunkVec = (int64_t)kNoAllocThreads * kNoAllocUnroll;  // 256-bit vectors per CTA
constexpr int64_t kNoAllocChunkElems = kNoAllocChunkVec * 4;                     // 4096 doubles
constexpr int kNoAllocMaxChunks = 8192;
constexpr int kNoAllocHeaderBytes = 256;
constexpr size_t kNoAllocScratchBytes = kNoAllocHeaderBytes + sizeof(double) * kNoAllocMaxChunks;
constexpr int64_t kNoAllocMaxN = (int64_t)kNoAllocMaxChunks * kNoAllocChunkElems;
constexpr int kMaxDevices = 64;

// The scratch.  In host code this name is only a stand-in for the symbol, not a device
// address: host code must go through cudaGetSymbolAddress (done once per device below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

// One CTA per chunk, partial per CTA, last-arriving CTA adds the partials in index order.
__global__ void __launch_bounds__(kNoAllocThreads, 2)
  daxpy_sum_noalloc_kernel(int64_t n4, int64_t n, double a, const double* __restrict__ x,
                           double* __restrict__ y, double* out_sum, unsigned char* scratch, bool vec)
{
  __shared__ double red[32];
  __shared__ bool is_last;
  unsigned int* arrived = reinterpret_cast<unsigned int*>(scratch);
  double* partials = reinterpret_cast<double*>(scratch + kNoAllocHeaderBytes);
  double acc = 0.0;
  if (vec) {
    const int64_t v0 = (int64_t)blockIdx.x * kNoAllocChunkVec;
    double4_t xv[kNoAllocUnroll], yv[kNoAllocUnroll];
#pragma unroll
    for (int u = 0; u < kNoAllocUnroll; ++u) {
      const int64_t i = v0 + threadIdx.x + (int64_t)u * kNoAllocThreads;
      if (i < n4) {
        xv[u] = ldg256_stream(x + 4 * i);
        yv[u] = ldg256(y + 4 * i);
      }
    }
#pragma unroll
    for (int u = 0; u < kNoAllocUnroll; ++u) {
      const int64_t i = v0 + threadIdx.x + (int64_t)u * kNoAllocThreads;
      if (i < n4) {
        yv[u].x = a * xv[u].x + yv[u].x;
        yv[u].y = a * xv[u].y + yv[u].y;
        yv[u].z = a * xv[u].z + yv[u].z;
        yv[u].w = a * xv[u].w + yv[u].w;
        stg256(y + 4 * i, yv[u]);
        acc += (yv[u].x + yv[u].y) + (yv[u].z + yv[u].w);
      }
    }
    if (blockIdx.x == gridDim.x - 1) {  // elements past the last full vector
      for (int64_t k = 4 * n4 + threadIdx.x; k < n; k += kNoAllocThreads) {
        const double v = a * x[k] + y[k];
        y[k] = v;
        acc += v;
      }
    }
  } else {
    // arrays not 32-byte aligned: same chunking, scalar accesses
    const int64_t e0 = (int64_t)blockIdx.x * kNoAllocChunkElems;
    const int64_t e1 = min(n, e0 + kNoAllocChunkElems);
    for (int64_t k = e0 + threadIdx.x; k < e1; k += kNoAllocThreads) {
      const double v = a * x[k] + y[k];
      y[k] = v;
      acc += v;
    }
  }
  const double part = block_sum(acc, red);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = part;
    __threadfence();
    is_last = (atomicAdd(arrived, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  double v = 0.0;
  for (unsigned int i = threadIdx.x; i < gridDim.x; i += kNoAllocThreads) v += __ldcg(&partials[i]);
  const double total = block_sum(v, red);
  if (threadIdx.x == 0) *out_sum = total;
}

// Device address of the scratch on the current device, looked up once per device.  The look-up
// (which may load the module on that device) must not happen inside a stream capture: a first call
// on a capturing stream is refused with a clear message instead of failing somewhere in the runtime.
// The address is cached for the life of the process; a cudaDeviceReset() in the host program
// invalidates it (and everything else this library holds on that device).
static int noalloc_scratch_ptr(unsigned char** out, cudaStream_t st)
{
  static std::mutex mu;
  static unsigned char* ptrs[kMaxDevices] = {nullptr};
  int dev = 0;
  GMT_CUDA(cudaGetDevice(&dev));
  GMT_REQUIRE(dev >= 0 && dev < kMaxDevices, GMT_EUNSUPPORTED, "gmt_daxpy_sum_noalloc_f64: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    GMT_CUDA(cudaStreamIsCapturing(st, &cap));
    GMT_REQUIRE(cap == cudaStreamCaptureStatusNone, GMT_EUNSUPPORTED,
                "gmt_daxpy_sum_noalloc_f64: the first call on device %d must come before any stream capture "
                "(it looks up the scratch array's device address)", dev);
    void* p = nullptr;
    // `dev` is current here, so this is the address of THIS device's copy of the array
    GMT_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    GMT_REQUIRE(p != nullptr, GMT_EUNSUPPORTED, "gmt_daxpy_sum_noalloc_f64: no device address for the scratch array");
    ptrs[dev] = static_cast<unsigned char*>(p);
  }
  *out = ptrs[dev];
  return 0;
}

}  // namespace gmt

extern "C" {

int64_t gmt_noalloc_max_n(void) { return gmt::kNoAllocMaxN; }

size_t gmt_noalloc_scratch_bytes(void) { return gmt::kNoAllocScratchBytes; }

int gmt_daxpy_sum_noalloc_f64(int64_t n, double a, const double* x, double* y, double* out_sum,
                              gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(n >= 0, GMT_EINVAL, "gmt_daxpy_sum_noalloc_f64: negative n");
  GMT_REQUIRE(out_sum && (n == 0 || (x && y)), GMT_EINVAL, "gmt_daxpy_sum_noalloc_f64: NULL argument");
  GMT_REQUIRE(n <= kNoAllocMaxN, GMT_ETOOBIG,
              "gmt_daxpy_sum_noalloc_f64: n = %lld exceeds the fixed scratch capacity (max %lld elements); "
              "use gmt_daxpy_sum_f64 with caller-provided scratch for larger vectors",
              (long long)n, (long long)kNoAllocMaxN);
  cudaStream_t st = as_stream(stream);
  unsigned char* scratch = nullptr;
  int rc = noalloc_scratch_ptr(&scratch, st);
  if (rc) return rc;
  // the arrival counter must start at zero; the array still holds the previous call's state
  GMT_CUDA(cudaMemsetAsync(scratch, 0, kNoAllocHeaderBytes, st));
  if (n == 0) {
    GMT_CUDA(cudaMemsetAsync(out_sum, 0, sizeof(double), st));
    return 0;
  }
  const bool vec = aligned(x, 32) && aligned(y, 32);
  const int64_t nchunks = (n + kNoAllocChunkElems - 1) / kNoAllocChunkElems;
  daxpy_sum_noalloc_kernel<<<(int)nchunks, kNoAllocThreads, 0, st>>>(n / 4, n, a, x, y, out_sum, scratch, vec);
  GMT_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
