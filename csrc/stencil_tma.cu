// stencil_tma.cu -- the 1-D stencil fed by the Tensor Memory Accelerator.
//
// Same arithmetic as stencil1d_5 (mpi_stencil_gt.cc:54-59).  A producer thread
// streams (tile + 4)-point windows of the input into a ring of shared-memory
// stages with cp.async.bulk (SASS: UBLKCP) and mbarrier transaction counts;
// eight consumer warps read each window from shared memory (the 4-point overlap
// between neighbouring threads costs LDS bandwidth instead of L1/LSU traffic)
// and store 256-bit results.  The ring gives each SM ~64-96 KB of loads in
// flight from two threads' worth of issue slots.
#include "common.cuh"
#include "tma.cuh"

namespace gmt {

constexpr int kTmaTile = 4096;   // outputs per stage
constexpr int kTmaStages = 3;
constexpr int kTmaConsumers = 256;
constexpr int kTmaStageBytes = ((kTmaTile + 4) * 8 + 127) / 128 * 128;
constexpr int kTmaSmemBytes = 128 + kTmaStages * kTmaStageBytes;

__global__ void __launch_bounds__(kTmaConsumers + 32)
  stencil1d_tma_kernel(int64_t n, const double* __restrict__ in, double* __restrict__ out, double scale)
{
  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty = full + kTmaStages;
  unsigned char* data = smem + 128;
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < kTmaStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kTmaConsumers / 32);
    }
    mbar_fence_init();
  }
  __syncthreads();
  const int64_t ntiles = (n + kTmaTile - 1) / kTmaTile;

  if (tid >= kTmaConsumers) {
    if (tid == kTmaConsumers) {
      int k = 0;
      for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x, ++k) {
        const int s = k % kTmaStages;
        const uint32_t ph = (k / kTmaStages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        const int64_t t0 = t * kTmaTile;
        const int64_t cnt = min((int64_t)kTmaTile, n - t0) + 4;
        const uint32_t bytes = (uint32_t)(cnt * 8);
        mbar_arrive_expect_tx(&full[s], bytes);
        bulk_g2s(data + s * kTmaStageBytes, in + t0, bytes, &full[s]);
      }
    }
    return;
  }

  int k = 0;
  for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x, ++k) {
    const int s = k % kTmaStages;
    const uint32_t ph = (k / kTmaStages) & 1;
    const int64_t t0 = t * kTmaTile;
    const double* w = reinterpret_cast<const double*>(data + s * kTmaStageBytes);
    mbar_wait(&full[s], ph);
#pragma unroll
    for (int u = 0; u < kTmaTile / (kTmaConsumers * 4); ++u) {
      const int idx = u * kTmaConsumers * 4 + 4 * tid;
      const int64_t i = t0 + idx;
      if (i + 3 < n) {
        const double2 a0 = *reinterpret_cast<const double2*>(w + idx);
        const double2 a1 = *reinterpret_cast<const double2*>(w + idx + 2);
        const double2 b0 = *reinterpret_cast<const double2*>(w + idx + 4);
        const double2 b1 = *reinterpret_cast<const double2*>(w + idx + 6);
        double4_t o;
        o.x = sten5(a0.x, a0.y, a1.y, b0.x) * scale;
        o.y = sten5(a0.y, a1.x, b0.x, b0.y) * scale;
        o.z = sten5(a1.x, a1.y, b0.y, b1.x) * scale;
        o.w = sten5(a1.y, b0.x, b1.x, b1.y) * scale;
        stg256(out + i, o);
      } else if (i < n) {
        for (int64_t q = i; q < n; ++q) {
          const int r = (int)(q - t0);
          out[q] = sten5(w[r], w[r + 1], w[r + 3], w[r + 4]) * scale;
        }
      }
    }
    __syncwarp();
    if ((tid & 31) == 0) mbar_arrive(&empty[s]);
  }
}

}  // namespace gmt

extern "C" int gmt_stencil1d_tma_f64(int64_t n, const double* in, double* out, double scale,
                                     gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(n > 0 && in && out, GMT_EINVAL, "gmt_stencil1d_tma_f64: bad argument");
  GMT_REQUIRE(n % 2 == 0 && aligned(in, 32) && aligned(out, 32), GMT_EUNSUPPORTED,
              "gmt_stencil1d_tma_f64: needs even n and 32-byte aligned arrays");
  static thread_local int configured_dev = -1;
  int dev = 0;
  GMT_CUDA(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    GMT_CUDA(cudaFuncSetAttribute(stencil1d_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  kTmaSmemBytes));
    configured_dev = dev;
  }
  int occ = 0;
  GMT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, stencil1d_tma_kernel, kTmaConsumers + 32,
                                                         kTmaSmemBytes));
  if (occ < 1) occ = 1;
  const int64_t ntiles = (n + kTmaTile - 1) / kTmaTile;
  const int grid = (int)std::min<int64_t>(ntiles, (int64_t)sm_count() * occ);
  stencil1d_tma_kernel<<<grid, kTmaConsumers + 32, kTmaSmemBytes, as_stream(stream)>>>(n, in, out, scale);
  GMT_LAUNCH_CHECK();
  return 0;
}
