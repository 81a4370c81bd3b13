// reduce.cu -- fp64 reductions, fills and strided block copies.
//
// Device-side replacements for work the reference does on the host or through
// generic gtensor / SYCL kernels:
//   gmt_sum_f64          host sum loops               mpi_daxpy_nvtx.cc:253-256,296-298
//   gmt_sumsq_diff*_f64  gt::sum_squares / diff_norm  mpi_stencil_gt.cc:222,
//                                                     mpi_stencil2d_sycl.cc:165-181
//   gmt_axis_sum_f64     gt::sum_axis_to              mpi_stencil2d_gt.cc:611,620
//   gmt_fill_f64         fill constructor             mpi_stencil2d_gt.cc:598
//   gmt_copy2d_f64       pack / unpack view assigns   mpi_stencil2d_gt.cc:166,172,239,251
//                        buf_from_view / buf_to_view  mpi_stencil2d_sycl.cc:82-116
// All reductions are deterministic for a fixed size (fixed grid, ordered final add).
#include "common.cuh"

namespace gmt {

constexpr int kRedThreads = 512;

// ---------------------------------------------------------------- full sums
// mode 0: sum(a)   mode 1: sum((a-b)^2)
// Persistent grid; chunks of kRedThreads x 2 vectors are drawn in address order from the
// counter in scratch word 1 (an in-order front streams HBM ~5% faster than a grid-stride
// loop, see daxpy.cu).
constexpr int64_t kRedChunk = (int64_t)kRedThreads * 2;

template <int kMode>
__global__ void __launch_bounds__(kRedThreads, 2)
  reduce_kernel(int64_t n, const double* __restrict__ a, const double* __restrict__ b, double* out,
                void* scratch, bool vec)
{
  __shared__ double red[32];
  __shared__ int64_t next[2];
  unsigned int* counter = reinterpret_cast<unsigned int*>(scratch) + 1;
  double acc0 = 0.0, acc1 = 0.0;
  if (vec) {
    const int64_t n4 = n / 4;
    const int64_t nchunks = (n4 + kRedChunk - 1) / kRedChunk;
    // next chunk drawn while the current one streams (see daxpy.cu)
    if (threadIdx.x == 0) next[0] = (int64_t)atomicAdd(counter, 1u);
    __syncthreads();
    int64_t c = next[0];
    int slot = 0;
    while (c < nchunks) {
      unsigned int prefetched = 0;
      if (threadIdx.x == 0) prefetched = atomicAdd(counter, 1u);
      const int64_t i0 = c * kRedChunk + threadIdx.x, i1 = i0 + kRedThreads;
      double4_t u0 = {0, 0, 0, 0}, u1 = u0, w0 = u0, w1 = u0;
      if (i0 < n4) u0 = ldg256_stream(a + 4 * i0);
      if (i1 < n4) u1 = ldg256_stream(a + 4 * i1);
      if (kMode == 1) {
        if (i0 < n4) w0 = ldg256_stream(b + 4 * i0);
        if (i1 < n4) w1 = ldg256_stream(b + 4 * i1);
        u0.x -= w0.x; u0.y -= w0.y; u0.z -= w0.z; u0.w -= w0.w;
        u1.x -= w1.x; u1.y -= w1.y; u1.z -= w1.z; u1.w -= w1.w;
        acc0 += (u0.x * u0.x + u0.y * u0.y) + (u0.z * u0.z + u0.w * u0.w);
        acc1 += (u1.x * u1.x + u1.y * u1.y) + (u1.z * u1.z + u1.w * u1.w);
      } else {
        acc0 += (u0.x + u0.y) + (u0.z + u0.w);
        acc1 += (u1.x + u1.y) + (u1.z + u1.w);
      }
      if (threadIdx.x == 0) next[slot ^ 1] = (int64_t)prefetched;
      __syncthreads();
      slot ^= 1;
      c = next[slot];
    }
    if (blockIdx.x == 0) {
      for (int64_t k = 4 * n4 + threadIdx.x; k < n; k += kRedThreads) {
        double d = kMode == 1 ? a[k] - b[k] : a[k];
        acc1 += kMode == 1 ? d * d : d;
      }
    }
  } else {
    const int64_t stride = (int64_t)gridDim.x * kRedThreads;
    for (int64_t i = (int64_t)blockIdx.x * kRedThreads + threadIdx.x; i < n; i += stride) {
      double d = kMode == 1 ? a[i] - b[i] : a[i];
      acc0 += kMode == 1 ? d * d : d;
    }
  }
  double part = block_sum(acc0 + acc1, red);
  double total;
  if (grid_sum_last_block(part, scratch, red, total, gridDim.x, blockIdx.x)) {
    if (threadIdx.x == 0) {
      *out = total;
      *counter = 0u;
    }
  }
}

// strided 2-D squared-difference sum: one CTA row-chunk at a time
__global__ void __launch_bounds__(kRedThreads, 2)
  sumsq_diff2d_kernel(int64_t nx, int64_t ny, const double* __restrict__ a, int64_t lda,
                      const double* __restrict__ b, int64_t ldb, double* out, void* scratch)
{
  __shared__ double red[32];
  double acc = 0.0;
  const int64_t total_pts = nx * ny;
  const int64_t stride = (int64_t)gridDim.x * kRedThreads;
  for (int64_t p = (int64_t)blockIdx.x * kRedThreads + threadIdx.x; p < total_pts; p += stride) {
    const int64_t j = p / nx, i = p - j * nx;
    const double d = a[i + j * lda] - b[i + j * ldb];
    acc += d * d;
  }
  double part = block_sum(acc, red);
  double total;
  if (grid_sum_last_block(part, scratch, red, total, gridDim.x, blockIdx.x)) {
    if (threadIdx.x == 0) *out = total;
  }
}

// ---------------------------------------------------------------- axis sums
// axis 0: out[j] = sum_i z(i,j): one warp per column, coalesced along i.
__global__ void __launch_bounds__(256)
  axis0_sum_kernel(int64_t nx, int64_t ld, const double* __restrict__ z, int64_t n_out, double* out)
{
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t j = warp; j < n_out; j += nwarps) {
    const double* col = z + j * ld;
    double a0 = 0.0, a1 = 0.0;
    int64_t i = lane;
    for (; i + 32 < nx; i += 64) {
      a0 += col[i];
      a1 += col[i + 32];
    }
    if (i < nx) a0 += col[i];
    double s = warp_sum(a0 + a1);
    if (lane == 0) out[j] = s;
  }
}

// axis 1: out[i] = sum_j z(i,j): a (32 x 8) thread tile walks the columns, lanes along i
// (coalesced), the 8 thread rows split j; partials combined through shared memory.
__global__ void __launch_bounds__(256)
  axis1_sum_kernel(int64_t ny, int64_t ld, const double* __restrict__ z, int64_t n_out, double* out)
{
  __shared__ double tile[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int64_t i0 = (int64_t)blockIdx.x * 32; i0 < n_out; i0 += (int64_t)gridDim.x * 32) {
    const int64_t i = i0 + tx;
    double acc = 0.0;
    if (i < n_out) {
      for (int64_t j = ty; j < ny; j += 8) acc += z[i + j * ld];
    }
    tile[ty][tx] = acc;
    __syncthreads();
    if (ty == 0 && i < n_out) {
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < 8; ++k) s += tile[k][tx];
      out[i] = s;
    }
    __syncthreads();
  }
}

// -------------------------------------------------------------- fills, copies
__global__ void __launch_bounds__(512) fill_kernel(int64_t n, double v, double* x)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) x[i] = v;
}

__global__ void __launch_bounds__(256) fill_u64_kernel(int64_t n, uint64_t v, uint64_t* x)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) x[i] = v;
}

__global__ void __launch_bounds__(512)
  iota_kernel(int64_t n, double start, double step, double mul, double* x)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    x[i] = (start + (double)i * step) * mul;
}

// x[i-1] = (rank * i) / N in INTEGER arithmetic, i = 1..n: the slot pattern of the reference's
// in-place gather test (mpigatherinplace.f90:34-37), done in 64 bits (the Fortran default
// integers overflow at >= 16 ranks).
__global__ void __launch_bounds__(512) slot_pattern_kernel(int64_t n, int64_t rank, int64_t N, double* x)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    x[i] = (double)((rank * (i + 1)) / N);
}

// dst(i,j) = src(i,j).  The interesting shapes are the halo strips: rows = 2 with half a
// million columns (dim 0: one 16-byte element pair every ld doubles) and rows = 524288 with
// 2 columns (dim 1: two long contiguous runs).  Threads are flattened over (i,j) with i
// fastest so both extremes stay as coalesced as the layout allows.
__global__ void __launch_bounds__(256)
  copy2d_kernel(int64_t rows, int64_t cols, const double* __restrict__ src, int64_t ld_src,
                double* __restrict__ dst, int64_t ld_dst)
{
  const int64_t total = rows * cols;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += stride) {
    const int64_t j = p / rows, i = p - j * rows;
    dst[i + j * ld_dst] = src[i + j * ld_src];
  }
}

// rows == 2 fast path: one 16-byte access per column when both sides are 16-byte aligned
__global__ void __launch_bounds__(256)
  copy2d_rows2_kernel(int64_t cols, const double* __restrict__ src, int64_t ld_src,
                      double* __restrict__ dst, int64_t ld_dst)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < cols; j += stride)
    stg128(dst + j * ld_dst, ldg128(src + j * ld_src));
}

static int grid_for(int64_t work_items, int threads, int per_sm)
{
  int64_t want = (work_items + threads - 1) / threads;
  return (int)std::min<int64_t>(std::max<int64_t>(want, 1), (int64_t)sm_count() * per_sm);
}

}  // namespace gmt

extern "C" {

size_t gmt_reduce_scratch_bytes(void) { return gmt::kReduceScratchBytes; }

int gmt_sum_f64(int64_t n, const double* x, double* out, void* scratch, gmt_stream_t stream)
{
  GMT_REQUIRE(n >= 0 && out && scratch && (n == 0 || x), GMT_EINVAL, "gmt_sum_f64: bad argument");
  const bool vec = gmt::aligned(x, 32);
  int grid = gmt::grid_for(vec ? n / 8 + 1 : n, gmt::kRedThreads, 2);
  gmt::reduce_kernel<0><<<grid, gmt::kRedThreads, 0, gmt::as_stream(stream)>>>(n, x, nullptr, out,
                                                                             scratch, vec);
  GMT_LAUNCH_CHECK();
  return 0;
}

int gmt_sumsq_diff_f64(int64_t n, const double* a, const double* b, double* out, void* scratch,
                       gmt_stream_t stream)
{
  GMT_REQUIRE(n >= 0 && out && scratch && (n == 0 || (a && b)), GMT_EINVAL,
              "gmt_sumsq_diff_f64: bad argument");
  const bool vec = gmt::aligned(a, 32) && gmt::aligned(b, 32);
  int grid = gmt::grid_for(vec ? n / 8 + 1 : n, gmt::kRedThreads, 2);
  gmt::reduce_kernel<1><<<grid, gmt::kRedThreads, 0, gmt::as_stream(stream)>>>(n, a, b, out, scratch,
                                                                             vec);
  GMT_LAUNCH_CHECK();
  return 0;
}

int gmt_sumsq_diff2d_f64(int64_t nx, int64_t ny, const double* a, int64_t lda, const double* b,
                         int64_t ldb, double* out, void* scratch, gmt_stream_t stream)
{
  GMT_REQUIRE(nx >= 0 && ny >= 0 && out && scratch, GMT_EINVAL, "gmt_sumsq_diff2d_f64: bad argument");
  GMT_REQUIRE(lda >= nx && ldb >= nx, GMT_EINVAL, "gmt_sumsq_diff2d_f64: leading dimension < nx");
  if (lda == nx && ldb == nx) return gmt_sumsq_diff_f64(nx * ny, a, b, out, scratch, stream);
  int grid = gmt::grid_for(nx * ny, gmt::kRedThreads, 2);
  gmt::sumsq_diff2d_kernel<<<grid, gmt::kRedThreads, 0, gmt::as_stream(stream)>>>(nx, ny, a, lda, b,
                                                                                ldb, out, scratch);
  GMT_LAUNCH_CHECK();
  return 0;
}

int gmt_axis_sum_f64(int64_t nx, int64_t ny, int64_t ld, const double* z, int axis, int64_t n_out,
                     double* out, gmt_stream_t stream)
{
  GMT_REQUIRE(nx > 0 && ny > 0 && ld >= nx && z && out, GMT_EINVAL, "gmt_axis_sum_f64: bad argument");
  GMT_REQUIRE(axis == 0 || axis == 1, GMT_EINVAL, "gmt_axis_sum_f64: axis must be 0 or 1");
  const int64_t surviving = axis == 0 ? ny : nx;
  GMT_REQUIRE(n_out >= 0 && n_out <= surviving, GMT_EINVAL,
              "gmt_axis_sum_f64: n_out %lld exceeds surviving extent %lld", (long long)n_out,
              (long long)surviving);
  if (n_out == 0) return 0;
  cudaStream_t st = gmt::as_stream(stream);
  if (axis == 0) {
    int grid = gmt::grid_for(n_out * 32, 256, 8);
    gmt::axis0_sum_kernel<<<grid, 256, 0, st>>>(nx, ld, z, n_out, out);
  } else {
    int grid = (int)std::min<int64_t>((n_out + 31) / 32, (int64_t)gmt::sm_count() * 8);
    gmt::axis1_sum_kernel<<<grid, 256, 0, st>>>(ny, ld, z, n_out, out);
  }
  GMT_LAUNCH_CHECK();
  return 0;
}

int gmt_fill_f64(int64_t n, double value, double* x, gmt_stream_t stream)
{
  GMT_REQUIRE(n >= 0 && (n == 0 || x), GMT_EINVAL, "gmt_fill_f64: bad argument");
  if (n == 0) return 0;
  gmt::fill_kernel<<<gmt::grid_for(n, 512, 4), 512, 0, gmt::as_stream(stream)>>>(n, value, x);
  GMT_LAUNCH_CHECK();
  return 0;
}

int gmt_fill_u64(int64_t n, uint64_t value, uint64_t* x, gmt_stream_t stream)
{
  GMT_REQUIRE(n >= 0 && (n == 0 || x), GMT_EINVAL, "gmt_fill_u64: bad argument");
  if (n == 0) return 0;
  gmt::fill_u64_kernel<<<gmt::grid_for(n, 256, 1), 256, 0, gmt::as_stream(stream)>>>(n, value, x);
  GMT_LAUNCH_CHECK();
  return 0;
}

int gmt_iota_f64(int64_t n, double start, double step, double mul, double* x, gmt_stream_t stream)
{
  GMT_REQUIRE(n >= 0 && (n == 0 || x), GMT_EINVAL, "gmt_iota_f64: bad argument");
  if (n == 0) return 0;
  gmt::iota_kernel<<<gmt::grid_for(n, 512, 4), 512, 0, gmt::as_stream(stream)>>>(n, start, step, mul,
                                                                               x);
  GMT_LAUNCH_CHECK();
  return 0;
}

int gmt_slot_pattern_f64(int64_t n, int64_t rank, int64_t N, double* x, gmt_stream_t stream)
{
  GMT_REQUIRE(n >= 0 && N > 0 && rank >= 0 && (n == 0 || x), GMT_EINVAL, "gmt_slot_pattern_f64: bad argument");
  if (n == 0) return 0;
  gmt::slot_pattern_kernel<<<gmt::grid_for(n, 512, 4), 512, 0, gmt::as_stream(stream)>>>(n, rank, N, x);
  GMT_LAUNCH_CHECK();
  return 0;
}

int gmt_copy2d_f64(int64_t rows, int64_t cols, const double* src, int64_t ld_src, double* dst,
                   int64_t ld_dst, gmt_stream_t stream)
{
  GMT_REQUIRE(rows >= 0 && cols >= 0, GMT_EINVAL, "gmt_copy2d_f64: negative extent");
  GMT_REQUIRE(ld_src >= rows && ld_dst >= rows, GMT_EINVAL, "gmt_copy2d_f64: leading dimension < rows");
  if (rows == 0 || cols == 0) return 0;
  GMT_REQUIRE(src && dst, GMT_EINVAL, "gmt_copy2d_f64: NULL array");
  cudaStream_t st = gmt::as_stream(stream);
  if (rows == 2 && gmt::aligned(src, 16) && gmt::aligned(dst, 16) && ld_src % 2 == 0 &&
      ld_dst % 2 == 0) {
    gmt::copy2d_rows2_kernel<<<gmt::grid_for(cols, 256, 8), 256, 0, st>>>(cols, src, ld_src, dst,
                                                                         ld_dst);
  } else {
    gmt::copy2d_kernel<<<gmt::grid_for(rows * cols, 256, 8), 256, 0, st>>>(rows, cols, src, ld_src,
                                                                          dst, ld_dst);
  }
  GMT_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
