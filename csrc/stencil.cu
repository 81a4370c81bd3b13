// stencil.cu -- fp64 5-point first-derivative stencils (single GPU) and the
// device-side analytic initialisers.
//
// Replaces, from the reference:
//   stencil1d_5            mpi_stencil_gt.cc:54-59      (gmt_stencil1d_f64)
//   stencil2d_1d_5_d0/_d1  mpi_stencil2d_gt.cc:84-110   (gmt_stencil2d_f64)
//   stencil2d_1d_5 (SYCL)  mpi_stencil2d_sycl.cc:53-75, mpi_stencil2d_sycl_oo.cc:289-310
//   host init loops        mpi_stencil_gt.cc:177-196, mpi_stencil2d_gt.cc:445-497
// and adds the cross stencil (d/dx + d/dy) BASELINE config 3 needs.
// The halo-fused versions of the same tiles are in stencil_fused.cu.
#include <algorithm>

#include "stencil_tiles.cuh"

extern "C" int gmt_stencil2d_d1_cluster_f64(int64_t nx, int64_t ny, const double* z, int64_t ldz, double* out,
                                            int64_t ldo, double scale, gmt_stream_t stream);

namespace gmt {

// ============================================================== variant 0
// One thread per output point, scalar loads: what a generic expression-template
// kernel does.  Fallback for shapes the vector kernels do not take, and the
// "un-fused generic kernel" figure in the benchmarks.
__global__ void __launch_bounds__(256)
  stencil1d_scalar_kernel(int64_t n, const double* __restrict__ in, double* __restrict__ out,
                          double scale)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = sten5(in[i], in[i + 1], in[i + 3], in[i + 4]) * scale;
}

__global__ void __launch_bounds__(256)
  stencil2d_scalar_kernel(int64_t nx, int64_t ny, const double* __restrict__ z, int64_t ldz,
                          double* __restrict__ out, int64_t ldo, int gx, int gy, double sx, double sy)
{
  const int64_t total = nx * ny;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += stride) {
    const int64_t j = p / nx, i = p - j * nx;
    double v = 0.0;
    if (gx) {
      const double* c = z + i + (j + gy) * ldz;
      v += sten5(c[0], c[1], c[3], c[4]) * sx;
    }
    if (gy) {
      const double* c = z + (i + gx) + j * ldz;
      v += sten5(c[0], c[ldz], c[3 * ldz], c[4 * ldz]) * sy;
    }
    out[i + j * ldo] = v;
  }
}

// ============================================================== variant 1
// One tile per CTA, CTAs dispatched by the hardware in blockIdx (= address) order: measured
// ~5% faster on B200 than a persistent grid-stride loop over the same tiles (the chip-wide
// access front stays tight, see tools/bw_sweep.cu and daxpy.cu).
__global__ void __launch_bounds__(kStThreads, 2) stencil1d_vec_kernel(Sten1d S)
{
  double err = 0.0;
  tile1d<false>(S, (int64_t)blockIdx.x * kTile1d, threadIdx.x, err);
}

template <int GX, int GY>
__global__ void __launch_bounds__(kStThreads, 2) stencil2d_vec_kernel(Sten2d S, int64_t tiles_i, int TJ)
{
  const int64_t t = blockIdx.x;
  const int64_t tj = t / tiles_i, ti = t - tj * tiles_i;
  const int64_t j0 = tj * TJ;
  double err = 0.0;
  tile2d<GX, GY, false>(S, ti * kTileI, j0, min(S.ny, j0 + (int64_t)TJ), threadIdx.x, err);
}

// ============================================================== initialisers
__global__ void __launch_bounds__(256)
  init1d_kernel(int64_t n, int nb, double x0, double dx, double* y, double* actual, int left, int right)
{
  const int64_t total = n + 2 * nb;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += stride) {
    const int64_t i = g - nb;
    const double x = x0 + (double)i * dx;
    const bool ghost = i < 0 || i >= n;
    if (!ghost) {
      y[g] = x * x * x;
      if (actual) actual[i] = 3 * x * x;
    } else if ((i < 0 && left) || (i >= n && right)) {
      y[g] = x * x * x;
    } else {
      y[g] = 0.0;
    }
  }
}

__global__ void __launch_bounds__(256)
  init2d_kernel(int64_t nx, int64_t ny, int gx, int gy, double x0, double y0, double dx, double dy,
                double* z, int64_t ldz, double* actual, int64_t lda, int mode, double sx, double sy,
                int ghost_mask, double poison)
{
  const int64_t tx = nx + 2 * gx, ty = ny + 2 * gy;
  const int64_t total = tx * ty;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += stride) {
    const int64_t gj = p / tx, gi = p - gj * tx;
    const int64_t i = gi - gx, j = gj - gy;
    const double x = x0 + (double)i * dx, y = y0 + (double)j * dy;
    const double v = x * x * x + y * y;
    int side = 0;
    if (i < 0) side |= 1;
    if (i >= nx) side |= 2;
    if (j < 0) side |= 4;
    if (j >= ny) side |= 8;
    if (side == 0) {
      z[gi + gj * ldz] = v;
      if (actual && mode) {
        double a = 0.0;
        if (mode == 1) a = 3 * x * x;
        if (mode == 2) a = 2 * y;
        if (mode == 3) a = sx * 3 * x * x + sy * 2 * y;
        actual[i + j * lda] = a;
      }
    } else {
      // corners are never read by the axis-aligned stencils
      const bool corner = (side & 3) && (side & 12);
      z[gi + gj * ldz] = (!corner && (side & ghost_mask)) ? v : poison;
    }
  }
}

// ============================================================== launch helpers
static bool vec2d_ok(int64_t nx, const double* z, int64_t ldz, const double* out, int64_t ldo)
{
  return nx % 4 == 0 && ldz % 4 == 0 && ldo % 4 == 0 && aligned(z, 32) && aligned(out, 32);
}

static int launch_plain2d(int gx, int gy, int64_t nx, int64_t ny, const double* z, int64_t ldz,
                          double* out, int64_t ldo, double sx, double sy, int variant, cudaStream_t st)
{
  GMT_REQUIRE(nx > 0 && ny > 0 && z && out, GMT_EINVAL, "stencil2d: bad argument");
  GMT_REQUIRE(ldz >= nx + 2 * gx && ldo >= nx, GMT_EINVAL, "stencil2d: leading dimension too small");
  GMT_REQUIRE(variant >= 0 && variant <= 2, GMT_EINVAL, "stencil2d: unknown variant %d", variant);
  if (variant == 2 && !gx && gy && vec2d_ok(nx, z, ldz, out, ldo))
    return gmt_stencil2d_d1_cluster_f64(nx, ny, z, ldz, out, ldo, sy, (gmt_stream_t)st);
  if (variant == 0 || !vec2d_ok(nx, z, ldz, out, ldo)) {
    int64_t want = (nx * ny + 255) / 256;
    int grid = (int)std::min<int64_t>(want, (int64_t)sm_count() * 8);
    stencil2d_scalar_kernel<<<grid, 256, 0, st>>>(nx, ny, z, ldz, out, ldo, gx, gy, sx, sy);
    GMT_LAUNCH_CHECK();
    return 0;
  }
  Sten2d S{z, out, nx, ny, ldz, ldo, sx, sy, 0, 0, 0, 0};
  const int64_t tiles_i = (nx + kTileI - 1) / kTileI;
  const int tj = runtime_tile_j(gx, gy, false);
  const int64_t nt = tiles_i * ((ny + tj - 1) / tj);
  GMT_REQUIRE(nt < (1LL << 31), GMT_ETOOBIG, "stencil2d: array too large for a one-shot grid");
  const int grid = (int)nt;
  if (gx && !gy) {
    stencil2d_vec_kernel<2, 0><<<grid, kStThreads, 0, st>>>(S, tiles_i, tj);
  } else if (!gx && gy) {
    stencil2d_vec_kernel<0, 2><<<grid, kStThreads, 0, st>>>(S, tiles_i, tj);
  } else {
    stencil2d_vec_kernel<2, 2><<<grid, kStThreads, 0, st>>>(S, tiles_i, tj);
  }
  GMT_LAUNCH_CHECK();
  return 0;
}

}  // namespace gmt

extern "C" {

int gmt_stencil1d_tma_f64(int64_t n, const double* in, double* out, double scale, gmt_stream_t stream);

int gmt_stencil1d_f64(int64_t n, const double* in, double* out, double scale, int variant,
                      gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(n >= 0 && (n == 0 || (in && out)), GMT_EINVAL, "gmt_stencil1d_f64: bad argument");
  GMT_REQUIRE(variant >= 0 && variant <= 2, GMT_EINVAL, "gmt_stencil1d_f64: unknown variant %d", variant);
  if (n == 0) return 0;
  cudaStream_t st = as_stream(stream);
  const bool vec = aligned(in, 32) && aligned(out, 32);
  if (variant == 2 && vec && n % 2 == 0 && n >= 4096) return gmt_stencil1d_tma_f64(n, in, out, scale, stream);
  if (variant == 0 || !vec) {
    int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)sm_count() * 8);
    stencil1d_scalar_kernel<<<grid, 256, 0, st>>>(n, in, out, scale);
    GMT_LAUNCH_CHECK();
    return 0;
  }
  Sten1d S{in, out, n, scale, 0.0, 0.0};
  const int64_t nt = (n + kTile1d - 1) / kTile1d;
  GMT_REQUIRE(nt < (1LL << 31), GMT_ETOOBIG, "gmt_stencil1d_f64: array too large for a one-shot grid");
  stencil1d_vec_kernel<<<(int)nt, kStThreads, 0, st>>>(S);
  GMT_LAUNCH_CHECK();
  return 0;
}

int gmt_stencil2d_f64(int dim, int64_t nx, int64_t ny, const double* z, int64_t ldz, double* dz,
                      int64_t lddz, double scale, int variant, gmt_stream_t stream)
{
  GMT_REQUIRE(dim == 0 || dim == 1, GMT_EINVAL, "gmt_stencil2d_f64: dim must be 0 or 1");
  return gmt::launch_plain2d(dim == 0 ? 2 : 0, dim == 1 ? 2 : 0, nx, ny, z, ldz, dz, lddz, scale, scale,
                             variant, gmt::as_stream(stream));
}

int gmt_stencil2d_xy_f64(int64_t nx, int64_t ny, const double* z, int64_t ldz, double* out,
                         int64_t ldo, double sx, double sy, gmt_stream_t stream)
{
  return gmt::launch_plain2d(2, 2, nx, ny, z, ldz, out, ldo, sx, sy, 1, gmt::as_stream(stream));
}

int gmt_init_stencil1d_f64(int64_t n, int nb, double x0, double dx, double* y, double* actual,
                           int fill_left_ghost, int fill_right_ghost, gmt_stream_t stream)
{
  GMT_REQUIRE(n > 0 && nb >= 0 && y, GMT_EINVAL, "gmt_init_stencil1d_f64: bad argument");
  int grid = (int)std::min<int64_t>((n + 2 * nb + 255) / 256, (int64_t)gmt::sm_count() * 8);
  gmt::init1d_kernel<<<grid, 256, 0, gmt::as_stream(stream)>>>(n, nb, x0, dx, y, actual, fill_left_ghost,
                                                              fill_right_ghost);
  GMT_LAUNCH_CHECK();
  return 0;
}

int gmt_init_stencil2d_f64(int64_t nx, int64_t ny, int gx, int gy, double x0, double y0, double dx,
                           double dy, double* z, int64_t ldz, double* actual, int64_t lda,
                           int deriv_mode, double sx, double sy, int ghost_mask, double poison,
                           gmt_stream_t stream)
{
  GMT_REQUIRE(nx > 0 && ny > 0 && z && ldz >= nx + 2 * gx, GMT_EINVAL, "gmt_init_stencil2d_f64: bad argument");
  GMT_REQUIRE(!actual || lda >= nx, GMT_EINVAL, "gmt_init_stencil2d_f64: lda < nx");
  const int64_t total = (nx + 2 * gx) * (ny + 2 * gy);
  int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)gmt::sm_count() * 8);
  gmt::init2d_kernel<<<grid, 256, 0, gmt::as_stream(stream)>>>(nx, ny, gx, gy, x0, y0, dx, dy, z, ldz,
                                                              actual, lda, deriv_mode, sx, sy, ghost_mask,
                                                              poison);
  GMT_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
