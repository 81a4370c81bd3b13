// stencil_tiles.cuh -- per-tile device code shared by the plain (stencil.cu) and
// the fused halo+stencil kernels (stencil_fused.cu).
//
// Design (B200, fp64, 16 B/point => pure HBM stream):
//   * every thread owns 4 consecutive points of the contiguous axis and moves
//     them with 256-bit accesses (LDG.E.256 / STG.E.256);
//   * along the strided axis a thread marches over columns keeping four partial
//     sums in registers, so every input byte is read from HBM once;
//   * loads are issued in hand-made batches of 4 columns (up to 8 x 32 B per
//     thread in flight) so that ~128 KB per SM is outstanding at 2 CTAs/SM.
#pragma once
#include <cstdlib>

#include "common.cuh"
#include "p2p.cuh"

namespace gmt {

constexpr int kStThreads = 256;           // threads per CTA
constexpr int kTileI = kStThreads * 4;    // points of the contiguous axis per tile
#ifndef GMT_KBATCH
#define GMT_KBATCH 4
#endif
constexpr int kBatch = GMT_KBATCH;        // columns (or sub-blocks) loaded per batch
constexpr int kTile1d = kTileI * kBatch;  // 1-D tile: 4096 points
constexpr int kTileJx = 32;               // columns per tile when only dim 0 is differentiated
constexpr int kTileJy = 128;              // columns per tile when dim 1 is differentiated
constexpr int kFlagCols = 1024;           // columns covered by one x-face flag word

template <int GY>
__host__ __device__ constexpr int tile_j()
{
  return GY ? kTileJy : kTileJx;
}

// Tile width in columns used by a launch.  Defaults measured on B200 (tools/tile_sweep.py):
// small tiles when columns are independent (dim 0: 8 columns = 128 KB of traffic, for the plain
// and the warp-specialised fused kernel alike), 16 for the plain both-dims kernel, long marches
// when columns are coupled (dim 1 and the fused both-dims kernel: 128).  $GMT_TILE_JX / $GMT_TILE_JY
// override (x-face flag chunks need kFlagCols % width == 0).
inline int runtime_tile_j(int gx, int gy, bool fused)
{
  int dflt;
  if (gy == 0) dflt = 8;
  else if (gx == 0) dflt = kTileJy;
  else dflt = fused ? kTileJy : 16;
  const char* e = getenv(gy ? "GMT_TILE_JY" : "GMT_TILE_JX");
  const int v = e ? atoi(e) : 0;
  if (v < 4 || v > kStThreads || (kFlagCols % v) != 0) return dflt;
  return v;
}

// four outputs from eight consecutive inputs: o_k = sten5(in[k..k+4]) * s
__device__ __forceinline__ double4_t sten_along(const double4_t& A, const double4_t& B, double s)
{
  double4_t o;
  o.x = sten5(A.x, A.y, A.w, B.x) * s;
  o.y = sten5(A.y, A.z, B.x, B.y) * s;
  o.z = sten5(A.z, A.w, B.y, B.z) * s;
  o.w = sten5(A.w, B.x, B.z, B.w) * s;
  return o;
}

__device__ __forceinline__ double sq(double v) { return v * v; }

// ============================================================== 1-D
struct Sten1d {
  const double* in;
  double* out;
  int64_t n;
  double scale;
  double x0, dx;
  // analytic d/dx of x^3 weighted like the numeric one: scale * sum c_k y_k ~ (scale*dx) * 3x^2
  __device__ __forceinline__ double an(double x) const { return (scale * dx) * 3.0 * x * x; }
};

// one tile = kBatch sub-blocks of kTileI points; all loads first, then math + stores
template <bool kErr>
__device__ __forceinline__ void tile1d(const Sten1d& S, int64_t t0, int tid, double& err)
{
  double4_t A[kBatch], B[kBatch];
#pragma unroll
  for (int u = 0; u < kBatch; ++u) {
    const int64_t i = t0 + (int64_t)u * kTileI + 4 * tid;
    if (i + 3 < S.n) {
      A[u] = ldg256(S.in + i);
      B[u] = ldg256(S.in + i + 4);
    }
  }
#pragma unroll
  for (int u = 0; u < kBatch; ++u) {
    const int64_t i = t0 + (int64_t)u * kTileI + 4 * tid;
    if (i + 3 < S.n) {
      const double4_t o = sten_along(A[u], B[u], S.scale);
      stg256(S.out + i, o);
      if (kErr) {
        const double x = S.x0 + (double)i * S.dx;
        const double x1 = x + S.dx, x2 = x + 2 * S.dx, x3 = x + 3 * S.dx;
        err += (sq(o.x - S.an(x)) + sq(o.y - S.an(x1))) + (sq(o.z - S.an(x2)) + sq(o.w - S.an(x3)));
      }
    } else if (i < S.n) {
      for (int64_t k = i; k < S.n; ++k) {
        const double o = sten5(S.in[k], S.in[k + 1], S.in[k + 3], S.in[k + 4]) * S.scale;
        S.out[k] = o;
        if (kErr) err += sq(o - S.an(S.x0 + (double)k * S.dx));
      }
    }
  }
}

// ============================================================== 2-D
struct Sten2d {
  const double* z;
  double* out;
  int64_t nx, ny, ldz, ldo;
  double sx, sy;
  double x0, y0, dx, dy;
  // analytic derivatives of x^3 + y^2 weighted like the numeric ones (sx = wx/dx, sy = wy/dy)
  __device__ __forceinline__ double sx_an(double x) const { return (sx * dx) * 3.0 * x * x; }
  __device__ __forceinline__ double sy_an(double y) const { return (sy * dy) * 2.0 * y; }
};

// Tile = rows [i0, i0+kTileI) x columns [j0, j1).  GX/GY: ghost width (0 or 2) of z in each
// dim, i.e. whether that dim is differentiated.  Requires nx % 4 == 0 and 32-byte aligned
// columns (ldz, ldo multiples of 4).
template <int GX, int GY, bool kErr>
__device__ __forceinline__ void tile2d(const Sten2d& S, int64_t i0, int64_t j0, int64_t j1, int tid,
                                       double& err)
{
  const int64_t i = i0 + 4 * tid;
  if (i >= S.nx) return;  // nx % 4 == 0: a thread is all in or all out
  // analytic x-derivative of this thread's four rows (the same for every column)
  double ex0 = 0, ex1 = 0, ex2 = 0, ex3 = 0;
  if (kErr && GX) {
    const double x = S.x0 + (double)i * S.dx;
    ex0 = S.sx_an(x);
    ex1 = S.sx_an(x + S.dx);
    ex2 = S.sx_an(x + 2 * S.dx);
    ex3 = S.sx_an(x + 3 * S.dx);
  }
  if (GY == 0) {
    // dim 0 only: columns are independent; z has no ghost columns
    for (int64_t j = j0; j < j1; j += kBatch) {
      double4_t A[kBatch], B[kBatch];
#pragma unroll
      for (int u = 0; u < kBatch; ++u) {
        if (j + u < j1) {
          const double* c = S.z + i + (j + u) * S.ldz;
          A[u] = ldg256(c);
          B[u] = ldg256(c + 4);
        }
      }
#pragma unroll
      for (int u = 0; u < kBatch; ++u) {
        if (j + u < j1) {
          const double4_t o = sten_along(A[u], B[u], S.sx);
          stg256(S.out + i + (j + u) * S.ldo, o);
          if (kErr) err += (sq(o.x - ex0) + sq(o.y - ex1)) + (sq(o.z - ex2) + sq(o.w - ex3));
        }
      }
    }
  } else {
    // dim 1 (and optionally dim 0): march over ghosted columns g = j0 .. j1+3 keeping the
    // partial sums of outputs g-1 .. g-4 in registers
    const double w0 = S.sy * c0(), w1 = S.sy * c1(), w3 = S.sy * c3(), w4 = S.sy * c4();
    double4_t a1 = {0, 0, 0, 0}, a2 = a1, a3 = a1, a4 = a1;
    const int64_t gend = j1 + 4;
    for (int64_t g = j0; g < gend; g += kBatch) {
      double4_t A[kBatch], B[kBatch];
#pragma unroll
      for (int u = 0; u < kBatch; ++u) {
        if (g + u < gend) {
          const double* c = S.z + i + (g + u) * S.ldz;
          A[u] = ldg256(c);
          if (GX) B[u] = ldg256(c + 4);
        }
      }
#pragma unroll
      for (int u = 0; u < kBatch; ++u) {
        if (g + u < gend) {
          double4_t m;
          if (GX) {
            m.x = A[u].z; m.y = A[u].w; m.z = B[u].x; m.w = B[u].y;
          } else {
            m = A[u];
          }
          // output column j = g+u-4 is complete with this contribution
          const int64_t j = g + u - 4;
          if (j >= j0) {
            double4_t o;
            o.x = a4.x + w4 * m.x; o.y = a4.y + w4 * m.y; o.z = a4.z + w4 * m.z; o.w = a4.w + w4 * m.w;
            stg256(S.out + i + j * S.ldo, o);
            if (kErr) {
              const double ey = S.sy_an(S.y0 + (double)j * S.dy);
              err += (sq(o.x - (ex0 + ey)) + sq(o.y - (ex1 + ey))) + (sq(o.z - (ex2 + ey)) + sq(o.w - (ex3 + ey)));
            }
          }
          a4.x = a3.x + w3 * m.x; a4.y = a3.y + w3 * m.y; a4.z = a3.z + w3 * m.z; a4.w = a3.w + w3 * m.w;
          if (GX) {
            const double4_t d = sten_along(A[u], B[u], S.sx);
            a3.x = a2.x + d.x; a3.y = a2.y + d.y; a3.z = a2.z + d.z; a3.w = a2.w + d.w;
          } else {
            a3 = a2;
          }
          a2.x = a1.x + w1 * m.x; a2.y = a1.y + w1 * m.y; a2.z = a1.z + w1 * m.z; a2.w = a1.w + w1 * m.w;
          a1.x = w0 * m.x; a1.y = w0 * m.y; a1.z = w0 * m.z; a1.w = w0 * m.w;
        }
      }
    }
  }
}

// dim-0 tile whose x-face ghost pairs come from shared memory (filled from the neighbours' inbox by the
// communication warp) instead of the array's own ghost cells: glo / ghi hold one pair per column of
// the tile, or are NULL when the tile does not touch that face (or the face is a physical boundary,
// whose analytic ghosts are in the array).  Only the thread that owns rows 0..3 (resp. nx-4..nx-1)
// patches two of its eight loaded values; everything else is the plain tile.
template <bool kErr>
__device__ __forceinline__ void tile2d_d0_inbox(const Sten2d& S, int64_t i0, int64_t j0, int64_t j1, int tid,
                                                double& err, const double2* glo, const double2* ghi)
{
  const int64_t i = i0 + 4 * tid;
  if (i >= S.nx) return;
  double ex0 = 0, ex1 = 0, ex2 = 0, ex3 = 0;
  if (kErr) {
    const double x = S.x0 + (double)i * S.dx;
    ex0 = S.sx_an(x);
    ex1 = S.sx_an(x + S.dx);
    ex2 = S.sx_an(x + 2 * S.dx);
    ex3 = S.sx_an(x + 3 * S.dx);
  }
  // 0: plain thread, 1: owns rows 0..3 (low ghosts), 2: owns rows nx-4..nx-1 (high ghosts), 3: both (nx == 4)
  const int patch = ((glo != nullptr && i == 0) ? 1 : 0) | ((ghi != nullptr && i == S.nx - 4) ? 2 : 0);
  for (int64_t j = j0; j < j1; j += kBatch) {
    double4_t A[kBatch], B[kBatch];
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      if (j + u < j1) {
        const double* c = S.z + i + (j + u) * S.ldz;
        A[u] = ldg256(c);
        B[u] = ldg256(c + 4);
      }
    }
    if (patch) {   // two threads of the CTA at most
#pragma unroll
      for (int u = 0; u < kBatch; ++u) {
        if (j + u < j1) {
          if (patch & 1) {
            const double2 g = glo[j + u - j0];
            A[u].x = g.x;
            A[u].y = g.y;
          }
          if (patch & 2) {
            const double2 g = ghi[j + u - j0];
            B[u].z = g.x;
            B[u].w = g.y;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      if (j + u < j1) {
        const double4_t o = sten_along(A[u], B[u], S.sx);
        stg256(S.out + i + (j + u) * S.ldo, o);
        if (kErr) err += (sq(o.x - ex0) + sq(o.y - ex1)) + (sq(o.z - ex2) + sq(o.w - ex3));
      }
    }
  }
}

// grid size of a persistent kernel: SMs x occupancy, shared between n_ranks virtual ranks
inline int persistent_grid(const void* kernel, int threads, int64_t ntiles, int n_ranks, int* out_grid)
{
  int occ = 0;
  GMT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, threads, 0));
  if (occ < 1) occ = 1;
  int64_t cap = (int64_t)sm_count() * occ;
  if (n_ranks > 1) cap /= n_ranks;
  GMT_REQUIRE(cap >= 1, GMT_ETOOBIG, "too many virtual ranks (%d) for one cooperative launch", n_ranks);
  *out_grid = (int)std::max<int64_t>(1, std::min<int64_t>(cap, ntiles));
  return 0;
}

}  // namespace gmt
