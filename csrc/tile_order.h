// tile_order.h -- the order in which the fused 2-D kernel visits its tiles (pure integer logic,
// usable from host code so that it can be unit-tested without a GPU: tests/test_cpu_units.py
// compiles it with g++ and checks that every tile is visited exactly once in every configuration).
//
// Tiles are (ti, tj): ti indexes 1024-row blocks along dim 0 (fastest in memory), tj blocks of TJ
// columns.  "Interior" tiles [a0,a1) x [b0,b1) come first in memory order (ti fastest), then the
// excluded low-tj column blocks, the high-tj ones, and finally the excluded ti rows of the
// interior tj range.  With natural_x the ti range is never cut (a0 = 0, a1 = tiles_i): x-face
// tiles stay where memory order puts them and only y-face column blocks are deferred.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define GMT_HD __host__ __device__ __forceinline__
#else
#define GMT_HD inline
#endif

namespace gmt {

struct TileMap {
  int64_t tiles_i, tiles_j, a0, a1, b0, b1, ni, nj, n_int;
};

// has_*: the side is fed by a neighbour or by the host (its tiles must wait for flags)
GMT_HD TileMap tile_map_make(int64_t tiles_i, int64_t tiles_j, bool has_lo_x, bool has_hi_x, bool has_lo_y,
                             bool has_hi_y, bool natural_x)
{
  TileMap M;
  M.tiles_i = tiles_i;
  M.tiles_j = tiles_j;
  M.a0 = (!natural_x && has_lo_x) ? 1 : 0;
  const int64_t a1 = tiles_i - ((!natural_x && has_hi_x) ? 1 : 0);
  M.a1 = a1 > M.a0 ? a1 : M.a0;
  M.b0 = has_lo_y ? 1 : 0;
  const int64_t b1 = tiles_j - (has_hi_y ? 1 : 0);
  M.b1 = b1 > M.b0 ? b1 : M.b0;
  M.ni = M.a1 - M.a0;
  M.nj = M.b1 - M.b0;
  M.n_int = M.ni * M.nj;
  return M;
}

// counter value -> tile, interior tiles first, boundary tiles after them
GMT_HD void decode_tile(const TileMap& M, int64_t c, int64_t& ti, int64_t& tj)
{
  if (c < M.n_int) {
    tj = M.b0 + c / M.ni;
    ti = M.a0 + c % M.ni;
    return;
  }
  int64_t k = c - M.n_int;
  if (M.n_int == 0) {
    tj = k / M.tiles_i;
    ti = k % M.tiles_i;
    return;
  }
  const int64_t nbot = M.b0 * M.tiles_i, ntop = (M.tiles_j - M.b1) * M.tiles_i;
  if (k < nbot) {
    tj = k / M.tiles_i;
    ti = k % M.tiles_i;
  } else if (k < nbot + ntop) {
    k -= nbot;
    tj = M.b1 + k / M.tiles_i;
    ti = k % M.tiles_i;
  } else {
    k -= nbot + ntop;
    const int64_t nside = M.tiles_i - M.ni;
    tj = M.b0 + k / nside;
    const int64_t idx = k % nside;
    ti = idx < M.a0 ? idx : M.a1 + (idx - M.a0);
  }
}

}  // namespace gmt
