// Host-side unit test of csrc/tile_order.h (the tile order of the fused 2-D kernel).
// Built and run by tests/test_cpu_units.py::test_tile_order_is_a_bijection with g++; no GPU.
#include <cstdio>
#include <set>
#include <utility>
#include <vector>

#include "tile_order.h"

static int fail(const char* what, int ti_n, int tj_n, int has, int nat)
{
  std::printf("FAIL %s: tiles %dx%d has=%d natural=%d\n", what, ti_n, tj_n, has, nat);
  return 1;
}

int main()
{
  long checked = 0;
  for (int ti_n = 1; ti_n <= 7; ++ti_n)
    for (int tj_n = 1; tj_n <= 7; ++tj_n)
      for (int has = 0; has < 16; ++has)
        for (int nat = 0; nat < 2; ++nat) {
          const bool lx = has & 1, hx = has & 2, ly = has & 4, hy = has & 8;
          const gmt::TileMap M = gmt::tile_map_make(ti_n, tj_n, lx, hx, ly, hy, nat != 0);
          const long n = (long)ti_n * tj_n;
          std::set<std::pair<long, long>> seen;
          std::vector<std::pair<long, long>> order;
          for (long c = 0; c < n; ++c) {
            int64_t ti = -1, tj = -1;
            gmt::decode_tile(M, c, ti, tj);
            if (ti < 0 || ti >= ti_n || tj < 0 || tj >= tj_n) return fail("tile out of range", ti_n, tj_n, has, nat);
            if (!seen.insert({ti, tj}).second) return fail("tile visited twice", ti_n, tj_n, has, nat);
            order.push_back({ti, tj});
          }
          // 1. every tile exactly once (size check completes the bijection)
          if ((long)seen.size() != n) return fail("tile missing", ti_n, tj_n, has, nat);
          // 2. a tile that needs no flag never comes after one that does, except inside the interior block
          //    when x-face tiles are kept in memory order (natural)
          auto waits = [&](long ti, long tj) {
            return (lx && ti == 0) || (hx && ti == ti_n - 1) || (ly && tj == 0) || (hy && tj == tj_n - 1);
          };
          auto waits_y = [&](long tj) { return (ly && tj == 0) || (hy && tj == tj_n - 1); };
          if (M.n_int > 0) {
            bool seen_wait = false;
            for (auto& t : order) {
              const bool w = nat ? waits_y(t.second) : waits(t.first, t.second);
              if (w) seen_wait = true;
              else if (seen_wait) return fail("free tile after a waiting tile", ti_n, tj_n, has, nat);
            }
          }
          // 3. with natural order and no y neighbours the order is plain memory order
          if (nat && !ly && !hy)
            for (long c = 0; c < n; ++c)
              if (order[c].first != c % ti_n || order[c].second != c / ti_n)
                return fail("natural order is not memory order", ti_n, tj_n, has, nat);
          ++checked;
        }
  std::printf("ok %ld configurations\n", checked);
  return 0;
}
