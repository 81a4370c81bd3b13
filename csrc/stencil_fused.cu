// stencil_fused.cu -- halo exchange fused into the stencil sweep, over peer memory.
//
// Replaces, from the reference:
//   boundary_exchange + stencil1d_5          mpi_stencil_gt.cc:83-122,206
//   boundary_exchange_x/_y + stencil2d_1d_5  mpi_stencil2d_gt.cc:136-373,530-532
//   (SYCL twins)                             mpi_stencil2d_sycl.cc:212-375, mpi_stencil2d_sycl_oo.cc:363-515
// and adds the 4-neighbour process-grid exchange of BASELINE config 3.
//
// The reference packs strips into staging buffers, hands them to MPI, waits, unpacks, and
// brackets all of it with device-wide syncs before the stencil may start.  Here one
// persistent kernel per rank does everything:
//   phase 1  a few CTAs store this rank's boundary values straight into the neighbours'
//            inboxes (ordinary stores on NVLink-mapped addresses), chunk by chunk in the order
//            the neighbour will need them, each chunk followed by a system-scope fence and a
//            st.release.sys flag; all other CTAs start computing immediately;
//   phase 2  CTAs draw tiles from an atomic counter: interior tiles first (they depend on
//            nobody), boundary tiles last; a boundary tile polls its chunk flags
//            (ld.acquire.sys, bounded), copies the inbox into the array's own ghost cells
//            and only then runs the same tile code as the plain kernel.
// No pack buffer, no MPI/NCCL call, no host synchronisation; flags carry a monotonically
// increasing epoch and inboxes are double-buffered by epoch parity, so a fast neighbour
// that is already one launch ahead can never clobber ghosts that are still being read.
#include <algorithm>

#include "p2p.cuh"
#include "stencil_tiles.cuh"
#include "tile_order.h"
#include "tma.cuh"

namespace gmt {

// Test hook (tests/test_gpu_stress.py): when non-zero, every halo flag store is preceded by a
// pseudo-random sleep of up to this many nanoseconds, so that chunks arrive late, out of order
// and at different times on different ranks -- the race test the reference never needed because
// it synchronises the whole device around every exchange (mpi_stencil2d_gt.cc:202,254).
__device__ unsigned int g_debug_push_delay_ns = 0;
static unsigned int g_debug_push_delay_host = 0;  // launchers pick the kDebug instantiation when set

__device__ __forceinline__ void debug_delay(uint64_t salt)
{
  const unsigned int mx = g_debug_push_delay_ns;
  if (mx == 0) return;
  unsigned int h = (unsigned int)(salt * 2654435761ull) ^ (unsigned int)clock64();
  h ^= h >> 13;
  h *= 0x5bd1e995u;
  h ^= h >> 15;
  unsigned int ns = h % mx;
  while (ns > 0) {
    const unsigned int step = ns < 100000u ? ns : 100000u;
    __nanosleep(step);
    ns -= step;
  }
}

template <typename RankT>
__device__ __forceinline__ void load_rank(RankT& dst, const RankT* src)
{
  const uint64_t* s = reinterpret_cast<const uint64_t*>(src);
  uint64_t* d = reinterpret_cast<uint64_t*>(&dst);
  for (int w = threadIdx.x; w < (int)(sizeof(RankT) / 8); w += blockDim.x) d[w] = s[w];
  __syncthreads();
}

// the same, with the epoch fetched straight from global memory beside the descriptor copy (one level of
// dependent loads less: the exchange-only kernels are a few microseconds long and latency-bound)
template <typename RankT>
__device__ __forceinline__ uint64_t load_rank_and_epoch(RankT& dst, const RankT* src, uint64_t* epoch_smem)
{
  const uint64_t* s = reinterpret_cast<const uint64_t*>(src);
  uint64_t* d = reinterpret_cast<uint64_t*>(&dst);
  for (int w = threadIdx.x; w < (int)(sizeof(RankT) / 8); w += blockDim.x) d[w] = s[w];
  if (threadIdx.x == blockDim.x - 1) *epoch_smem = ld_relaxed_sys(src->epoch);
  __syncthreads();
  return *epoch_smem + 1;
}

// last CTA of this rank: reset the counters and advance the epoch
__device__ __forceinline__ void finish_epoch(uint32_t* done, uint64_t* epoch, uint64_t e, int n_counters)
{
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int t = atomicAdd(done, 1u);
    if (t == gridDim.x - 1) {
      for (int k = 0; k < n_counters; ++k) done[k] = 0u;
      __threadfence();
      st_relaxed_sys(epoch, e);
    }
  }
}

// ====================================================================== 1-D
template <bool kErr, bool kDebug>
__global__ void __launch_bounds__(kStThreads, 2) fused1d_kernel(const gmt_halo1d_rank* __restrict__ ranks)
{
  __shared__ gmt_halo1d_rank R;
  __shared__ double red[32];
  __shared__ int64_t next_tile[2];
  load_rank(R, ranks + blockIdx.y);
  const uint64_t e = ld_relaxed_sys(R.epoch) + 1;
  const int par = (int)(e & 1);
  const int tid = threadIdx.x;

  // 1. push my two boundary values to each neighbour (16 B each) and publish
  if (blockIdx.x == 0 && (tid == 0 || tid == 32)) {
    const int s = tid >> 5;  // 0: to the low neighbour, 1: to the high neighbour
    if (R.has[s]) {
      const double2 v = ldg128(s == 0 ? R.z + 2 : R.z + R.n);
      stg128(R.peer_inbox[s][par], v);
      fence_sys();
      if (kDebug) debug_delay(e * 131 + blockIdx.y * 7 + s);
      st_release_sys(R.peer_flag[s], e);
    }
  }

  Sten1d S{R.z, R.out, R.n, R.scale, R.x0, R.dx};
  // Work is drawn in order from the counter in done[1]: interior groups first -- they depend
  // on nobody -- then the group(s) holding the edge tiles that need a neighbour's values.
  // The draw for the NEXT group is issued before the current one is processed, so the
  // atomic's round trip to L2 is hidden behind the tile instead of stalling the CTA.
  constexpr int kGrab = 1;
  const int64_t ntiles = (R.n + kTile1d - 1) / kTile1d;
  const int64_t ngroups = (ntiles + kGrab - 1) / kGrab;
  const int64_t first_int = R.has[0] ? 1 : 0;
  const int64_t end_int = max(first_int, ngroups - (R.has[1] ? 1 : 0));
  const int64_t n_int = end_int - first_int;
  uint32_t* counter = R.done + 1;
  double err = 0.0;
  if (tid == 0) next_tile[0] = (int64_t)atomicAdd(counter, 1u);
  __syncthreads();
  int64_t c = next_tile[0];
  int slot = 0;
  for (;;) {
    if (c >= ngroups) break;
    unsigned int prefetched = 0;
    if (tid == 0) prefetched = atomicAdd(counter, 1u);
    int64_t g;
    if (c < n_int) {
      g = first_int + c;
    } else {
      // edge groups in index order: those below first_int, then those from end_int up
      const int64_t k = c - n_int;
      g = k < first_int ? k : end_int + (k - first_int);
    }
    const bool lo = (g == 0 && R.has[0]), hi = (g == ngroups - 1 && R.has[1]);
    if (lo || hi) {
      // wait for the neighbour, land the ghosts in the array, then compute like any tile
      if (tid == 0 || tid == 32) {
        const int s = tid >> 5;
        if ((s == 0 && lo) || (s == 1 && hi)) {
          if (!wait_flag_ge(R.my_flag[s], e, R.spin_limit)) atomicExch(R.status, 1u + s);
          const double2 v = ld_peer_f64x2(R.my_inbox[s][par]);
          stg128(s == 0 ? R.z : R.z + R.n + 2, v);
        }
      }
      __syncthreads();
    }
    for (int q = 0; q < kGrab; ++q) {
      const int64_t t = g * kGrab + q;
      if (t < ntiles) tile1d<kErr>(S, t * kTile1d, tid, err);
    }
    // hand the prefetched draw to the CTA through the other slot (its last readers are a
    // barrier behind), one barrier per tile
    if (tid == 0) next_tile[slot ^ 1] = (int64_t)prefetched;
    __syncthreads();
    slot ^= 1;
    c = next_tile[slot];
  }
  // 4. epilogue: error norm, then advance the epoch once every CTA of this rank is done
  if (kErr) {
    const double part = block_sum(err, red);
    double total;
    if (grid_sum_last_block(part, R.err_scratch, red, total, gridDim.x, blockIdx.x)) {
      if (tid == 0) *R.err_sumsq = total;
    }
  }
  finish_epoch(R.done, R.epoch, e, 2);
}

// ---------------- export of the boundary chunks.  x faces: kFlagCols columns x 2 rows
// (strided 16-byte pieces gathered into a contiguous inbox); y faces: kTileI rows x 2 columns.
// Only the first kPushCtas CTAs export (then join the tile loop late); everybody else starts
// computing at once.  Chunks go out in the order the neighbour's tile loop will want them,
// each followed by its own fence + flag, so the exchange streams underneath the sweep: the
// export front moves ~10x faster than the compute front.
// x-face chunks can be exported in groups of `group` chunks per system fence ($GMT_PUSH_GROUP, dim-0
// kernel).  Measured on 2 GPUs with the reference's dim-0 tile (2 x 512 chunks, tools/dim0_variants.py,
// profiles/dim0_variants_n2.jsonl): 1 chunk per fence 1.302 ms, 4: 1.308, 8: 1.319, 15: 1.353 -- the
// fences are NOT what the exchange costs there; later flags are.  The default stays one fence per chunk.
template <int GX, int GY, bool kDebug, bool kNamedBar = false>
__device__ __noinline__ void push_faces(const gmt_halo2d_rank& R, uint64_t e, int par, int64_t tiles_i,
                                           int max_push_ctas = 32, int group = 0)
{
  const int tid = threadIdx.x;
  const int64_t nx = R.nx, ny = R.ny, ldz = R.ldz;
  const int64_t nfx = GX ? (ny + kFlagCols - 1) / kFlagCols : 0;
  const int64_t nfy = GY ? tiles_i : 0;
  const int grp = group > 0 ? group : 1;
  const int64_t ngx = (nfx + grp - 1) / grp;
  const int64_t nunits = 2 * (ngx + nfy);
  const int npush = min((int)gridDim.x, max_push_ctas);
  if ((int)blockIdx.x >= npush) return;
  auto cta_bar = [&]() {
    if (kNamedBar)
      asm volatile("bar.sync 1, %0;" ::"n"(kStThreads) : "memory");  // the 256 exporting threads only
    else
      __syncthreads();
  };
  // unit u -> (side, first chunk); units of the two sides of a direction alternate
  auto decode = [&](int64_t u, int& side, int64_t& c) {
    if (u < 2 * ngx) {
      side = (int)(u & 1);
      c = (u >> 1) * grp;
    } else {
      side = 2 + (int)((u - 2 * ngx) & 1);
      c = (u - 2 * ngx) >> 1;
    }
  };
  // this CTA's next unit at or after u whose side has a neighbour (has == 2: fed by the host, nothing to export)
  auto next_live = [&](int64_t u) {
    for (; u < nunits; u += npush) {
      int side;
      int64_t c;
      decode(u, side, c);
      if (R.has[side] == 1) break;
    }
    return u;
  };
  // one chunk = four 16-byte pieces per thread: an x-face chunk is kFlagCols columns x 2 rows (strided
  // pieces gathered into a contiguous inbox), a y-face chunk kTileI rows x 2 columns
  auto load_chunk = [&](int side, int64_t c, double2* v) {
    if (side < 2) {
      const double* src = R.z + (side == 0 ? 2 : nx) + GY * ldz;
      const int64_t j0 = c * kFlagCols;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t j = j0 + tid + q * kStThreads;
        if (j < ny) v[q] = ldg128(src + j * ldz);
      }
    } else {
      const int64_t i0 = c * kTileI, i1 = min(nx, i0 + (int64_t)kTileI);
      const double* src = R.z + GX + (side == 2 ? 2 : ny) * ldz;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t i = i0 + 2 * tid + (q & 1) * 2 * kStThreads;
        if (i < i1) v[q] = ldg128(src + i + (q >> 1) * ldz);
      }
    }
  };
  auto store_chunk = [&](int side, int64_t c, const double2* v) {
    double* box = R.peer_inbox[side][par];
    if (side < 2) {
      const int64_t j0 = c * kFlagCols;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t j = j0 + tid + q * kStThreads;
        if (j < ny) stg128(box + 2 * j, v[q]);
      }
    } else {
      const int64_t i0 = c * kTileI, i1 = min(nx, i0 + (int64_t)kTileI);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t i = i0 + 2 * tid + (q & 1) * 2 * kStThreads;
        if (i < i1) stg128(box + (q >> 1) * nx + i, v[q]);
      }
    }
  };
  static_assert(kFlagCols == 4 * kStThreads && kTileI == 4 * kStThreads, "a chunk is four pieces per thread");

  if (group == 0) {
    // software pipeline: the loads of the NEXT chunk are issued before the fence of this one, so their
    // latency runs concurrently with the round trip the fence waits for -- and no flag is delayed
    double2 cur[4], nxt[4];
    int64_t u = next_live(blockIdx.x);
    int side = 0;
    int64_t c = 0;
    if (u < nunits) {
      decode(u, side, c);
      load_chunk(side, c, cur);
    }
    while (u < nunits) {
      store_chunk(side, c, cur);
      const int64_t u2 = next_live(u + npush);
      int side2 = 0;
      int64_t c2 = 0;
      if (u2 < nunits) {
        decode(u2, side2, c2);
        load_chunk(side2, c2, nxt);
      }
      cta_bar();
      if (tid == 0) {
        fence_sys();
        if (kDebug) debug_delay(e * 8191 + (uint64_t)u * 31 + blockIdx.y);
        st_relaxed_sys(R.peer_flag[side] + c, e);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) cur[q] = nxt[q];
      u = u2, side = side2, c = c2;
    }
    return;
  }
  // several x-face chunks per fence (tuning knob only)
  for (int64_t u = next_live(blockIdx.x); u < nunits; u = next_live(u + npush)) {
    int side;
    int64_t c;
    decode(u, side, c);
    const int64_t c_end = side < 2 ? min(nfx, c + grp) : c + 1;
#pragma unroll 1
    for (int64_t cc = c; cc < c_end; ++cc) {
      double2 v[4];
      load_chunk(side, cc, v);
      store_chunk(side, cc, v);
    }
    cta_bar();
    if (tid == 0) {
      fence_sys();
      for (int64_t cc = c; cc < c_end; ++cc) {
        if (kDebug) debug_delay(e * 8191 + (uint64_t)(u + cc) * 31 + blockIdx.y);
        st_relaxed_sys(R.peer_flag[side] + cc, e);
      }
    }
  }
}

// ====================================================================== 2-D
// (TileMap / decode_tile: tile_order.h)

// kOneShot = false: persistent grid, tiles drawn from a counter -- needed when several virtual
//   ranks share one cooperative launch (every CTA must be resident for the spin-waits to be safe).
// kOneShot = true: one tile per CTA in blockIdx order (the plain kernels' fastest pattern);
//   opt-in (GMT_FUSED_ONESHOT=1) for launches that hold ONE rank.  Measured SLOWER than the
//   persistent mode for this kernel (16384^2: 0.779 vs 0.669 ms; dim 0: 1.388 vs 1.26 ms) -- the
//   interior-first tile order breaks the in-order front that makes one-shot grids fast, and the
//   per-CTA prologue (descriptor load, epoch read) is paid per tile.  Safe without
//   co-residency: the exporting CTAs are the lowest block indices, hence scheduled first, and
//   never wait; the CTAs that wait (boundary tiles, highest indices) wait only for ANOTHER GPU's
//   exporting CTAs, never for a later CTA of their own grid.
template <int GX, int GY, bool kErr, bool kDebug, bool kOneShot>
__global__ void __launch_bounds__(kStThreads, 2) fused2d_kernel(const gmt_halo2d_rank* __restrict__ ranks, int tj_flags)
{
  const int TJ = tj_flags & 0xFFFF;
  const bool natural = (tj_flags >> 16) & 1;   // x-face tiles in memory order (they wait in place)
  const int push_group = (tj_flags >> 17) & 1 ? 1 : 0;   // 1: plain export loop (A/B knob), 0: pipelined
  __shared__ gmt_halo2d_rank R;
  __shared__ double red[32];
  __shared__ int64_t next_tile[2];
  load_rank(R, ranks + blockIdx.y);
  const uint64_t e = ld_relaxed_sys(R.epoch) + 1;
  const int par = (int)(e & 1);
  const int tid = threadIdx.x;
  const int64_t nx = R.nx, ny = R.ny, ldz = R.ldz;
  // x-face tiles (first / last row-tile of every column block) stay where memory order puts them:
  // pulling them out leaves 8 KB holes in every column for a later pass and costs more bandwidth
  // than the few waits at the very start of the kernel (measured 0.6725 vs 0.6796 ms at 2 GPUs).
  // y-face tiles are whole column blocks (contiguous), so they are deferred to the end for free.
  const TileMap M = tile_map_make((nx + kTileI - 1) / kTileI, (ny + TJ - 1) / TJ, GX && R.has[0], GX && R.has[1],
                                  GY && R.has[2], GY && R.has[3], natural);
  const int64_t ntiles = M.tiles_i * M.tiles_j;

  push_faces<GX, GY, kDebug>(R, e, par, M.tiles_i, 32, push_group);

  // ---------------- phase 2: tiles from the shared counter
  Sten2d S{R.z, R.out, nx, ny, ldz, R.ldo, R.sx, R.sy, R.x0, R.y0, R.dx, R.dy};
  uint32_t* counter = R.done + 1;
  double err = 0.0;
  // (no prefetched draw here, unlike the 1-D kernel: 2-D tiles are 0.5-2 MB, the atomic's
  // latency is noise, and a tile pre-claimed by a still-busy CTA would lengthen the tail)
  for (int trip = 0;; ++trip) {
    int64_t c;
    if (kOneShot) {
      if (trip > 0) break;
      c = blockIdx.x;
    } else {
      __syncthreads();
      if (tid == 0) next_tile[0] = (int64_t)atomicAdd(counter, 1u);
      __syncthreads();
      c = next_tile[0];
    }
    if (c >= ntiles) break;
    int64_t ti, tj;
    decode_tile(M, c, ti, tj);
    const int64_t i0 = ti * kTileI, i1 = min(nx, i0 + (int64_t)kTileI);
    const int64_t j0 = tj * TJ, j1 = min(ny, j0 + (int64_t)TJ);
    int sides = 0;
    if (GX && ti == 0 && R.has[0]) sides |= 1;
    if (GX && ti == M.tiles_i - 1 && R.has[1]) sides |= 2;
    if (GY && tj == 0 && R.has[2]) sides |= 4;
    if (GY && tj == M.tiles_j - 1 && R.has[3]) sides |= 8;
    if (sides) {
      if (tid < 4 && ((sides >> tid) & 1)) {
        const uint64_t* f = R.my_flag[tid] + (tid < 2 ? j0 / kFlagCols : ti);
        if (!wait_flag_ge(f, e, R.spin_limit)) atomicExch(R.status, 1u + tid);
      }
      __syncthreads();
      if (sides & 3) {  // x-face ghost rows: inbox -> the array's own ghost cells
        for (int64_t j = j0 + tid; j < j1; j += kStThreads) {
          double* col = R.z + (j + GY) * ldz;
          if (sides & 1) stg128(col, ld_peer_f64x2(R.my_inbox[0][par] + 2 * j));
          if (sides & 2) stg128(col + nx + 2, ld_peer_f64x2(R.my_inbox[1][par] + 2 * j));
        }
      }
      if (sides & 12) {  // y-face ghost columns are landed in the array first (few tiles)
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          for (int64_t i = i0 + 2 * tid; i < i1; i += 2 * kStThreads) {
            if (sides & 4) stg128(R.z + (i + GX) + r * ldz, ld_peer_f64x2(R.my_inbox[2][par] + r * nx + i));
            if (sides & 8)
              stg128(R.z + (i + GX) + (ny + 2 + r) * ldz, ld_peer_f64x2(R.my_inbox[3][par] + r * nx + i));
          }
        }
      }
      __syncthreads();
    }
    tile2d<GX, GY, kErr>(S, i0, j0, j1, tid, err);
  }
  if (kErr && !kOneShot) {
    const double part = block_sum(err, red);
    double total;
    if (grid_sum_last_block(part, R.err_scratch, red, total, gridDim.x, blockIdx.x)) {
      if (tid == 0) *R.err_sumsq = total;
    }
  }
  if (kOneShot) {
    // thousands of CTAs: their partials go into an accumulator (two of them, used by alternating
    // steps, so the one for the next step can be cleared while this one is still being read)
    double* acc = reinterpret_cast<double*>(reinterpret_cast<char*>(R.err_scratch) + kScratchHeaderBytes);
    if (kErr) {
      const double part = block_sum(err, red);
      if (tid == 0) atomicAdd(acc + par, part);
    }
    __syncthreads();
    if (tid == 0) {
      __threadfence();
      const unsigned int t = atomicAdd(R.done, 1u);
      if (t == gridDim.x - 1) {
        __threadfence();
        if (kErr) {
          *R.err_sumsq = __ldcg(acc + par);
          acc[par ^ 1] = 0.0;
        }
        R.done[0] = 0u;
        R.done[1] = 0u;
        __threadfence();
        st_relaxed_sys(R.epoch, e);
      }
    }
    return;
  }
  finish_epoch(R.done, R.epoch, e, 2);
}

// ====================================================================== 2-D, dim-0 slabs
// In the reference's dim-0 test (1024 x 524288 tile, halos = two strided 2 x 524288 strips)
// EVERY tile touches both x faces, so the latency of "poll the chunk flag, fetch the ghost
// pairs, put them into the array" would sit on the critical path of every tile.  Here it is
// moved off it by warp specialisation: a ninth warp per CTA draws tiles from the shared
// counter kDepth ahead of the compute warps, does the flag wait (ld.acquire.sys) and lands the
// ghosts for each, then hands the tile over through an mbarrier ring.  The eight compute warps
// never execute a CTA barrier in the loop: each waits on the ring slot, sweeps its rows of the
// tile and releases the slot, so a fast warp can run ahead of a slow one.
constexpr int kWsDepth = 2;
constexpr int kWsMaxTj = 64;   // widest dim-0 tile whose ghost pairs fit the shared-memory hand-over

// kLand = true : the communication warp copies the inbox into the array's own ghost cells (the array
//                ends up as after the reference's boundary_exchange_x), the compute warps read them there;
// kLand = false: the inbox IS the ghost storage: the communication warp hands the
//                tile's ghost pairs over in shared memory and the two threads that own the face rows
//                patch them into their registers.  No strided 16-byte writes into the array (each one
//                a read-modify-write of a 32-byte sector under ECC), no re-read: `out` is identical,
//                the array's ghost cells are NOT refreshed (call the exchange-only kernel for that).
template <bool kErr, bool kDebug, bool kLand>
__global__ void __launch_bounds__(kStThreads + 32, 2)
  fused_dim0_ws_kernel(const gmt_halo2d_rank* __restrict__ ranks, int tj_knobs)
{
  const int TJ = tj_knobs & 0xFFFF;
  const int push_group = (tj_knobs >> 16) & 0xF;            // 0 = default rule
  const int push_ctas = ((tj_knobs >> 20) & 0xFF) * 8;      // 0 = default (32)
  __shared__ gmt_halo2d_rank R;
  __shared__ double red[32];
  __shared__ int64_t ring[kWsDepth];
  __shared__ __align__(8) uint64_t full[kWsDepth];
  __shared__ __align__(8) uint64_t empty[kWsDepth];
  __shared__ __align__(16) double2 ghost[kLand ? 1 : kWsDepth][2][kLand ? 1 : kWsMaxTj];
  load_rank(R, ranks + blockIdx.y);
  const uint64_t e = ld_relaxed_sys(R.epoch) + 1;
  const int par = (int)(e & 1);
  const int tid = threadIdx.x;
  const int64_t nx = R.nx, ny = R.ny, ldz = R.ldz;
  const int64_t tiles_i = (nx + kTileI - 1) / kTileI, tiles_j = (ny + TJ - 1) / TJ;
  const int64_t ntiles = tiles_i * tiles_j;
  if (tid == 0) {
    for (int s = 0; s < kWsDepth; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kStThreads / 32);
    }
    mbar_fence_init();
  }
  __syncthreads();
  uint32_t* counter = R.done + 1;
  double err = 0.0;

  if (tid >= kStThreads) {
    // ---------------- communication warp
    const int lane = tid - kStThreads;
    for (int k = 0;; ++k) {
      const int slot = k % kWsDepth;
      const uint32_t ph = (k / kWsDepth) & 1;
      if (lane == 0) mbar_wait(&empty[slot], ph ^ 1);
      __syncwarp();
      long long t = 0;
      if (lane == 0) t = (long long)atomicAdd(counter, 1u);
      t = __shfl_sync(0xffffffffu, t, 0);
      if (t >= ntiles) {
        if (lane == 0) {
          ring[slot] = -1;  // sentinel: the compute warps leave
          mbar_arrive(&full[slot]);
        }
        break;
      }
      const int64_t tj = t / tiles_i, ti = t - tj * tiles_i;
      const int64_t j0 = tj * TJ, j1 = min(ny, j0 + (int64_t)TJ);
      const int sides = ((ti == 0 && R.has[0]) ? 1 : 0) | ((ti == tiles_i - 1 && R.has[1]) ? 2 : 0);
      if (sides) {
        if (lane < 2 && ((sides >> lane) & 1)) {
          if (!wait_flag_ge(R.my_flag[lane] + j0 / kFlagCols, e, R.spin_limit)) atomicExch(R.status, 1u + lane);
        }
        __syncwarp();
        if (kLand) {
          for (int64_t j = j0 + lane; j < j1; j += 32) {
            double* col = R.z + j * ldz;
            if (sides & 1) stg128(col, ld_peer_f64x2(R.my_inbox[0][par] + 2 * j));
            if (sides & 2) stg128(col + nx + 2, ld_peer_f64x2(R.my_inbox[1][par] + 2 * j));
          }
        } else {
          // lanes 0-15 fetch the low face's pairs, lanes 16-31 the high face's
          const int sd = lane >> 4;
          if ((sides >> sd) & 1)
            for (int64_t j = j0 + (lane & 15); j < j1; j += 16)
              ghost[slot][sd][j - j0] = ld_peer_f64x2(R.my_inbox[sd][par] + 2 * j);
        }
        __syncwarp();
      }
      if (lane == 0) {
        ring[slot] = (int64_t)t;
        mbar_arrive(&full[slot]);  // release: ring entry and landed / handed-over ghosts are visible to the waiters
      }
    }
  } else {
    // ---------------- compute warps
    push_faces<2, 0, kDebug, true>(R, e, par, tiles_i, push_ctas > 0 ? push_ctas : 32, push_group);
    Sten2d S{R.z, R.out, nx, ny, ldz, R.ldo, R.sx, R.sy, R.x0, R.y0, R.dx, R.dy};
    for (int k = 0;; ++k) {
      const int slot = k % kWsDepth;
      const uint32_t ph = (k / kWsDepth) & 1;
      mbar_wait(&full[slot], ph);
      const int64_t t = ring[slot];
      if (t < 0) break;
      const int64_t tj = t / tiles_i, ti = t - tj * tiles_i;
      const int64_t j0 = tj * TJ;
      if (kLand) {
        tile2d<2, 0, kErr>(S, ti * kTileI, j0, min(ny, j0 + (int64_t)TJ), tid, err);
      } else {
        const double2* glo = (ti == 0 && R.has[0]) ? ghost[slot][0] : nullptr;
        const double2* ghi = (ti == tiles_i - 1 && R.has[1]) ? ghost[slot][1] : nullptr;
        tile2d_d0_inbox<kErr>(S, ti * kTileI, j0, min(ny, j0 + (int64_t)TJ), tid, err, glo, ghi);
      }
      __syncwarp();
      if ((tid & 31) == 0) mbar_arrive(&empty[slot]);
    }
  }
  __syncthreads();
  if (kErr) {
    const double part = block_sum(err, red);
    double total;
    if (grid_sum_last_block(part, R.err_scratch, red, total, gridDim.x, blockIdx.x)) {
      if (tid == 0) *R.err_sumsq = total;
    }
  }
  finish_epoch(R.done, R.epoch, e, 2);
}

// ====================================================================== exchange only
// The reference's central callable -- boundary_exchange (mpi_stencil_gt.cc:83-122),
// boundary_exchange_x / _y (mpi_stencil2d_gt.cc:136-373) -- and the quantity it times
// (mpi_stencil_gt.cc:200-204, mpi_stencil2d_gt.cc:512-522), as one small kernel per rank on the
// same inboxes, flags and epochs as the fused steps: export the boundary values into the
// neighbours' inboxes, wait for theirs, put them into the array's own ghost cells.  No sweep;
// afterwards the array is what it is after the reference's exchange, ready for ANY compute step.
// Exchange-only calls and fused steps can be mixed freely on one plan (each advances the epoch).
// Every CTA exports all of its share before it waits for anything, and waits only for another
// rank's exports, so the kernel cannot block itself whatever part of the grid is resident.
// 1-D: the two boundary values of a side travel as two 16-byte "LL" cells (p2p.cuh) that fill the
// side's 32-byte inbox exactly: data and "arrived" come in one store -- no fence, no separate flag, the
// receiver polls the cells themselves.  (The fused 1-D kernel keeps the plain-values + flag layout in
// the same inbox; both ranks of a pair always run the same kind of call in the same epoch.)
template <bool kDebug>
__global__ void __launch_bounds__(64) halo_exchange1d_kernel(const gmt_halo1d_rank* __restrict__ ranks)
{
  __shared__ gmt_halo1d_rank R;
  __shared__ uint64_t epoch0;
  const uint64_t e = load_rank_and_epoch(R, ranks + blockIdx.y, &epoch0);
  const uint64_t e32 = e & 0xffffffffull;
  const int par = (int)(e & 1);
  const int tid = threadIdx.x;
  if (tid == 0 || tid == 32) {
    const int s = tid >> 5;  // 0: low neighbour, 1: high neighbour
    if (R.has[s]) {
      const double2 v = ldg128(s == 0 ? R.z + 2 : R.z + R.n);
      uint64_t* out = reinterpret_cast<uint64_t*>(R.peer_inbox[s][par]);
      if (kDebug) debug_delay(e * 131 + blockIdx.y * 7 + s);
      st_ll_cell(out, v.x, e32);
      st_ll_cell(out + 2, v.y, e32);
      const uint64_t* in = reinterpret_cast<const uint64_t*>(R.my_inbox[s][par]);
      const int64_t limit = R.spin_limit > 0 ? R.spin_limit : kDefaultSpinLimit;
      double2 g = {0.0, 0.0};
      bool ok = false;
      for (int64_t it = 0; it < limit; ++it) {
        if (ld_ll_cell(in, e32, g.x) && ld_ll_cell(in + 2, e32, g.y)) {
          ok = true;
          break;
        }
        if (it > 64) __nanosleep(32);
      }
      if (!ok) atomicExch(R.status, 1u + s);
      stg128(s == 0 ? R.z : R.z + R.n + 2, g);
    }
  }
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    st_relaxed_sys(R.epoch, e);
  }
}

// 2-D.  Phase A: every CTA exports its share of the live faces -- all loads in flight, then all
// stores, ONE system fence per CTA, then the flags of everything it exported.  (The fused kernels flag
// chunk by chunk so that the exchange streams under the sweep; here nothing runs underneath, and one
// fence instead of one per chunk is what counts.)  Phase B: for its share of (side, chunk) units the CTA
// waits for the neighbour's flag and copies the inbox into the array's ghost cells, four 16-byte
// pieces per thread in flight.  Only live sides are enumerated, so all CTAs share the work even when
// a rank has a single neighbour.
template <int GX, int GY, bool kDebug>
__global__ void __launch_bounds__(kStThreads, 2) halo_exchange2d_kernel(const gmt_halo2d_rank* __restrict__ ranks)
{
  __shared__ gmt_halo2d_rank R;
  __shared__ uint64_t epoch0;
  const uint64_t e = load_rank_and_epoch(R, ranks + blockIdx.y, &epoch0);
  const int par = (int)(e & 1);
  const int tid = threadIdx.x;
  const int64_t nx = R.nx, ny = R.ny, ldz = R.ldz;
  const int64_t tiles_i = (nx + kTileI - 1) / kTileI;
  const int64_t nfx = GX ? (ny + kFlagCols - 1) / kFlagCols : 0;
  const int64_t nfy = GY ? tiles_i : 0;
  // live sides: out[] = sides I export to (a neighbour), in[] = sides I receive on (neighbour or host)
  int out_side[4], in_side[4], n_out = 0, n_in = 0;
#pragma unroll
  for (int sd = 0; sd < 4; ++sd) {
    if (sd < 2 ? !GX : !GY) continue;
    if (R.has[sd] == 1) out_side[n_out++] = sd;
    if (R.has[sd] != 0) in_side[n_in++] = sd;
  }
  auto chunks_of = [&](int sd) { return sd < 2 ? nfx : nfy; };
  // units of a side list: side-major blocks, [0, chunks(side0)), then side1's, ...
  auto decode = [&](const int* sides, int n_sides, int64_t u, int& sd, int64_t& c) {
    for (int k = 0; k < n_sides; ++k) {
      const int64_t nc = chunks_of(sides[k]);
      if (u < nc) {
        sd = sides[k];
        c = u;
        return;
      }
      u -= nc;
    }
    sd = -1;
    c = 0;
  };
  int64_t total_out = 0, total_in = 0;
  for (int k = 0; k < n_out; ++k) total_out += chunks_of(out_side[k]);
  for (int k = 0; k < n_in; ++k) total_in += chunks_of(in_side[k]);

  // ---------------- phase A: export
  for (int64_t u = blockIdx.x; u < total_out; u += gridDim.x) {
    int sd;
    int64_t c;
    decode(out_side, n_out, u, sd, c);
    double* box = R.peer_inbox[sd][par];
    double2 v[4];
    if (sd < 2) {
      const int64_t j0 = c * kFlagCols;
      const double* src = R.z + (sd == 0 ? 2 : nx) + GY * ldz;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t j = j0 + tid + q * kStThreads;
        if (j < ny) v[q] = ldg128(src + j * ldz);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t j = j0 + tid + q * kStThreads;
        if (j < ny) stg128(box + 2 * j, v[q]);
      }
    } else {
      const int64_t i0 = c * kTileI, i1 = min(nx, i0 + (int64_t)kTileI);
      const double* src = R.z + GX + (sd == 2 ? 2 : ny) * ldz;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t i = i0 + 2 * tid + (q & 1) * 2 * kStThreads;
        if (i < i1) v[q] = ldg128(src + i + (q >> 1) * ldz);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t i = i0 + 2 * tid + (q & 1) * 2 * kStThreads;
        if (i < i1) stg128(box + (q >> 1) * nx + i, v[q]);
      }
    }
  }
  __syncthreads();
  if (tid == 0 && (int64_t)blockIdx.x < total_out) {
    fence_sys();   // one fence: every store of this CTA above is ordered before every flag below
    for (int64_t u = blockIdx.x; u < total_out; u += gridDim.x) {
      int sd;
      int64_t c;
      decode(out_side, n_out, u, sd, c);
      if (kDebug) debug_delay(e * 8191 + (uint64_t)u * 31 + blockIdx.y);
      st_relaxed_sys(R.peer_flag[sd] + c, e);
    }
  }

  // ---------------- phase B: land
  for (int64_t u = blockIdx.x; u < total_in; u += gridDim.x) {
    int sd;
    int64_t c;
    decode(in_side, n_in, u, sd, c);
    if (tid == 0) {
      if (!wait_flag_ge(R.my_flag[sd] + c, e, R.spin_limit)) atomicExch(R.status, 1u + sd);
    }
    __syncthreads();   // the acquire of thread 0 + this barrier order the plain loads below after the flag
    const double* box = R.my_inbox[sd][par];
    double2 v[4];
    if (sd < 2) {
      const int64_t j0 = c * kFlagCols;
      double* dst = R.z + (sd == 0 ? 0 : nx + 2) + GY * ldz;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t j = j0 + tid + q * kStThreads;
        if (j < ny) v[q] = ld_peer_f64x2(box + 2 * j);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t j = j0 + tid + q * kStThreads;
        if (j < ny) stg128(dst + j * ldz, v[q]);
      }
    } else {
      const int64_t i0 = c * kTileI, i1 = min(nx, i0 + (int64_t)kTileI);
      double* dst = R.z + GX + (sd == 2 ? 0 : ny + 2) * ldz;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t i = i0 + 2 * tid + (q & 1) * 2 * kStThreads;
        if (i < i1) v[q] = ld_peer_f64x2(box + (q >> 1) * nx + i);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t i = i0 + 2 * tid + (q & 1) * 2 * kStThreads;
        if (i < i1) stg128(dst + i + (q >> 1) * ldz, v[q]);
      }
    }
    __syncthreads();
  }
  finish_epoch(R.done, R.epoch, e, 2);
}

// flags[i] <- *epoch + 1 with release semantics: "the host-fed ghost values of the step now
// running are in the inbox".  Runs on the stream that carried the host-to-device copy, next to
// the fused kernel of that step (which bumps *epoch only when it finishes, i.e. after it has
// seen these flags).
__global__ void publish_flags_kernel(uint64_t* flags, int64_t n, const uint64_t* epoch)
{
  const uint64_t e = ld_relaxed_sys(epoch) + 1;
  __threadfence_system();
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) st_release_sys(flags + i, e);
}

template <typename RankT>
static int launch_fused(const void* kernel, const RankT* ranks_dev, int n_ranks, int64_t ntiles,
                        cudaStream_t st, int tile_j_arg = 0, int threads = kStThreads, bool oneshot = false)
{
  int grid = 0;
  if (oneshot) {
    GMT_REQUIRE(ntiles < (1LL << 31), GMT_ETOOBIG, "fused stencil: array too large for a one-shot grid");
    grid = (int)std::max<int64_t>(1, ntiles);
  } else {
    int rc = persistent_grid(kernel, threads, ntiles, n_ranks, &grid);
    if (rc) return rc;
  }
  void* args[] = {(void*)&ranks_dev, (void*)&tile_j_arg};
  if (n_ranks > 1) {
    // several ranks spin on one another inside one launch: co-residency must be guaranteed
    GMT_CUDA(cudaLaunchCooperativeKernel(kernel, dim3(grid, n_ranks), dim3(threads), args, 0, st));
  } else {
    GMT_CUDA(cudaLaunchKernel(kernel, dim3(grid, 1), dim3(threads), args, 0, st));
  }
  return 0;
}

}  // namespace gmt

extern "C" {

int gmt_fused_stencil1d_f64(const gmt_halo1d_rank* ranks_dev, int n_ranks, int64_t n, int with_err,
                            gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(ranks_dev && n_ranks >= 1 && n > 0, GMT_EINVAL, "gmt_fused_stencil1d_f64: bad argument");
  GMT_REQUIRE(n % 2 == 0, GMT_EUNSUPPORTED, "gmt_fused_stencil1d_f64: n must be even (16-byte halo stores)");
  const bool dbg = g_debug_push_delay_host != 0;
  const void* k = dbg ? (with_err ? (const void*)fused1d_kernel<true, true> : (const void*)fused1d_kernel<false, true>)
                      : (with_err ? (const void*)fused1d_kernel<true, false> : (const void*)fused1d_kernel<false, false>);
  return launch_fused(k, ranks_dev, n_ranks, (n + 4 * kTile1d - 1) / (4 * kTile1d), as_stream(stream));
}

int gmt_fused_stencil2d_f64(const gmt_halo2d_rank* ranks_dev, int n_ranks, int64_t nx, int64_t ny,
                            int gx, int gy, int with_err_flags, gmt_stream_t stream)
{
  using namespace gmt;
  const int with_err = with_err_flags & 1;
  const bool from_inbox = (with_err_flags & 2) != 0;   // dim-0 slabs: ghosts consumed straight from the inbox
  GMT_REQUIRE(ranks_dev && n_ranks >= 1 && nx > 0 && ny > 0, GMT_EINVAL, "gmt_fused_stencil2d_f64: bad argument");
  GMT_REQUIRE((gx == 0 || gx == 2) && (gy == 0 || gy == 2) && (gx || gy), GMT_EINVAL,
              "gmt_fused_stencil2d_f64: ghost widths must be 0 or 2 and not both 0");
  GMT_REQUIRE(nx % 4 == 0, GMT_EUNSUPPORTED, "gmt_fused_stencil2d_f64: nx must be a multiple of 4");
  const void* k = nullptr;
  const int64_t tiles_i = (nx + kTileI - 1) / kTileI;
  // one rank per launch (production): one-shot grid with the plain kernels' tile widths
  const char* os_env = getenv("GMT_FUSED_ONESHOT");
  const bool oneshot = n_ranks == 1 && os_env && os_env[0] == '1';  // opt-in: measured slower (0.779 vs 0.669 ms)
  const int tj = runtime_tile_j(gx, gy, !oneshot);
  int64_t nt = 0;
  const bool dbg = g_debug_push_delay_host != 0;
#define GMT_PICK1(GX_, GY_, OS_)                                                                             \
  (dbg ? (with_err ? (const void*)fused2d_kernel<GX_, GY_, true, true, OS_> : (const void*)fused2d_kernel<GX_, GY_, false, true, OS_>) \
       : (with_err ? (const void*)fused2d_kernel<GX_, GY_, true, false, OS_> : (const void*)fused2d_kernel<GX_, GY_, false, false, OS_>))
#define GMT_PICK(GX_, GY_) (oneshot ? GMT_PICK1(GX_, GY_, true) : GMT_PICK1(GX_, GY_, false))
  int threads = kStThreads;
  if (gx && !gy && oneshot) {
    k = GMT_PICK(2, 0);
    nt = tiles_i * ((ny + tj - 1) / tj);
  } else if (gx && !gy) {
    if (from_inbox && tj <= kWsMaxTj)
      k = dbg ? (with_err ? (const void*)fused_dim0_ws_kernel<true, true, false> : (const void*)fused_dim0_ws_kernel<false, true, false>)
              : (with_err ? (const void*)fused_dim0_ws_kernel<true, false, false> : (const void*)fused_dim0_ws_kernel<false, false, false>);
    else
      k = dbg ? (with_err ? (const void*)fused_dim0_ws_kernel<true, true, true> : (const void*)fused_dim0_ws_kernel<false, true, true>)
              : (with_err ? (const void*)fused_dim0_ws_kernel<true, false, true> : (const void*)fused_dim0_ws_kernel<false, false, true>);
    threads = kStThreads + 32;
    nt = tiles_i * ((ny + tj - 1) / tj);
  } else if (!gx && gy) {
    k = GMT_PICK(0, 2);
    nt = tiles_i * ((ny + tj - 1) / tj);
  } else {
    k = GMT_PICK(2, 2);
    nt = tiles_i * ((ny + tj - 1) / tj);
  }
#undef GMT_PICK
#undef GMT_PICK1
  // tile order of the general kernel: y-face column blocks always come last; x-face tiles stay in
  // memory order by default (GMT_FUSED_ORDER=interior moves them to the end as well)
  const char* ord = getenv("GMT_FUSED_ORDER");
  const bool ws = gx && !gy && !oneshot;
  int tj_flags = tj | ((!ws && !(ord && ord[0] == 'i')) ? (1 << 16) : 0);
  {
    const char* pp = getenv("GMT_PUSH_PIPELINE");
    if (!ws && pp && pp[0] == '0') tj_flags |= 1 << 17;   // A/B knob: export without the software pipeline
  }
  if (ws) {   // tuning knobs of the dim-0 export (tools/dim0_variants.py): chunks per fence, exporting CTAs
    const char* pg = getenv("GMT_PUSH_GROUP");
    const char* pc = getenv("GMT_PUSH_CTAS");
    const int g = pg ? atoi(pg) : 0, c = pc ? atoi(pc) : 0;
    if (g >= 1 && g <= 15) tj_flags |= g << 16;
    if (c >= 8 && c <= 2040) tj_flags |= (c / 8) << 20;
  }
  return launch_fused(k, ranks_dev, n_ranks, nt, as_stream(stream), tj_flags, threads, oneshot);
}

int gmt_halo_exchange1d_f64(const gmt_halo1d_rank* ranks_dev, int n_ranks, gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(ranks_dev && n_ranks >= 1, GMT_EINVAL, "gmt_halo_exchange1d_f64: bad argument");
  const void* k = g_debug_push_delay_host ? (const void*)halo_exchange1d_kernel<true> : (const void*)halo_exchange1d_kernel<false>;
  void* args[] = {(void*)&ranks_dev};
  if (n_ranks > 1)
    GMT_CUDA(cudaLaunchCooperativeKernel(k, dim3(1, n_ranks), dim3(64), args, 0, as_stream(stream)));
  else
    GMT_CUDA(cudaLaunchKernel(k, dim3(1, 1), dim3(64), args, 0, as_stream(stream)));
  return 0;
}

int gmt_halo_exchange2d_f64(const gmt_halo2d_rank* ranks_dev, int n_ranks, int64_t nx, int64_t ny, int gx, int gy,
                            gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(ranks_dev && n_ranks >= 1 && nx > 0 && ny > 0, GMT_EINVAL, "gmt_halo_exchange2d_f64: bad argument");
  GMT_REQUIRE((gx == 0 || gx == 2) && (gy == 0 || gy == 2) && (gx || gy), GMT_EINVAL,
              "gmt_halo_exchange2d_f64: ghost widths must be 0 or 2 and not both 0");
  GMT_REQUIRE(nx % 4 == 0, GMT_EUNSUPPORTED, "gmt_halo_exchange2d_f64: nx must be a multiple of 4");
  const bool dbg = g_debug_push_delay_host != 0;
  const void* k = nullptr;
#define GMT_PICKX(GX_, GY_) (dbg ? (const void*)halo_exchange2d_kernel<GX_, GY_, true> : (const void*)halo_exchange2d_kernel<GX_, GY_, false>)
  if (gx && gy) k = GMT_PICKX(2, 2);
  else if (gx) k = GMT_PICKX(2, 0);
  else k = GMT_PICKX(0, 2);
#undef GMT_PICKX
  // one landing unit per (side, 1024-column chunk | 1024-row block); a face of a few chunks needs a
  // few CTAs, the reference's 8 MiB strided faces (2 x 512 chunks) fill the GPU
  const int64_t units = 2 * ((gx ? gmt_halo2d_flag_count(0, nx, ny) : 0) + (gy ? gmt_halo2d_flag_count(2, nx, ny) : 0));
  int grid = 0;
  int rc = persistent_grid(k, kStThreads, units, n_ranks, &grid);
  if (rc) return rc;
  void* args[] = {(void*)&ranks_dev};
  if (n_ranks > 1)
    GMT_CUDA(cudaLaunchCooperativeKernel(k, dim3(grid, n_ranks), dim3(kStThreads), args, 0, as_stream(stream)));
  else
    GMT_CUDA(cudaLaunchKernel(k, dim3(grid, 1), dim3(kStThreads), args, 0, as_stream(stream)));
  return 0;
}

int gmt_publish_epoch_flags(uint64_t* flags, int64_t n, const uint64_t* epoch, gmt_stream_t stream)
{
  GMT_REQUIRE(flags && epoch && n >= 0, GMT_EINVAL, "gmt_publish_epoch_flags: bad argument");
  if (n == 0) return 0;
  gmt::publish_flags_kernel<<<1, 256, 0, gmt::as_stream(stream)>>>(flags, n, epoch);
  GMT_LAUNCH_CHECK();
  return 0;
}

int gmt_debug_set_push_delay(unsigned int max_ns)
{
  GMT_CUDA(cudaMemcpyToSymbol(gmt::g_debug_push_delay_ns, &max_ns, sizeof(max_ns)));
  gmt::g_debug_push_delay_host = max_ns;
  return 0;
}

int64_t gmt_halo2d_flag_count(int side, int64_t nx, int64_t ny)
{
  if (side < 2) return (ny + gmt::kFlagCols - 1) / gmt::kFlagCols;
  return (nx + gmt::kTileI - 1) / gmt::kTileI;
}

}  // extern "C"
