// stencil_tma.cu -- the stencils fed by the Tensor Memory Accelerator (variant = "tma").
//
// Same arithmetic as stencil1d_5 (mpi_stencil_gt.cc:54-59) and stencil2d_1d_5_d1
// (mpi_stencil2d_gt.cc:101-110); different data path:
//
//  1-D  a producer thread draws tiles in address order from an atomic counter and streams
//       (tile + 4)-point windows into a ring of shared-memory stages with cp.async.bulk
//       (SASS: UBLKCP) and mbarrier transaction counts; eight consumer warps read each window
//       from shared memory (the 4-point overlap between neighbouring threads costs LDS
//       bandwidth instead of L1/LSU traffic) and store 256-bit results.
//
//  2-D  (derivative along dim 1, the strided axis) thread-block CLUSTERS of two CTAs: each
//       CTA loads the 8 columns only it needs with 2-D tensor-map copies (cp.async.bulk.tensor,
//       SASS: UTMALDG) and its half of the 4 halo columns BOTH tiles need with a MULTICAST
//       copy that lands in both CTAs' shared memory and completes on both CTAs' mbarriers,
//       so the halo between the two tiles is fetched once and shared through SMEM.  Two
//       mbarriers per stage: bytes landed (full) and both CTAs done reading (empty, local +
//       remote relaxed arrives).  No thread-side fence anywhere in the loop.
//
// Measured on B200 these are NOT faster than the LSU kernels in stencil.cu for this
// 16 B/point stream, so they are opt-in variants, kept as the TMA/cluster
// reference implementation and cross-checked by the same tests.
#include <cuda.h>

#include <algorithm>

#include "common.cuh"
#include "p2p.cuh"
#include "tma.cuh"

namespace gmt {

// ===================================================================== 1-D
constexpr int kTmaTile = 4096;  // outputs per stage
constexpr int kTmaStages = 3;
constexpr int kTmaConsumers = 256;
constexpr int kTmaStageBytes = ((kTmaTile + 4) * 8 + 127) / 128 * 128;
constexpr int kTmaSmemBytes = 256 + kTmaStages * kTmaStageBytes;

__global__ void __launch_bounds__(kTmaConsumers + 32)
  stencil1d_tma_kernel(int64_t n, const double* __restrict__ in, double* __restrict__ out, double scale,
                       unsigned int* counters)
{
  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty = full + kTmaStages;
  int64_t* tile_of = reinterpret_cast<int64_t*>(smem + 128);
  unsigned char* data = smem + 256;
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
      // producer: tiles in address order from the shared counter (an in-order front streams
      // HBM faster than a grid-stride assignment, see daxpy.cu)
      for (int k = 0;; ++k) {
        const int64_t t = (int64_t)atomicAdd(counters, 1u);
        const int s = k % kTmaStages;
        const uint32_t ph = (k / kTmaStages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        if (t >= ntiles) {
          tile_of[s] = -1;  // sentinel: consumers leave
          mbar_arrive(&full[s]);
          break;
        }
        tile_of[s] = t;
        const int64_t t0 = t * kTmaTile;
        const int64_t cnt = min((int64_t)kTmaTile, n - t0) + 4;
        const uint32_t bytes = (uint32_t)(cnt * 8);
        mbar_arrive_expect_tx(&full[s], bytes);
        bulk_g2s(data + s * kTmaStageBytes, in + t0, bytes, &full[s]);
      }
      if (atomicAdd(counters + 1, 1u) == gridDim.x - 1) {  // last CTA out resets the counters
        counters[0] = 0u;
        counters[1] = 0u;
      }
    }
    return;
  }

  for (int k = 0;; ++k) {
    const int s = k % kTmaStages;
    const uint32_t ph = (k / kTmaStages) & 1;
    mbar_wait(&full[s], ph);
    const int64_t t = tile_of[s];
    if (t < 0) break;
    const int64_t t0 = t * kTmaTile;
    const double* w = reinterpret_cast<const double*>(data + s * kTmaStageBytes);
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

// ===================================================================== 1-D, fused halo exchange
// The TMA pipeline above with the halo exchange of gmt_fused_stencil1d_f64 folded in (replaces
// boundary_exchange + stencil1d_5, mpi_stencil_gt.cc:83-122,206): the producer thread of CTA 0 stores
// the rank's two boundary values per side into the neighbours' inboxes and flags them; every
// producer draws tiles interior-first, and the one that draws an edge tile first waits for the
// neighbour's flag, puts the two ghost values into the array, makes them visible to the async proxy
// (fence.proxy.async) and only then issues the bulk copy of that window.  The consumer warps never
// see any of it.  Same descriptors, inboxes, flags and epoch as the LSU kernel (calls can be mixed).
__global__ void __launch_bounds__(kTmaConsumers + 32)
  fused1d_tma_kernel(const gmt_halo1d_rank* __restrict__ ranks)
{
  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty = full + kTmaStages;
  int64_t* tile_of = reinterpret_cast<int64_t*>(smem + 128);
  unsigned char* data = smem + 256;
  __shared__ gmt_halo1d_rank R;
  const int tid = threadIdx.x;
  {
    const uint64_t* src = reinterpret_cast<const uint64_t*>(ranks + blockIdx.y);
    uint64_t* dst = reinterpret_cast<uint64_t*>(&R);
    for (int w = tid; w < (int)(sizeof(gmt_halo1d_rank) / 8); w += blockDim.x) dst[w] = src[w];
  }
  if (tid == 0) {
    for (int s = 0; s < kTmaStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kTmaConsumers / 32);
    }
    mbar_fence_init();
  }
  __syncthreads();
  const uint64_t e = ld_relaxed_sys(R.epoch) + 1;
  const int par = (int)(e & 1);
  const int64_t n = R.n;
  const double* __restrict__ in = R.z;
  double* __restrict__ out = R.out;
  const double scale = R.scale;
  const int64_t ntiles = (n + kTmaTile - 1) / kTmaTile;
  const int64_t first_int = R.has[0] ? 1 : 0;
  const int64_t end_int = max(first_int, ntiles - (R.has[1] ? 1 : 0));
  const int64_t n_int = end_int - first_int;

  if (tid >= kTmaConsumers) {
    if (tid == kTmaConsumers) {
      if (blockIdx.x == 0) {   // export: 16 bytes per side, one fence, two flags
        bool any = false;
        for (int s = 0; s < 2; ++s)
          if (R.has[s]) {
            stg128(R.peer_inbox[s][par], ldg128(s == 0 ? R.z + 2 : R.z + n));
            any = true;
          }
        if (any) fence_sys();
        for (int s = 0; s < 2; ++s)
          if (R.has[s]) st_relaxed_sys(R.peer_flag[s], e);
      }
      uint32_t* counter = R.done + 1;
      for (int k = 0;; ++k) {
        const int64_t c = (int64_t)atomicAdd(counter, 1u);
        const int s = k % kTmaStages;
        const uint32_t ph = (k / kTmaStages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        if (c >= ntiles) {
          tile_of[s] = -1;  // sentinel: consumers leave
          mbar_arrive(&full[s]);
          break;
        }
        // interior tiles first (they depend on nobody), then the edge tiles in index order
        int64_t t;
        if (c < n_int) {
          t = first_int + c;
        } else {
          const int64_t q = c - n_int;
          t = q < first_int ? q : end_int + (q - first_int);
        }
        const bool lo = (t == 0 && R.has[0]), hi = (t == ntiles - 1 && R.has[1]);
        if (lo || hi) {
          for (int sd = 0; sd < 2; ++sd) {
            if (!(sd == 0 ? lo : hi)) continue;
            if (!wait_flag_ge(R.my_flag[sd], e, R.spin_limit)) atomicExch(R.status, 1u + sd);
            stg128(sd == 0 ? R.z : R.z + n + 2, ld_peer_f64x2(R.my_inbox[sd][par]));
          }
          __threadfence();
          asm volatile("fence.proxy.async.global;" ::: "memory");  // my generic-proxy stores before the bulk copy's reads
        }
        tile_of[s] = t;
        const int64_t t0 = t * kTmaTile;
        const int64_t cnt = min((int64_t)kTmaTile, n - t0) + 4;
        const uint32_t bytes = (uint32_t)(cnt * 8);
        mbar_arrive_expect_tx(&full[s], bytes);
        bulk_g2s(data + s * kTmaStageBytes, in + t0, bytes, &full[s]);
      }
    }
  } else {
    for (int k = 0;; ++k) {
      const int s = k % kTmaStages;
      const uint32_t ph = (k / kTmaStages) & 1;
      mbar_wait(&full[s], ph);
      const int64_t t = tile_of[s];
      if (t < 0) break;
      const int64_t t0 = t * kTmaTile;
      const double* w = reinterpret_cast<const double*>(data + s * kTmaStageBytes);
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
  // last CTA of this rank: reset the counters and advance the epoch
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    if (atomicAdd(R.done, 1u) == gridDim.x - 1) {
      R.done[0] = 0u;
      R.done[1] = 0u;
      __threadfence();
      st_relaxed_sys(R.epoch, e);
    }
  }
}

// ===================================================================== 2-D, clusters of 2
// Work item of a cluster: rows [i0, i0+1024) x output columns [j0, j0+16).  The 20 ghosted
// columns it reads are split as: CTA0 alone 0..7, shared 8..11, CTA1 alone 12..19.  Each CTA
// loads its own 8 columns (unicast) and HALF of the shared halo (2 columns) with a MULTICAST
// tensor copy that lands in both CTAs' shared memory at the same offset and signals both
// CTAs' mbarriers -- the halo between the two tiles is fetched from L2/HBM exactly once and
// no thread ever issues a fence: the only thread-side signalling is the relaxed "stage free"
// arrive (local + remote).
constexpr int kC2Rows = 1024;             // rows per tile = 4 TMA boxes of 256 rows
constexpr int kC2Box = 256;               // TMA box extent along dim 0 (hardware limit 256)
constexpr int kC2Ty = 8;                  // output columns per CTA per work item (= unicast columns)
constexpr int kC2Halo = 4;                // shared columns per work item, 2 loaded by each CTA
constexpr int kC2Cols = kC2Ty + kC2Halo;  // columns resident per CTA per stage
constexpr int kC2Stages = 2;
constexpr int kC2StageBytes = kC2Rows * kC2Cols * 8;  // 96 KB
constexpr int kC2SmemBytes = 1024 + kC2Stages * kC2StageBytes;
constexpr int kC2Consumers = 256;

__device__ __forceinline__ uint32_t cluster_ctarank()
{
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all()
{
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t cta_rank)
{
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(cta_rank));
  return r;
}
// "I am done reading this stage": orders nothing but the reads that already fed my arithmetic
__device__ __forceinline__ void mbar_arrive_relaxed_local(uint64_t* bar)
{
  asm volatile("mbarrier.arrive.relaxed.cta.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_relaxed_remote(uint32_t cluster_addr)
{
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity)
{
  uint32_t ok = 0;
  while (!ok) {
    asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  }
}
__device__ __forceinline__ void tma_load_2d_mcast(void* smem_dst, const void* tmap, int32_t c0, int32_t c1,
                                                  uint64_t* bar, uint16_t cta_mask)
{
  asm volatile(
    "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster "
    "[%0], [%1, {%2, %3}], [%4], %5;" ::"r"(smem_u32(smem_dst)),
    "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar)), "h"(cta_mask)
    : "memory");
}
__device__ __forceinline__ double4_t lds256(uint32_t addr)
{
  double4_t v;
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr) : "memory");
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(v.z), "=d"(v.w) : "r"(addr + 16) : "memory");
  return v;
}

// out(i, j) = scale * sum_k c_k z(i, j + k): z is (nx, ny + 4); map8 has box 256 x 8, map2 box 256 x 2
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kC2Consumers + 32, 1)
  stencil2d_d1_cluster_kernel(const __grid_constant__ CUtensorMap map8, const __grid_constant__ CUtensorMap map2,
                              int64_t nx, int64_t ny, double* __restrict__ out, int64_t ldo, double scale)
{
  extern __shared__ __align__(1024) unsigned char smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty = full + kC2Stages;
  unsigned char* data = smem + 1024;
  const int tid = threadIdx.x;
  const uint32_t q = cluster_ctarank();  // 0: low half of the column window, 1: high half
  if (tid == 0) {
    for (int s = 0; s < kC2Stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 2 * (kC2Consumers / 32));  // both CTAs' consumer warps release a stage
    }
    mbar_fence_init();
    tma_prefetch_desc(&map8);
    tma_prefetch_desc(&map2);
  }
  cluster_sync_all();  // barriers initialised before the partner's multicast or arrives can land

  const int64_t tiles_i = (nx + kC2Rows - 1) / kC2Rows;
  const int64_t pairs_j = (ny + 2 * kC2Ty - 1) / (2 * kC2Ty);
  const int64_t nwork = tiles_i * pairs_j;
  const int64_t ncl = gridDim.x / 2, cl = blockIdx.x / 2;
  // stage layout (doubles): 4 row-boxes x [8 unicast columns | 4 shared columns] x 256 rows
  constexpr int kBoxU = kC2Box * kC2Ty;      // doubles of one unicast box
  constexpr int kBoxS = kC2Box * 2;          // doubles of one multicast box (2 columns)
  constexpr int kOffS = (kC2Rows / kC2Box) * kBoxU;  // shared columns start after all unicast boxes

  if (tid >= kC2Consumers) {
    if (tid == kC2Consumers) {
      int k = 0;
      for (int64_t w = cl; w < nwork; w += ncl, ++k) {
        const int s = k % kC2Stages;
        const uint32_t ph = (k / kC2Stages) & 1;
        mbar_wait_cluster(&empty[s], ph ^ 1);  // my consumers AND the partner's are done with stage s
        const int64_t pj = w / tiles_i, ti = w - pj * tiles_i;
        const int32_t c_row = (int32_t)(ti * kC2Rows);
        const int32_t g0 = (int32_t)(pj * 2 * kC2Ty);  // first ghosted column of the 20-column window
        double* st = reinterpret_cast<double*>(data + s * kC2StageBytes);
        mbar_arrive_expect_tx(&full[s], kC2StageBytes);  // 8 own + 2 own-multicast + 2 partner-multicast columns
#pragma unroll
        for (int b = 0; b < kC2Rows / kC2Box; ++b) {
          tma_load_2d(st + b * kBoxU, &map8, c_row + b * kC2Box, g0 + (q ? kC2Ty + kC2Halo : 0), &full[s]);
          // my half of the shared halo goes to BOTH CTAs (same offset, both full barriers)
          tma_load_2d_mcast(st + kOffS + (b * 2 + q) * kBoxS, &map2, c_row + b * kC2Box, g0 + kC2Ty + 2 * q,
                            &full[s], (uint16_t)0x3);
        }
      }
    }
    __syncwarp();  // the producer warp reconverges before the (aligned) cluster barrier below
  } else {
    const int lane = tid & 31;
    const int b = tid / 64, r = 4 * (tid % 64);  // row-box and row within it
    const double w0 = scale * c0(), w1 = scale * c1(), w3 = scale * c3(), w4 = scale * c4();
    int k = 0;
    for (int64_t w = cl; w < nwork; w += ncl, ++k) {
      const int s = k % kC2Stages;
      const uint32_t ph = (k / kC2Stages) & 1;
      const int64_t pj = w / tiles_i, ti = w - pj * tiles_i;
      const int64_t i = ti * kC2Rows + 4 * tid;
      mbar_wait(&full[s], ph);  // all 12 columns I need have landed (mine and the partner's multicast)
      const uint32_t base = smem_u32(data + s * kC2StageBytes);
      const uint32_t uni = base + (uint32_t)(b * kBoxU + r) * 8;               // + c * 256 * 8, c < 8
      const uint32_t sh = base + (uint32_t)(kOffS + b * 2 * kBoxS + r) * 8;    // + c * 256 * 8, c < 4
      double4_t a1 = {0, 0, 0, 0}, a2 = a1, a3 = a1, a4 = a1;
      const int64_t jbase = pj * 2 * kC2Ty + q * kC2Ty;  // first output column of this CTA
      // marching order of the 12 columns: q == 0: unicast 0..7 then shared 0..3
      //                                   q == 1: shared 0..3 then unicast 0..7
#pragma unroll
      for (int g = 0; g < kC2Ty + 4; ++g) {
        double4_t m;
        if (q == 0) {
          m = g < kC2Ty ? lds256(uni + g * kC2Box * 8) : lds256(sh + (g - kC2Ty) * kC2Box * 8);
        } else {
          m = g < kC2Halo ? lds256(sh + g * kC2Box * 8) : lds256(uni + (g - kC2Halo) * kC2Box * 8);
        }
        if (g >= 4) {
          const int64_t j = jbase + g - 4;
          if (i < nx && j < ny) {
            double4_t o;
            o.x = a4.x + w4 * m.x; o.y = a4.y + w4 * m.y; o.z = a4.z + w4 * m.z; o.w = a4.w + w4 * m.w;
            stg256(out + i + j * ldo, o);
          }
        }
        a4.x = a3.x + w3 * m.x; a4.y = a3.y + w3 * m.y; a4.z = a3.z + w3 * m.z; a4.w = a3.w + w3 * m.w;
        a3 = a2;
        a2.x = a1.x + w1 * m.x; a2.y = a1.y + w1 * m.y; a2.z = a1.z + w1 * m.z; a2.w = a1.w + w1 * m.w;
        a1.x = w0 * m.x; a1.y = w0 * m.y; a1.z = w0 * m.z; a1.w = w0 * m.w;
      }
      __syncwarp();
      if (lane == 0) {
        // both producers refill this stage (mine directly, the partner's by multicast)
        mbar_arrive_relaxed_local(&empty[s]);
        mbar_arrive_relaxed_remote(map_to_cta(smem_u32(&empty[s]), q ^ 1));
      }
    }
  }
  cluster_sync_all();  // nobody leaves while the partner's multicast may still target its shared memory
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int encode_tiled_2d(CUtensorMap* map, const double* base, int64_t d0, int64_t d1, int64_t ld, int box0,
                           int box1)
{
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    GMT_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
    GMT_REQUIRE(p && qres == cudaDriverEntryPointSuccess, GMT_EUNSUPPORTED, "cuTensorMapEncodeTiled not available");
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  const cuuint64_t dims[2] = {(cuuint64_t)d0, (cuuint64_t)d1};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * 8};
  const cuuint32_t box[2] = {(cuuint32_t)box0, (cuuint32_t)box1};
  const cuuint32_t estr[2] = {1, 1};
  CUresult rc = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  GMT_REQUIRE(rc == CUDA_SUCCESS, GMT_EUNSUPPORTED, "cuTensorMapEncodeTiled failed with %d", (int)rc);
  return 0;
}

}  // namespace gmt

extern "C" {

int gmt_stencil1d_tma_f64(int64_t n, const double* in, double* out, double scale, gmt_stream_t stream)
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
  unsigned int* counters = nullptr;
  int rc = stream_counters(as_stream(stream), &counters);
  if (rc) return rc;
  int occ = 0;
  GMT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, stencil1d_tma_kernel, kTmaConsumers + 32,
                                                         kTmaSmemBytes));
  if (occ < 1) occ = 1;
  const int64_t ntiles = (n + kTmaTile - 1) / kTmaTile;
  const int grid = (int)std::min<int64_t>(ntiles, (int64_t)sm_count() * occ);
  stencil1d_tma_kernel<<<grid, kTmaConsumers + 32, kTmaSmemBytes, as_stream(stream)>>>(n, in, out, scale, counters);
  GMT_LAUNCH_CHECK();
  return 0;
}

int gmt_fused_stencil1d_tma_f64(const gmt_halo1d_rank* ranks_dev, int n_ranks, int64_t n, gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(ranks_dev && n_ranks >= 1 && n > 0, GMT_EINVAL, "gmt_fused_stencil1d_tma_f64: bad argument");
  GMT_REQUIRE(n % 2 == 0, GMT_EUNSUPPORTED, "gmt_fused_stencil1d_tma_f64: n must be even (16-byte halo stores, bulk copies)");
  static thread_local int configured_dev = -1;
  int dev = 0;
  GMT_CUDA(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    GMT_CUDA(cudaFuncSetAttribute(fused1d_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTmaSmemBytes));
    configured_dev = dev;
  }
  int occ = 0;
  GMT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fused1d_tma_kernel, kTmaConsumers + 32, kTmaSmemBytes));
  if (occ < 1) occ = 1;
  int64_t cap = (int64_t)sm_count() * occ;
  if (n_ranks > 1) cap /= n_ranks;
  GMT_REQUIRE(cap >= 1, GMT_ETOOBIG, "too many virtual ranks (%d) for one cooperative launch", n_ranks);
  const int64_t ntiles = (n + kTmaTile - 1) / kTmaTile;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(ntiles, cap));
  void* args[] = {(void*)&ranks_dev};
  if (n_ranks > 1)
    GMT_CUDA(cudaLaunchCooperativeKernel((const void*)fused1d_tma_kernel, dim3(grid, n_ranks), dim3(kTmaConsumers + 32),
                                         args, kTmaSmemBytes, as_stream(stream)));
  else
    GMT_CUDA(cudaLaunchKernel((const void*)fused1d_tma_kernel, dim3(grid, 1), dim3(kTmaConsumers + 32), args,
                              kTmaSmemBytes, as_stream(stream)));
  return 0;
}

/* dim-1 derivative through 2-D TMA tensor maps and thread-block clusters (DSMEM halo sharing).
 * z is (nx, ny + 4), out is (nx, ny); needs nx % 4 == 0, 16-byte aligned z with ldz % 2 == 0 and
 * 32-byte aligned out with ldo % 4 == 0.                                                    */
int gmt_stencil2d_d1_cluster_f64(int64_t nx, int64_t ny, const double* z, int64_t ldz, double* out,
                                 int64_t ldo, double scale, gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(nx > 0 && ny > 0 && z && out && ldz >= nx && ldo >= nx, GMT_EINVAL,
              "gmt_stencil2d_d1_cluster_f64: bad argument");
  GMT_REQUIRE(nx % 4 == 0 && ldo % 4 == 0 && ldz % 2 == 0 && aligned(z, 16) && aligned(out, 32), GMT_EUNSUPPORTED,
              "gmt_stencil2d_d1_cluster_f64: unsupported shape or alignment");
  {
    // TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from
    // another GPU's memory hung the GPU: the source must be a plain device allocation of the current device
    // (not a peer mapping, an IPC mapping, a symmetric-memory view of another rank, or managed memory).
    int dev = -1;
    GMT_CUDA(cudaGetDevice(&dev));
    cudaPointerAttributes attr;
    GMT_CUDA(cudaPointerGetAttributes(&attr, z));
    GMT_REQUIRE(attr.type == cudaMemoryTypeDevice && attr.device == dev, GMT_EUNSUPPORTED,
                "gmt_stencil2d_d1_cluster_f64: z must be device memory of the current GPU (device %d); "
                "use the LSU variant for anything else", dev);
  }
  CUtensorMap map8, map2;
  int rc = encode_tiled_2d(&map8, z, nx, ny + 4, ldz, kC2Box, kC2Ty);
  if (rc) return rc;
  rc = encode_tiled_2d(&map2, z, nx, ny + 4, ldz, kC2Box, 2);
  if (rc) return rc;
  static thread_local int configured_dev = -1;
  int dev = 0;
  GMT_CUDA(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    GMT_CUDA(cudaFuncSetAttribute(stencil2d_d1_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  kC2SmemBytes));
    configured_dev = dev;
  }
  const int64_t tiles_i = (nx + kC2Rows - 1) / kC2Rows;
  const int64_t nwork = tiles_i * ((ny + 2 * kC2Ty - 1) / (2 * kC2Ty));
  const int64_t ncl = std::max<int64_t>(1, std::min<int64_t>(nwork, sm_count() / 2));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(2 * ncl));
  cfg.blockDim = dim3(kC2Consumers + 32);
  cfg.dynamicSmemBytes = kC2SmemBytes;
  cfg.stream = as_stream(stream);
  GMT_CUDA(cudaLaunchKernelEx(&cfg, stencil2d_d1_cluster_kernel, map8, map2, nx, ny, out, ldo, scale));
  return 0;
}

}  // extern "C"
