// collectives.cu -- allgather / gather / all-reduce / barrier and the two fused
// compute+collective ops (daxpy + global sum, axis-sum + all-reduce) as kernels over
// NVLink 5 / NVSwitch peer memory.
//
// Replaces, from the reference:
//   MPI_Allgather(MPI_IN_PLACE) / MPI_Allgather   mpi_daxpy_nvtx.cc:285,288; mpigatherinplace.f90:39
//   MPI_Allreduce(MPI_IN_PLACE, SUM)              mpi_stencil2d_gt.cc:615,624
//   cublasDaxpy + allgather + host sums           mpi_daxpy_nvtx.cc:245-310  (daxpy_allreduce)
//   gt::sum_axis_to + MPI_Allreduce               mpi_stencil2d_gt.cc:611-625 (axis_sum_allreduce)
//   MPI_Barrier                                   mpi_daxpy_nvtx.cc:277
// and adds the gather-to-root BASELINE config 5 asks for.
//
// NVSwitch gives every GPU the same bandwidth to every peer, so nothing here uses rings or
// trees: data goes straight to its final place (one store per peer, or one multimem.st for
// all of them), small reductions are one-shot (every rank reads every peer's vector), larger
// ones two-shot.  Every wait is a bounded spin on a system-scope acquire.
#include <algorithm>

#include "common.cuh"
#include "p2p.cuh"

namespace gmt {

constexpr int kCollThreads = 512;
constexpr int kRoundStride = 32;  // words per signalling round in a flag block
constexpr int kLLBase = 64;       // LL slots: [parity][src rank][2 words] after the two rounds
constexpr int64_t kLLMaxCells = 1LL << 18;  // n x world up to here: vectors travel as LL cells by default
//   (same rule as parallel/collectives.py: each double costs 16 bytes per peer, so at 8 ranks the
//   switch to one-/two-shot comes at 32 Ki doubles, at 2 ranks at 128 Ki)
constexpr int64_t kLLGatherMax = 65536;     // allgather slices up to 512 KiB can travel as LL cells (given stage_cap)
constexpr uint32_t kStatusTooBig = 0x1000u; // status word: the call needed more LL cells than stage_cap

struct CollShared {
  gmt_coll_rank R;
  double* buf[GMT_COLL_MAX_WORLD];
  double* stage[GMT_COLL_MAX_WORLD];
  uint64_t* flag[GMT_COLL_MAX_WORLD];
  double red[40];
  int64_t next[2];   // [1]: the epoch read by coll_enter
  bool last;
};

// Prologue of every collective: the descriptor into shared memory, the three peer tables and the
// epoch.  The table entries and the epoch are fetched straight from global memory in parallel with
// the descriptor copy (two dependent loads deep, one barrier) instead of descriptor -> tables ->
// epoch one after the other: every collective here is latency-bound and each level is ~0.5 us.
__device__ __forceinline__ uint64_t coll_enter(CollShared& S, const gmt_coll_rank* ranks)
{
  const gmt_coll_rank* Rg = ranks + blockIdx.y;
  {
    const uint64_t* s = reinterpret_cast<const uint64_t*>(Rg);
    uint64_t* d = reinterpret_cast<uint64_t*>(&S.R);
    for (int w = threadIdx.x; w < (int)(sizeof(gmt_coll_rank) / 8); w += blockDim.x) d[w] = s[w];
  }
  if (threadIdx.x < GMT_COLL_MAX_WORLD) {
    const int t = threadIdx.x, W = Rg->world;
    double* const* pb = Rg->peer_buf;
    double* const* ps = Rg->peer_stage;
    uint64_t* const* pf = Rg->peer_flag;
    if (t < W) {
      S.buf[t] = pb ? pb[t] : nullptr;
      S.stage[t] = ps ? ps[t] : nullptr;
      S.flag[t] = pf[t];
    }
  } else if (threadIdx.x == GMT_COLL_MAX_WORLD) {
    S.next[1] = (int64_t)ld_relaxed_sys(Rg->epoch);
  }
  __syncthreads();
  return (uint64_t)S.next[1] + 1;
}

// "rank R.rank has reached `round` of epoch e" to every rank (or to one).  Call after a CTA
// barrier: the release is cumulative over what the CTA wrote before the barrier.
// publishes = true: the signal makes data this grid wrote visible -- a release.  In a multi-CTA grid
// arrive_last has just executed its system fences on thread 0 of the last CTA, so that thread stores all the
// flags itself, relaxed (fence + relaxed store is a release pattern; a second MEMBAR.SYS costs ~1.4 us).  A
// single-CTA grid has no such fence: its threads store one flag each with release semantics (the CTA barrier
// before makes the release cumulative over what the CTA wrote).  publishes = false: the signal only says
// "I am in this call" (entry round) -- nothing of this kernel's to publish: relaxed, one flag per thread.
__device__ __forceinline__ void signal_all(CollShared& S, int round, uint64_t e, bool publishes = true)
{
  const int W = S.R.world;
  if (publishes && gridDim.x > 1) {
    if (threadIdx.x == 0)
      for (int p = 0; p < W; ++p) st_relaxed_sys(S.flag[p] + round * kRoundStride + S.R.rank, e);
    return;
  }
  if ((int)threadIdx.x < W) {
    uint64_t* f = S.flag[threadIdx.x] + round * kRoundStride + S.R.rank;
    if (publishes) st_release_sys(f, e);
    else st_relaxed_sys(f, e);
  }
}
__device__ __forceinline__ void signal_one(CollShared& S, int dst, int round, uint64_t e)
{
  if (threadIdx.x == 0) st_release_sys(S.flag[dst] + round * kRoundStride + S.R.rank, e);
}
// wait until every rank (or one rank) has signalled `round` of epoch e; ends with a CTA barrier
__device__ __forceinline__ void wait_all(CollShared& S, int round, uint64_t e)
{
  if ((int)threadIdx.x < S.R.world) {
    if (!wait_flag_ge(S.flag[S.R.rank] + round * kRoundStride + threadIdx.x, e, S.R.spin_limit))
      atomicExch(S.R.status, 1u + threadIdx.x);
  }
  __syncthreads();
}
__device__ __forceinline__ void wait_one(CollShared& S, int src, int round, uint64_t e)
{
  if (threadIdx.x == 0) {
    if (!wait_flag_ge(S.flag[S.R.rank] + round * kRoundStride + src, e, S.R.spin_limit))
      atomicExch(S.R.status, 1u + src);
  }
  __syncthreads();
}

// True in exactly one CTA of this rank: the last one to get here.  With publish = true,
// everything the other CTAs wrote before is visible (system scope) to whoever synchronises
// with that CTA afterwards (fence / relaxed atomic / fence); publish = false only counts.
// A single-CTA grid needs neither the atomic nor the fences: the CTA barrier orders its own
// writes before the release stores that follow.
__device__ __forceinline__ bool arrive_last(CollShared& S, uint32_t* counter, bool publish = true)
{
  __syncthreads();
  if (gridDim.x == 1) return true;
  if (threadIdx.x == 0) {
    if (publish) fence_sys();
    S.last = (atomicAdd(counter, 1u) == gridDim.x - 1);
    if (S.last && publish) fence_sys();
  }
  __syncthreads();
  return S.last;
}

__device__ __forceinline__ void coll_finish(CollShared& S, uint64_t e)
{
  if (threadIdx.x == 0) {
    S.R.done[0] = S.R.done[1] = S.R.done[2] = S.R.done[3] = 0u;
    __threadfence();
    st_relaxed_sys(S.R.epoch, e);
  }
}

// weak (unqualified) multimem.st: ordering comes from the fence + flag that follow anyway
// A vector longer than the staging area holds (the raw entry points cannot see stage_cap from the host):
// write nothing, say so in the status word, still advance the epoch so that the rank stays in step.
// CTA-uniform, and every rank takes the same branch (same n, same stage_cap).
__device__ __forceinline__ bool coll_too_big(CollShared& S, uint64_t e, int64_t n)
{
  if (n <= S.R.stage_cap) return false;
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(S.R.status, kStatusTooBig);
  if (arrive_last(S, S.R.done + 2, false)) coll_finish(S, e);
  return true;
}

__device__ __forceinline__ void multimem_st_32B_weak(double* mc, const double4_t& v)
{
  const float4* f = reinterpret_cast<const float4*>(&v);
  asm volatile("multimem.st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(f[0].x), "f"(f[0].y), "f"(f[0].z),
               "f"(f[0].w)
               : "memory");
  asm volatile("multimem.st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc + 2), "f"(f[1].x), "f"(f[1].y), "f"(f[1].z),
               "f"(f[1].w)
               : "memory");
}

__device__ __forceinline__ void multimem_st_32B(double* mc, const double4_t& v)
{
  // the bytes are what matters: 2 x multimem.st.v4.f32 carry 4 doubles to every GPU of the group
  const float4* f = reinterpret_cast<const float4*>(&v);
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(f[0].x), "f"(f[0].y),
               "f"(f[0].z), "f"(f[0].w)
               : "memory");
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc + 2), "f"(f[1].x), "f"(f[1].y),
               "f"(f[1].z), "f"(f[1].w)
               : "memory");
}

// ============================================================ allgather / gather
// root < 0: allgather (everyone receives).  root >= 0: gather to root.
__global__ void __launch_bounds__(kCollThreads, 2)
  gather_kernel(const gmt_coll_rank* __restrict__ ranks, int64_t n, int from_in, int use_mc, int root)
{
  __shared__ CollShared S;
  const uint64_t e = coll_enter(S, ranks);
  const int W = S.R.world, r = S.R.rank;
  const bool all = root < 0;
  // round 0: nobody's buffer may be written before its owner has entered this epoch
  if (blockIdx.x == 0) {
    if (all) signal_all(S, 0, e, false);
    else if (r == root) signal_all(S, 0, e, false);
  }
  if (all) wait_all(S, 0, e);
  else if (r != root) wait_one(S, root, 0, e);

  const bool sender = all || r != root;
  if (sender) {
    double* own = S.buf[r];
    const double* src = from_in ? S.R.in : own + (int64_t)r * n;
    const int64_t off = (int64_t)r * n;
    const bool vec = (n % 4 == 0) && aligned(src, 32) && aligned(own, 32);
    const int64_t stride = (int64_t)gridDim.x * kCollThreads;
    const int64_t t0 = (int64_t)blockIdx.x * kCollThreads + threadIdx.x;
    if (vec) {
      const int64_t nv = n / 4;
      for (int64_t v = t0; v < nv; v += 2 * stride) {
        const bool two = v + stride < nv;
        const double4_t a = ldg256(src + 4 * v);
        double4_t b;
        if (two) b = ldg256(src + 4 * (v + stride));
        if (all && use_mc == 1) {
          multimem_st_32B(S.R.mc_buf + off + 4 * v, a);
          if (two) multimem_st_32B(S.R.mc_buf + off + 4 * (v + stride), b);
        } else if (all && use_mc == 3) {
          multimem_st_32B_weak(S.R.mc_buf + off + 4 * v, a);
          if (two) multimem_st_32B_weak(S.R.mc_buf + off + 4 * (v + stride), b);
        } else if (all && use_mc == 4) {
          // an ordinary store on the multicast mapping is replicated by the switch as well
          stg256(S.R.mc_buf + off + 4 * v, a);
          if (two) stg256(S.R.mc_buf + off + 4 * (v + stride), b);
        } else if (all) {
          // peers visited starting at r+1: at any moment the ranks target a permutation of
          // destinations instead of all hammering the same GPU's ingress port
          for (int k = from_in ? 0 : 1; k < W; ++k) {
            const int p = (r + k) % W;
            stg256(S.buf[p] + off + 4 * v, a);
            if (two) stg256(S.buf[p] + off + 4 * (v + stride), b);
          }
        } else {
          stg256(S.buf[root] + off + 4 * v, a);
          if (two) stg256(S.buf[root] + off + 4 * (v + stride), b);
        }
      }
    } else {
      for (int64_t i = t0; i < n; i += stride) {
        const double a = src[i];
        if (all) {
          for (int p = 0; p < W; ++p)
            if (p != r || from_in) S.buf[p][off + i] = a;
        } else {
          S.buf[root][off + i] = a;
        }
      }
    }
  } else if (from_in) {
    // root of a gather with a separate send buffer keeps its own contribution too
    double* own = S.buf[r] + (int64_t)r * n;
    for (int64_t i = (int64_t)blockIdx.x * kCollThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kCollThreads)
      own[i] = S.R.in[i];
  }
  // round 1: my slice has landed everywhere; return only when everyone's slice has landed here
  // (a gather to root ends with the same all-to-all round although only the root needs the data:
  // every collective here must leave nobody more than one call behind, because the staging areas and
  // LL cells of the reductions are double-buffered by epoch parity -- a sender that could run through
  // gathers without waiting for anybody would reach the same parity again while a slow rank is still
  // reading what the sender put there two calls ago)
  if (arrive_last(S, S.R.done + 0)) {
    signal_all(S, 1, e);
    wait_all(S, 1, e);
    coll_finish(S, e);
  }
}

// LL cells live behind the four staging vectors: word index 4*cap + ((parity*W + src)*cap + i)*2 of a
// rank's staging area (see the LL all-reduce below).
__device__ __forceinline__ uint64_t* ll_cell(const CollShared& S, int owner, int par, int src, int64_t i)
{
  return reinterpret_cast<uint64_t*>(S.stage[owner]) + 4 * S.R.stage_cap +
         (((int64_t)par * S.R.world + src) * S.R.stage_cap + i) * 2;
}

// ---- small allgather through LL cells (slices up to 512 KiB when the staging area has the cells).  The flag-based kernel above pays two
// signalling rounds (entry: nobody's user buffer may be written before its owner is in the call;
// exit: my slice has landed everywhere) = two NVLink round trips plus two fences.  Here nobody
// writes anybody's USER buffer: every value goes into the peers' LL cell area (staging memory that
// only collectives touch) as a 16-byte cell carrying the epoch in both words, and every rank copies
// what arrives into its own buffer -- one one-way trip, no fence, no entry round.  The thread that
// waits for a cell waits for another GPU only, so no co-residency is assumed.
__global__ void __launch_bounds__(kCollThreads, 2)
  gather_ll_kernel(const gmt_coll_rank* __restrict__ ranks, int64_t n, int from_in)
{
  __shared__ CollShared S;
  const uint64_t e = coll_enter(S, ranks);
  const int W = S.R.world, r = S.R.rank, par = (int)(e & 1);
  if (n > S.R.stage_cap) {   // CTA-uniform; every rank takes the same branch
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(S.R.status, kStatusTooBig);
    if (arrive_last(S, S.R.done + 2, false)) coll_finish(S, e);
    return;
  }
  double* own = S.buf[r];
  const double* src = from_in ? S.R.in : own + (int64_t)r * n;
  const int64_t stride = (int64_t)gridDim.x * kCollThreads;
  const int64_t t0 = (int64_t)blockIdx.x * kCollThreads + threadIdx.x;
  const uint64_t e32 = e & 0xffffffffull;
  for (int64_t i = t0; i < n; i += stride) {
    const double v = src[i];
    for (int k = 1; k < W; ++k) st_ll_cell(ll_cell(S, (r + k) % W, par, r, i), v, e32);
    if (from_in) own[(int64_t)r * n + i] = v;
  }
  const int64_t limit = S.R.spin_limit > 0 ? S.R.spin_limit : kDefaultSpinLimit;
  const int64_t total = (int64_t)(W - 1) * n;
  // four cells per thread in flight; whatever has not arrived yet is then polled one by one
  for (int64_t q0 = t0; q0 < total; q0 += 4 * stride) {
    uint64_t w0[4], w1[4];
    int64_t dst[4];
    const uint64_t* cell[4];
    int src_rank[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t q = q0 + (int64_t)k * stride;
      dst[k] = -1;
      if (q < total) {
        const int kk = 1 + (int)(q / n);
        const int64_t i = q - (int64_t)(kk - 1) * n;
        const int p = (r + kk) % W;
        src_rank[k] = p;
        cell[k] = ll_cell(S, r, par, p, i);
        dst[k] = (int64_t)p * n + i;
        ld_ll_raw(cell[k], w0[k], w1[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (dst[k] < 0) continue;
      double v = 0.0;
      bool ok = ll_unpack(w0[k], w1[k], e32, v);
      for (int64_t it = 0; !ok && it < limit; ++it) {
        ok = ld_ll_cell(cell[k], e32, v);
        if (it > 64) __nanosleep(32);
      }
      if (!ok) atomicExch(S.R.status, 1u + src_rank[k]);
      own[dst[k]] = v;
    }
  }
  if (arrive_last(S, S.R.done + 2, false)) coll_finish(S, e);
}

// ============================================================ all-reduce building blocks
// After this call every rank's input vector of epoch e sits in its staging area and is visible.
__device__ __forceinline__ void publish_inputs(CollShared& S, uint64_t e)
{
  if (arrive_last(S, S.R.done + 0)) signal_all(S, 0, e);
  wait_all(S, 0, e);
}

// out[i] = sum over ranks (rank order) of their staged inputs; i in this CTA's share
template <int kAlgo>
__device__ __forceinline__ void reduce_tail(CollShared& S, uint64_t e, int64_t n)
{
  const int W = S.R.world, r = S.R.rank, par = (int)(e & 1);
  const int64_t cap = S.R.stage_cap;
  const int64_t in_off = (int64_t)par * cap, res_off = (int64_t)(2 + par) * cap;
  const int64_t stride = (int64_t)gridDim.x * kCollThreads;
  const int64_t t0 = (int64_t)blockIdx.x * kCollThreads + threadIdx.x;
  double* out = S.R.out;
  if (kAlgo == 0) {
    for (int64_t i = t0; i < n; i += stride) {
      double s = 0.0;
      for (int p = 0; p < W; ++p) s += ld_peer_f64(S.stage[p] + in_off + i);
      out[i] = s;
    }
  } else if (kAlgo == 2) {
    const double* mc = S.R.mc_stage + in_off;
    for (int64_t i = t0; i < n; i += stride) out[i] = multimem_ld_reduce_add_f64(mc + i);
  } else {
    // two-shot: reduce my slice, store it into everyone's result area, then collect
    const int64_t lo = n * r / W, hi = n * (r + 1) / W;
    for (int64_t i = lo + t0; i < hi; i += stride) {
      double s = 0.0;
      for (int p = 0; p < W; ++p) s += ld_peer_f64(S.stage[p] + in_off + i);
      for (int k = 0; k < W; ++k) S.stage[(r + k) % W][res_off + i] = s;
    }
    if (arrive_last(S, S.R.done + 1)) signal_all(S, 1, e);
    wait_all(S, 1, e);
    const double* res = S.stage[r] + res_off;
    for (int64_t i = t0; i < n; i += stride) out[i] = ld_peer_f64(res + i);
  }
  if (arrive_last(S, S.R.done + 2, false)) coll_finish(S, e);
}

// ---- "LL" all-reduce: no grid-wide rendezvous, no flag round trip.  Every rank stores each of its
// values as an LL cell into the slot [parity][its rank][i] of EVERY rank (its own included), then
// sums, in rank order, the W cells of each i as they turn up.  The cells live behind the four
// staging vectors: word index 4*cap + ((parity*W + src)*cap + i)*2 of a rank's staging area.
// Parity double-buffering is enough for the same reason as for the halo inboxes: a rank can only be
// two calls ahead of a peer after that peer has read everything of the call in between.
__device__ __forceinline__ void ll_push(const CollShared& S, uint64_t e, int64_t i, double v)
{
  const int W = S.R.world, r = S.R.rank, par = (int)(e & 1);
  for (int k = 0; k < W; ++k) st_ll_cell(ll_cell(S, (r + k) % W, par, r, i), v, e & 0xffffffffull);
}
__device__ __forceinline__ double ll_collect(CollShared& S, uint64_t e, int64_t i)
{
  const int W = S.R.world, r = S.R.rank, par = (int)(e & 1);
  const int64_t limit = S.R.spin_limit > 0 ? S.R.spin_limit : kDefaultSpinLimit;
  const uint64_t e32 = e & 0xffffffffull;
  double s = 0.0;
  // the cells of up to eight source ranks are fetched together (one memory latency instead of eight);
  // whatever has not arrived yet is then polled one by one.  Adds stay in rank order: identical bits everywhere.
  for (int base = 0; base < W; base += 8) {
    uint64_t w0[8], w1[8];
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (base + k < W) ld_ll_raw(ll_cell(S, r, par, base + k, i), w0[k], w1[k]);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (base + k >= W) break;
      double v = 0.0;
      bool ok = ll_unpack(w0[k], w1[k], e32, v);
      if (!ok) {
        const uint64_t* cell = ll_cell(S, r, par, base + k, i);
        for (int64_t it = 0; it < limit; ++it) {
          if (ld_ll_cell(cell, e32, v)) {
            ok = true;
            break;
          }
          if (it > 64) __nanosleep(32);
        }
        if (!ok) atomicExch(S.R.status, 1u + base + k);
      }
      s += v;
    }
  }
  return s;
}
// every thread collects the indices of its grid-stride share; then the epoch advances
__device__ __forceinline__ void ll_tail(CollShared& S, uint64_t e, int64_t n)
{
  double* out = S.R.out;
  const int64_t stride = (int64_t)gridDim.x * kCollThreads;
  for (int64_t i = (int64_t)blockIdx.x * kCollThreads + threadIdx.x; i < n; i += stride) out[i] = ll_collect(S, e, i);
  if (arrive_last(S, S.R.done + 2, false)) coll_finish(S, e);
}

__global__ void __launch_bounds__(kCollThreads, 2)
  allreduce_ll_kernel(const gmt_coll_rank* __restrict__ ranks, int64_t n)
{
  __shared__ CollShared S;
  const uint64_t e = coll_enter(S, ranks);
  if (coll_too_big(S, e, n)) return;
  const double* in = S.R.in;
  const int64_t stride = (int64_t)gridDim.x * kCollThreads;
  for (int64_t i = (int64_t)blockIdx.x * kCollThreads + threadIdx.x; i < n; i += stride) ll_push(S, e, i, in[i]);
  ll_tail(S, e, n);
}

template <int kAlgo>
__global__ void __launch_bounds__(kCollThreads, 2)
  allreduce_kernel(const gmt_coll_rank* __restrict__ ranks, int64_t n)
{
  __shared__ CollShared S;
  const uint64_t e = coll_enter(S, ranks);
  if (coll_too_big(S, e, n)) return;
  double* mine = S.stage[S.R.rank] + (int64_t)(e & 1) * S.R.stage_cap;
  const double* in = S.R.in;
  for (int64_t i = (int64_t)blockIdx.x * kCollThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kCollThreads)
    mine[i] = in[i];
  publish_inputs(S, e);
  reduce_tail<kAlgo>(S, e, n);
}

// lanes 0..W-1 of a warp each fetch one rank's cell of index i; every lane returns the rank-ordered sum
__device__ __forceinline__ double ll_collect_warp(CollShared& S, uint64_t e, int64_t i)
{
  const int W = S.R.world, r = S.R.rank, par = (int)(e & 1), lane = threadIdx.x & 31;
  const int64_t limit = S.R.spin_limit > 0 ? S.R.spin_limit : kDefaultSpinLimit;
  double v = 0.0;
  if (lane < W) {
    const uint64_t* cell = ll_cell(S, r, par, lane, i);
    bool ok = false;
    for (int64_t it = 0; it < limit; ++it) {
      if (ld_ll_cell(cell, e & 0xffffffffull, v)) {
        ok = true;
        break;
      }
      if (it > 64) __nanosleep(32);
    }
    if (!ok) atomicExch(S.R.status, 1u + lane);
  }
  double s = 0.0;
  for (int src = 0; src < W; ++src) s += __shfl_sync(0xffffffffu, v, src);  // rank order: same bits everywhere
  return s;
}

// axis sum straight into the staging area (or out as LL cells), then the all-reduce.  The LL /
// one-shot choice is made in the kernel from the descriptor's world size, with the same rule as the
// Python layer (n_out x world <= 2^18 cells and n_out <= stage_cap).
// In the LL flavour the thread (warp) that pushed an index is the one that later collects it, so a
// wait is only ever for ANOTHER rank's push, never for a sibling CTA of this grid: the kernel is safe
// without a cooperative launch even when only part of the grid is resident (ranks sharing a GPU
// across processes -- the MPS case the reference was written for, daxpy.cu:6-7 -- or another stream
// holding SMs).
__global__ void __launch_bounds__(kCollThreads, 2)
  axis_sum_allreduce_kernel(const gmt_coll_rank* __restrict__ ranks, int64_t nx, int64_t ny, int64_t ld,
                            int axis, int64_t n_out)
{
  __shared__ CollShared S;
  __shared__ double tile[16][33];
  const uint64_t e = coll_enter(S, ranks);
  if (coll_too_big(S, e, n_out)) return;
  const bool ll = n_out * S.R.world <= kLLMaxCells;
  double* mine = S.stage[S.R.rank] + (int64_t)(e & 1) * S.R.stage_cap;
  const double* z = S.R.in;
  double* out = S.R.out;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (axis == 0) {
    // out[j] = sum_i z(i,j): one warp per column, lanes along the contiguous axis
    const int64_t nwarps = (int64_t)gridDim.x * (kCollThreads / 32);
    const int64_t jw = (int64_t)blockIdx.x * (kCollThreads / 32) + warp;
    for (int64_t j = jw; j < n_out; j += nwarps) {
      const double* col = z + j * ld;
      double a0 = 0.0, a1 = 0.0;
      int64_t i = lane;
      for (; i + 32 < nx; i += 64) {
        a0 += col[i];
        a1 += col[i + 32];
      }
      if (i < nx) a0 += col[i];
      const double s = warp_sum(a0 + a1);
      if (lane == 0) {
        if (ll) ll_push(S, e, j, s);
        else mine[j] = s;
      }
    }
    if (ll) {
      for (int64_t j = jw; j < n_out; j += nwarps) {   // the columns this warp pushed
        const double s = ll_collect_warp(S, e, j);
        if (lane == 0) out[j] = s;
      }
    }
  } else {
    // out[i] = sum_j z(i,j): 32 rows per CTA trip, the 16 warps split the columns
    for (int64_t i0 = (int64_t)blockIdx.x * 32; i0 < n_out; i0 += (int64_t)gridDim.x * 32) {
      const int64_t i = i0 + lane;
      double acc = 0.0;
      if (i < n_out)
        for (int64_t j = warp; j < ny; j += kCollThreads / 32) acc += z[i + j * ld];
      tile[warp][lane] = acc;
      __syncthreads();
      if (warp == 0 && i < n_out) {
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < kCollThreads / 32; ++k) s += tile[k][lane];
        if (ll) ll_push(S, e, i, s);
        else mine[i] = s;
      }
      __syncthreads();
    }
    if (ll && warp == 0) {
      for (int64_t i0 = (int64_t)blockIdx.x * 32; i0 < n_out; i0 += (int64_t)gridDim.x * 32) {
        const int64_t i = i0 + lane;
        if (i < n_out) out[i] = ll_collect(S, e, i);   // the rows this thread pushed
      }
    }
  }
  if (ll) {
    if (arrive_last(S, S.R.done + 2, false)) coll_finish(S, e);
    return;
  }
  publish_inputs(S, e);
  reduce_tail<0>(S, e, n_out);
}

// ============================================================ daxpy + global sum
// One-shot grid in address order (the pattern that gives the plain daxpy kernel 7 TB/s): CTA b
// streams its contiguous share, reduces it, and stores the partial as an LL cell [parity][rank][b] of
// its own rank's cell area -- no ticket, no fence, no atomic.  The CTA with the highest index sums the
// rank's cells as they turn up, in a fixed order, and the W rank totals then cross NVLink as one LL
// cell each (one one-way trip) and are added in rank order: every rank gets the same bits and the
// result is reproducible run to run.  The collecting CTA waits only for cells, i.e. for CTAs that
// themselves wait for nothing, so no co-residency is assumed.
// (A first version sent every CTA's cell to every rank and summed world x grid cells at the end:
// 38 us at n = 6 Mi on 8 GPUs -- 12288 cells polled by one CTA -- against 33-36 us before.)
// (Round 1: persistent grid drawing 4096-element chunks from a counter with a barrier per chunk,
// then ticket + partials re-read + one LL scalar per rank: 9.2 us at n = 2^20 on ONE GPU where the
// stream itself takes 3.8 us.)
constexpr int kDaUnroll = 2;
constexpr int64_t kDaChunk = (int64_t)kCollThreads * kDaUnroll;   // 256-bit vectors per chunk
constexpr int kDaMaxCtas = 2048;                                 // partial cells per rank (<= stage_cap)

__global__ void __launch_bounds__(kCollThreads, 2)
  daxpy_allreduce_kernel(const gmt_coll_rank* __restrict__ ranks, int64_t n, double a, int chunks_per_cta)
{
  __shared__ double red[40];
  __shared__ double* stage_of[GMT_COLL_MAX_WORLD];
  const gmt_coll_rank* Rg = ranks + blockIdx.y;
  // operand pointers first (one broadcast load each), so the stream starts after ONE dependent load;
  // everything the epilogue needs is fetched in the shadow of the stream
  const double* __restrict__ x = Rg->in;
  double* __restrict__ y = Rg->y;
  const int tid = threadIdx.x;
  const int64_t n4 = n / 4;
  const bool vec = aligned(x, 32) && aligned(y, 32);
  double acc = 0.0;
  const int64_t c0 = (int64_t)blockIdx.x * chunks_per_cta;
  if (vec) {
    for (int q = 0; q < chunks_per_cta; ++q) {
      const int64_t v0 = (c0 + q) * kDaChunk;
      if (v0 >= n4) break;
      double4_t xv[kDaUnroll], yv[kDaUnroll];
#pragma unroll
      for (int u = 0; u < kDaUnroll; ++u) {
        const int64_t i = v0 + tid + (int64_t)u * kCollThreads;
        if (i < n4) {
          xv[u] = ldg256_stream(x + 4 * i);
          yv[u] = ldg256(y + 4 * i);
        }
      }
#pragma unroll
      for (int u = 0; u < kDaUnroll; ++u) {
        const int64_t i = v0 + tid + (int64_t)u * kCollThreads;
        if (i < n4) {
          yv[u].x = a * xv[u].x + yv[u].x;
          yv[u].y = a * xv[u].y + yv[u].y;
          yv[u].z = a * xv[u].z + yv[u].z;
          yv[u].w = a * xv[u].w + yv[u].w;
          stg256(y + 4 * i, yv[u]);
          acc += (yv[u].x + yv[u].y) + (yv[u].z + yv[u].w);
        }
      }
    }
    if (blockIdx.x == gridDim.x - 1) {
      for (int64_t k = 4 * n4 + tid; k < n; k += kCollThreads) {
        const double v = a * x[k] + y[k];
        y[k] = v;
        acc += v;
      }
    }
  } else {
    const int64_t lo = c0 * kDaChunk * 4, hi = min(n, (c0 + chunks_per_cta) * kDaChunk * 4);
    const int64_t hi2 = blockIdx.x == gridDim.x - 1 ? n : hi;
    for (int64_t i = lo + tid; i < hi2; i += kCollThreads) {
      const double v = a * x[i] + y[i];
      y[i] = v;
      acc += v;
    }
  }
  // descriptor fields of the epilogue
  const int W = Rg->world, r = Rg->rank;
  const int64_t cap = Rg->stage_cap;
  const uint64_t e = ld_relaxed_sys(Rg->epoch) + 1;
  const uint64_t e32 = e & 0xffffffffull;
  const int par = (int)(e & 1);
  if (tid < W) stage_of[tid] = Rg->peer_stage[tid];
  const double part = block_sum(acc, red);   // (barriers inside: stage_of is visible afterwards)
  if (tid == 0) red[32] = part;
  __syncthreads();
  const int G = (int)gridDim.x;
  if ((int64_t)G > cap) {   // more partial cells than the staging area holds: report, do not scribble
    if (blockIdx.x == G - 1 && tid == 0) {
      atomicOr(Rg->status, kStatusTooBig);
      *Rg->out = __longlong_as_double(0x7ff8000000000000LL);
      st_relaxed_sys(Rg->epoch, e);
    }
    return;
  }
  auto cell_at = [&](int owner, int src, int64_t b) {
    return reinterpret_cast<uint64_t*>(stage_of[owner]) + 4 * cap + (((int64_t)par * W + src) * cap + b) * 2;
  };
  // level 1: my partial, as a cell in my OWN rank's area [parity][r][b] (an ordinary local store)
  if (tid == 0) st_ll_cell(cell_at(r, r, blockIdx.x), red[32], e32);
  if (blockIdx.x != G - 1) return;
  // the collecting CTA: this rank's G partials in a fixed thread-strided order (four polls in flight
  // per thread), the block tree, then the rank totals travel as LL cells [parity][r][0] to the peers
  const int64_t limit = Rg->spin_limit > 0 ? Rg->spin_limit : kDefaultSpinLimit;
  double s = 0.0;
  for (int64_t q0 = tid; q0 < G; q0 += 4 * kCollThreads) {
    double v[4] = {0.0, 0.0, 0.0, 0.0};
    bool ok[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t q = q0 + (int64_t)k * kCollThreads;
      ok[k] = q >= G || ld_ll_cell(cell_at(r, r, q), e32, v[k]);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t q = q0 + (int64_t)k * kCollThreads;
      for (int64_t it = 0; !ok[k] && it < limit; ++it) {
        ok[k] = ld_ll_cell(cell_at(r, r, q), e32, v[k]);
        if (it > 64) __nanosleep(32);
      }
      if (!ok[k]) atomicExch(Rg->status, 1u + r);
      s += v[k];
    }
  }
  const double mine = block_sum(s, red);
  if (tid == 0) red[33] = mine;
  __syncthreads();
  double theirs = red[33];
  if (tid < W && tid != r) {
    st_ll_cell(cell_at(tid, r, 0), red[33], e32);     // level 2: my total into every peer's area
    const uint64_t* cell = cell_at(r, tid, 0);
    bool ok = false;
    for (int64_t it = 0; it < limit; ++it) {
      if (ld_ll_cell(cell, e32, theirs)) {
        ok = true;
        break;
      }
      if (it > 64) __nanosleep(32);
    }
    if (!ok) atomicExch(Rg->status, 1u + tid);
  }
  if (tid < W) red[tid] = theirs;
  __syncthreads();
  if (tid == 0) {
    double total = 0.0;
    for (int p = 0; p < W; ++p) total += red[p];   // rank order: the same bits on every rank
    *Rg->out = total;
    st_relaxed_sys(Rg->epoch, e);   // every CTA of this launch read the epoch before it pushed its cell
  }
}

// One warp: flags and epoch are fetched side by side (two dependent loads instead of the generic
// prologue's three), the release stores leave at once, and nothing is reset afterwards because the
// barrier dirties no counter.
__global__ void __launch_bounds__(32) barrier_kernel(const gmt_coll_rank* __restrict__ ranks)
{
  const gmt_coll_rank* Rg = ranks + blockIdx.y;
  const int t = threadIdx.x;
  const int W = Rg->world, r = Rg->rank;
  uint64_t* const* tab = Rg->peer_flag;
  uint64_t* ep = Rg->epoch;
  const int64_t spin = Rg->spin_limit;
  uint64_t* theirs = t < W ? tab[t] : nullptr;
  uint64_t* mine = tab[r];
  const uint64_t e = ld_relaxed_sys(ep) + 1;
  if (t < W) {
    // relaxed is enough: the barrier kernel itself wrote nothing, and the stream's earlier kernels
    // completed (their stores performed at system scope) before this one started
    st_relaxed_sys(theirs + r, e);   // round 0, word [src]
    if (!wait_flag_ge(mine + t, e, spin)) atomicExch(Rg->status, 1u + t);
  }
  __syncwarp();
  if (t == 0) st_relaxed_sys(ep, e);
}

// ============================================================ launch helpers
// Test hook: single-rank launches get up to `factor` times more CTAs than can be resident at once, so
// that a kernel which (wrongly) waited for a sibling CTA of its own grid would expire its bounded spin.
static int g_debug_grid_factor = 1;

static int coll_grid(const void* kernel, int64_t work_ctas, int n_ranks, int* grid)
{
  int occ = 0;
  GMT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kCollThreads, 0));
  if (occ < 1) occ = 1;
  int64_t cap = (int64_t)sm_count() * occ;
  if (n_ranks > 1) cap /= n_ranks;
  else cap *= g_debug_grid_factor;
  GMT_REQUIRE(cap >= 1, GMT_ETOOBIG, "too many virtual ranks (%d) for one cooperative launch", n_ranks);
  *grid = (int)std::max<int64_t>(1, std::min<int64_t>(cap, work_ctas));
  return 0;
}

static int coll_launch(const void* kernel, int grid, int n_ranks, void** args, cudaStream_t st,
                       int threads = kCollThreads)
{
  if (n_ranks > 1)
    GMT_CUDA(cudaLaunchCooperativeKernel(kernel, dim3(grid, n_ranks), dim3(threads), args, 0, st));
  else
    GMT_CUDA(cudaLaunchKernel(kernel, dim3(grid, 1), dim3(threads), args, 0, st));
  return 0;
}

}  // namespace gmt

extern "C" {

int gmt_allgather_f64(const gmt_coll_rank* ranks_dev, int n_ranks, int64_t n, int from_in, int use_multicast,
                      gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(ranks_dev && n_ranks >= 1 && n >= 0, GMT_EINVAL, "gmt_allgather_f64: bad argument");
  GMT_REQUIRE(use_multicast >= 0 && use_multicast <= 4, GMT_EINVAL, "gmt_allgather_f64: unknown mode %d", use_multicast);
  int grid = 0, root = -1;
  if (use_multicast == 2) {   // LL cells
    GMT_REQUIRE(n <= kLLGatherMax, GMT_ETOOBIG, "gmt_allgather_f64: LL mode is for slices of up to %lld doubles",
                (long long)kLLGatherMax);
    const void* kl = (const void*)gather_ll_kernel;
    int rc = coll_grid(kl, (n + kCollThreads - 1) / kCollThreads, n_ranks, &grid);
    if (rc) return rc;
    void* largs[] = {(void*)&ranks_dev, (void*)&n, (void*)&from_in};
    return coll_launch(kl, grid, n_ranks, largs, as_stream(stream));
  }
  const void* k = (const void*)gather_kernel;
  int rc = coll_grid(k, (n / 4 + kCollThreads - 1) / kCollThreads + 1, n_ranks, &grid);
  if (rc) return rc;
  void* args[] = {(void*)&ranks_dev, (void*)&n, (void*)&from_in, (void*)&use_multicast, (void*)&root};
  return coll_launch(k, grid, n_ranks, args, as_stream(stream));
}

int gmt_gather_f64(const gmt_coll_rank* ranks_dev, int n_ranks, int64_t n, int root, gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(ranks_dev && n_ranks >= 1 && n >= 0 && root >= 0, GMT_EINVAL, "gmt_gather_f64: bad argument");
  int grid = 0, from_in = 0, use_mc = 0;
  const void* k = (const void*)gather_kernel;
  int rc = coll_grid(k, (n / 4 + kCollThreads - 1) / kCollThreads + 1, n_ranks, &grid);
  if (rc) return rc;
  void* args[] = {(void*)&ranks_dev, (void*)&n, (void*)&from_in, (void*)&use_mc, (void*)&root};
  return coll_launch(k, grid, n_ranks, args, as_stream(stream));
}

int gmt_allreduce_f64(const gmt_coll_rank* ranks_dev, int n_ranks, int64_t n, int algo, gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(ranks_dev && n_ranks >= 1 && n >= 0, GMT_EINVAL, "gmt_allreduce_f64: bad argument");
  GMT_REQUIRE(algo >= 0 && algo <= 3, GMT_EINVAL, "gmt_allreduce_f64: unknown algorithm %d", algo);
  const void* k = algo == 0   ? (const void*)allreduce_kernel<0>
                  : algo == 1 ? (const void*)allreduce_kernel<1>
                  : algo == 2 ? (const void*)allreduce_kernel<2>
                              : (const void*)allreduce_ll_kernel;
  int grid = 0;
  // LL needs no grid-wide rendezvous, so it can afford one element per thread; the others keep few CTAs
  const int64_t per_cta = algo == 3 ? kCollThreads : kCollThreads * 4;
  int rc = coll_grid(k, (n + per_cta - 1) / per_cta, n_ranks, &grid);
  if (rc) return rc;
  void* args[] = {(void*)&ranks_dev, (void*)&n};
  return coll_launch(k, grid, n_ranks, args, as_stream(stream));
}

int gmt_daxpy_allreduce_f64(const gmt_coll_rank* ranks_dev, int n_ranks, int64_t n, double a, gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(ranks_dev && n_ranks >= 1 && n >= 0, GMT_EINVAL, "gmt_daxpy_allreduce_f64: bad argument");
  const void* k = (const void*)daxpy_allreduce_kernel;
  // one chunk per CTA in address order; beyond kDaMaxCtas chunks (or the co-resident share of a
  // cooperative multi-rank launch) every CTA takes several consecutive chunks
  const int64_t nchunks = std::max<int64_t>(1, (n / 4 + kDaChunk - 1) / kDaChunk);
  int64_t cap = kDaMaxCtas;
  if (n_ranks > 1) {
    int g = 0;
    int rc = coll_grid(k, nchunks, n_ranks, &g);
    if (rc) return rc;
    cap = std::min<int64_t>(cap, g);
  }
  GMT_REQUIRE((nchunks + cap - 1) / cap < (1LL << 31), GMT_ETOOBIG, "gmt_daxpy_allreduce_f64: n too large");
  int per_cta = (int)((nchunks + cap - 1) / cap);
  int grid = (int)((nchunks + per_cta - 1) / per_cta);
  void* args[] = {(void*)&ranks_dev, (void*)&n, (void*)&a, (void*)&per_cta};
  return coll_launch(k, grid, n_ranks, args, as_stream(stream));
}

int gmt_axis_sum_allreduce_f64(const gmt_coll_rank* ranks_dev, int n_ranks, int64_t nx, int64_t ny, int64_t ld,
                               int axis, int64_t n_out, gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(ranks_dev && n_ranks >= 1 && nx > 0 && ny > 0 && ld >= nx, GMT_EINVAL,
              "gmt_axis_sum_allreduce_f64: bad argument");
  GMT_REQUIRE(axis == 0 || axis == 1, GMT_EINVAL, "gmt_axis_sum_allreduce_f64: axis must be 0 or 1");
  GMT_REQUIRE(n_out >= 0 && n_out <= (axis == 0 ? ny : nx), GMT_EINVAL,
              "gmt_axis_sum_allreduce_f64: n_out exceeds the surviving extent");
  // (LL cells while n_out x world is small, one-shot beyond: decided in the kernel, which knows the world size)
  const void* k = (const void*)axis_sum_allreduce_kernel;
  int grid = 0;
  const int64_t work = axis == 0 ? (n_out + kCollThreads / 32 - 1) / (kCollThreads / 32) : (n_out + 31) / 32;
  int rc = coll_grid(k, work, n_ranks, &grid);
  if (rc) return rc;
  void* args[] = {(void*)&ranks_dev, (void*)&nx, (void*)&ny, (void*)&ld, (void*)&axis, (void*)&n_out};
  return coll_launch(k, grid, n_ranks, args, as_stream(stream));
}

int gmt_debug_set_grid_factor(int factor)
{
  GMT_REQUIRE(factor >= 1 && factor <= 64, GMT_EINVAL, "gmt_debug_set_grid_factor: factor must be 1..64");
  gmt::g_debug_grid_factor = factor;
  return 0;
}

size_t gmt_coll_stage_bytes(int world, int64_t stage_cap)
{
  if (world < 1 || stage_cap < 1) return 0;
  // 2 input + 2 result vectors (parity double-buffered) and 2 x world x stage_cap 16-byte LL cells
  return size_t(4 + 4 * (int64_t)world) * size_t(stage_cap) * sizeof(double);
}

int64_t gmt_coll_min_stage_cap(void) { return gmt::kDaMaxCtas; }
int64_t gmt_coll_ll_gather_max(void) { return gmt::kLLGatherMax; }

int gmt_barrier(const gmt_coll_rank* ranks_dev, int n_ranks, gmt_stream_t stream)
{
  using namespace gmt;
  GMT_REQUIRE(ranks_dev && n_ranks >= 1, GMT_EINVAL, "gmt_barrier: bad argument");
  void* args[] = {(void*)&ranks_dev};
  return coll_launch((const void*)barrier_kernel, 1, n_ranks, args, as_stream(stream), 32);
}

}  // extern "C"
