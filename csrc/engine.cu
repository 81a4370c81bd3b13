// engine.cu -- native exchange engine: a whole process grid driven from C++.
//
// The reference's host orchestration is C++: every program computes its neighbours, allocates
// the ghosted array and the staging buffers, calls boundary_exchange and the stencil per
// iteration and reduces the timings (mpi_stencil2d_gt.cc:385-572).  This file is the
// B200-native equivalent for programs that do not want a Python layer: ONE process owns every
// GPU of the node, a "rank" is a tile of the px x py process grid placed on a device with the
// reference's block rule (set_rank_device, mpi_daxpy.cc:36-62: several ranks may share a GPU),
// neighbours are wired through CUDA peer access (same address space, so a neighbour's inbox
// pointer is simply used), and a step is one fused halo+stencil launch per device
// (stencil_fused.cu).  Ranks that share a device go into the same cooperative launch, so
// mutually waiting kernels are never separate launches on one GPU.
//
// What it owns per rank: the ghosted array z, the result out, a block holding the four inboxes
// (double-buffered by epoch parity) and their flag words, the device-resident epoch, the
// done/status words and the error-norm scratch.  Nothing is allocated after create().
#include <algorithm>
#include <cmath>
#include <cstring>
#include <memory>
#include <vector>

#include "common.cuh"
#include "engine_util.h"

namespace {

struct RankMem {
  int device = 0;
  int cx = 0, cy = 0;
  double* z = nullptr;
  double* out = nullptr;
  char* block = nullptr;  // inboxes + flags + control words + scratch, zero-initialised
  size_t off_inbox[4][2] = {};
  size_t off_flag[4] = {};
  size_t off_epoch = 0, off_done = 0, off_status = 0, off_err = 0, off_scratch = 0;
};

struct DeviceGroup {
  int device = 0;
  std::vector<int> ranks;
  gmt_halo2d_rank* desc = nullptr;  // device array, one entry per hosted rank
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
};

}  // namespace

struct gmt_engine2d {
  int px = 1, py = 1, gx = 0, gy = 0, with_err = 0;
  int64_t nx = 0, ny = 0, ldz = 0, spin_limit = 0, steps = 0;
  double dx = 0, dy = 0, sx = 0, sy = 0;
  std::vector<RankMem> ranks;
  std::vector<DeviceGroup> groups;
  std::unique_ptr<gmt::DeviceExecutor> exec;   // one launching thread per GPU
};

namespace {

int neighbour(const gmt_engine2d* e, int rank, int side)
{
  int cx = rank % e->px, cy = rank / e->px;
  if (side == 0) --cx;
  if (side == 1) ++cx;
  if (side == 2) --cy;
  if (side == 3) ++cy;
  return (cx >= 0 && cx < e->px && cy >= 0 && cy < e->py) ? cx + cy * e->px : -1;
}

using gmt::DeviceRestore;
using gmt::take;

int free_all(gmt_engine2d* e)
{
  e->exec.reset();
  for (auto& g : e->groups) {
    if (cudaSetDevice(g.device) != cudaSuccess) continue;
    if (g.stream) cudaStreamSynchronize(g.stream);
    if (g.desc) cudaFree(g.desc);
    if (g.ev0) cudaEventDestroy(g.ev0);
    if (g.ev1) cudaEventDestroy(g.ev1);
    if (g.stream) cudaStreamDestroy(g.stream);
  }
  for (auto& r : e->ranks) {
    if (cudaSetDevice(r.device) != cudaSuccess) continue;
    if (r.z) cudaFree(r.z);
    if (r.out) cudaFree(r.out);
    if (r.block) cudaFree(r.block);
  }
  (void)cudaGetLastError();
  delete e;
  return 0;
}

}  // namespace

extern "C" {

int gmt_engine_default_device(int n_ranks, int rank, int n_devices, int* device)
{
  GMT_REQUIRE(device && n_ranks > 0 && n_devices > 0 && rank >= 0 && rank < n_ranks, GMT_EINVAL,
              "gmt_engine_default_device: bad arguments");
  if (n_ranks > n_devices) {
    GMT_REQUIRE(n_ranks % n_devices == 0, GMT_EINVAL,
                "ERROR: Number of ranks (%d) not a multiple of number of GPUs (%d)", n_ranks, n_devices);
    *device = rank / (n_ranks / n_devices);
  } else {
    *device = rank;
  }
  return 0;
}

int gmt_engine2d_create(int px, int py, const int* devices, int64_t nx, int64_t ny, int gx, int gy, int with_err,
                        int64_t spin_limit, gmt_engine2d** out)
{
  DeviceRestore restore_;
  GMT_REQUIRE(out, GMT_EINVAL, "gmt_engine2d_create: NULL out");
  *out = nullptr;
  GMT_REQUIRE(px >= 1 && py >= 1 && px * py <= 1024, GMT_EINVAL, "gmt_engine2d_create: bad grid %dx%d", px, py);
  GMT_REQUIRE((gx == 0 || gx == 2) && (gy == 0 || gy == 2) && (gx || gy), GMT_EINVAL,
              "gmt_engine2d_create: ghost widths must be 0 or 2 and not both 0");
  GMT_REQUIRE(nx >= 8 && ny >= 4 && nx % 4 == 0, GMT_EINVAL,
              "gmt_engine2d_create: need nx >= 8, ny >= 4, nx %% 4 == 0 (got %lld x %lld)", (long long)nx,
              (long long)ny);
  GMT_REQUIRE(gx || px == 1, GMT_EINVAL, "gmt_engine2d_create: px > 1 needs ghosts in x");
  GMT_REQUIRE(gy || py == 1, GMT_EINVAL, "gmt_engine2d_create: py > 1 needs ghosts in y");
  int n_dev = 0;
  GMT_CUDA(cudaGetDeviceCount(&n_dev));
  GMT_REQUIRE(n_dev > 0, GMT_EUNSUPPORTED, "gmt_engine2d_create: no CUDA device");
  const int world = px * py;

  gmt_engine2d* e = new gmt_engine2d;
  e->px = px, e->py = py, e->gx = gx, e->gy = gy, e->with_err = with_err != 0;
  e->nx = nx, e->ny = ny, e->ldz = nx + 2 * gx, e->spin_limit = spin_limit;
  // domain [0, 8) in each decomposed direction, weak scaling: the per-rank tile is fixed
  // (mpi_stencil2d_gt.cc:395,427-431)
  e->dx = 8.0 / double(nx * px), e->dy = 8.0 / double(ny * py);
  // slab decompositions are the reference's tests: ONE spacing, that of the differentiated axis,
  // is used for both coordinates (delta = 8 / n_global_deriv, mpi_stencil2d_gt.cc:427-448)
  if (gx && !gy) e->dy = e->dx;
  if (gy && !gx) e->dx = e->dy;
  e->sx = gx ? 1.0 / e->dx : 0.0, e->sy = gy ? 1.0 / e->dy : 0.0;
  e->ranks.resize(world);

  int rc = 0;
  auto bail = [&](int code) {
    free_all(e);
    return code;
  };
  for (int r = 0; r < world; ++r) {
    int d = 0;
    if (devices) {
      d = devices[r];
      if (d < 0 || d >= n_dev) return bail(gmt::fail(GMT_EINVAL, "gmt_engine2d_create: rank %d on device %d of %d", r, d, n_dev));
    } else if ((rc = gmt_engine_default_device(world, r, n_dev, &d)) != 0) {
      return bail(rc);
    }
    e->ranks[r].device = d;
    e->ranks[r].cx = r % px, e->ranks[r].cy = r / px;
  }
  {
    std::vector<int> dev_of(world);
    for (int r = 0; r < world; ++r) dev_of[r] = e->ranks[r].device;
    for (const auto& dr : gmt::group_by_device(dev_of)) {
      e->groups.emplace_back();
      e->groups.back().device = dr.device;
      e->groups.back().ranks = dr.ranks;
    }
  }
  // peer access between every pair of devices that host neighbouring ranks
  for (int r = 0; r < world; ++r)
    for (int side = 0; side < 4; ++side) {
      const int nb = neighbour(e, r, side);
      if (nb < 0 || e->ranks[nb].device == e->ranks[r].device) continue;
      if (cudaSetDevice(e->ranks[r].device) != cudaSuccess) return bail(gmt::fail(GMT_EINVAL, "cudaSetDevice failed"));
      if ((rc = gmt_enable_peer_access(e->ranks[nb].device)) != 0) return bail(rc);
    }

  const int64_t nfx = gmt_halo2d_flag_count(0, nx, ny), nfy = gmt_halo2d_flag_count(2, nx, ny);
  const size_t box_x = size_t(2) * ny * 8, box_y = size_t(2) * nx * 8;
  const size_t scratch_bytes = gmt_reduce_scratch_bytes();

  for (int r = 0; r < world; ++r) {
    RankMem& m = e->ranks[r];
    cudaError_t ce = cudaSetDevice(m.device);
    size_t cur = 0;
    for (int side = 0; side < 4; ++side) {
      if (!(side < 2 ? gx : gy)) continue;
      for (int p = 0; p < 2; ++p) m.off_inbox[side][p] = take(cur, side < 2 ? box_x : box_y);
      m.off_flag[side] = take(cur, size_t(side < 2 ? nfx : nfy) * 8);
    }
    m.off_epoch = take(cur, 64), m.off_done = take(cur, 64), m.off_status = take(cur, 64);
    m.off_err = take(cur, 64), m.off_scratch = take(cur, scratch_bytes);
    if (ce == cudaSuccess) ce = cudaMalloc(&m.z, size_t(e->ldz) * (ny + 2 * gy) * 8);
    if (ce == cudaSuccess) ce = cudaMalloc(&m.out, size_t(nx) * ny * 8);
    if (ce == cudaSuccess) ce = cudaMalloc(&m.block, cur);
    if (ce == cudaSuccess) ce = cudaMemset(m.block, 0, cur);
    if (ce != cudaSuccess) return bail(gmt::cuda_fail(ce, "engine allocation", __FILE__, __LINE__));
  }

  for (auto& g : e->groups) {
    std::vector<gmt_halo2d_rank> desc(g.ranks.size());
    for (size_t k = 0; k < g.ranks.size(); ++k) {
      const int r = g.ranks[k];
      const RankMem& m = e->ranks[r];
      gmt_halo2d_rank& s = desc[k];
      memset(&s, 0, sizeof(s));
      s.z = m.z, s.out = m.out;
      s.nx = nx, s.ny = ny, s.ldz = e->ldz, s.ldo = nx, s.gx = gx, s.gy = gy, s.sx = e->sx, s.sy = e->sy;
      for (int side = 0; side < 4; ++side) {
        const bool used = side < 2 ? gx : gy;
        const int nb = used ? neighbour(e, r, side) : -1;
        s.has[side] = nb >= 0;
        const int opp = side ^ 1;
        for (int p = 0; p < 2; ++p) {
          s.my_inbox[side][p] = used ? (double*)(m.block + m.off_inbox[side][p]) : nullptr;
          s.peer_inbox[side][p] =
              nb >= 0 ? (double*)(e->ranks[nb].block + e->ranks[nb].off_inbox[opp][p]) : nullptr;
        }
        s.my_flag[side] = used ? (uint64_t*)(m.block + m.off_flag[side]) : nullptr;
        s.peer_flag[side] = nb >= 0 ? (uint64_t*)(e->ranks[nb].block + e->ranks[nb].off_flag[opp]) : nullptr;
      }
      s.epoch = (uint64_t*)(m.block + m.off_epoch);
      s.done = (uint32_t*)(m.block + m.off_done);
      s.status = (uint32_t*)(m.block + m.off_status);
      s.err_scratch = m.block + m.off_scratch;
      s.err_sumsq = (double*)(m.block + m.off_err);
      s.x0 = m.cx * nx * e->dx, s.y0 = m.cy * ny * e->dy, s.dx = e->dx, s.dy = e->dy;
      s.spin_limit = spin_limit;
    }
    cudaError_t ce = cudaSetDevice(g.device);
    if (ce == cudaSuccess) ce = cudaMalloc(&g.desc, desc.size() * sizeof(gmt_halo2d_rank));
    if (ce == cudaSuccess)
      ce = cudaMemcpy(g.desc, desc.data(), desc.size() * sizeof(gmt_halo2d_rank), cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking);
    if (ce == cudaSuccess) ce = cudaEventCreate(&g.ev0);
    if (ce == cudaSuccess) ce = cudaEventCreate(&g.ev1);
    if (ce != cudaSuccess) return bail(gmt::cuda_fail(ce, "engine group set-up", __FILE__, __LINE__));
  }
  {
    std::vector<int> devs;
    for (auto& g : e->groups) devs.push_back(g.device);
    e->exec.reset(new gmt::DeviceExecutor(devs));
  }
  *out = e;
  return 0;
}

int gmt_engine2d_destroy(gmt_engine2d* e)
{
  DeviceRestore restore_;
  if (!e) return 0;
  return free_all(e);
}

int gmt_engine2d_world(const gmt_engine2d* e) { return e ? e->px * e->py : 0; }
int gmt_engine2d_device_groups(const gmt_engine2d* e) { return e ? (int)e->groups.size() : 0; }

int gmt_engine2d_rank_device(const gmt_engine2d* e, int rank, int* device)
{
  GMT_REQUIRE(e && device && rank >= 0 && rank < e->px * e->py, GMT_EINVAL, "gmt_engine2d_rank_device: bad rank");
  *device = e->ranks[rank].device;
  return 0;
}

int gmt_engine2d_array(const gmt_engine2d* e, int rank, int which, double** ptr, int64_t* cols, int64_t* ld)
{
  GMT_REQUIRE(e && ptr && rank >= 0 && rank < e->px * e->py && (which == 0 || which == 1), GMT_EINVAL,
              "gmt_engine2d_array: bad arguments");
  const RankMem& m = e->ranks[rank];
  *ptr = which == 0 ? m.z : m.out;
  if (cols) *cols = which == 0 ? e->ny + 2 * e->gy : e->ny;
  if (ld) *ld = which == 0 ? e->ldz : e->nx;
  return 0;
}

int gmt_engine2d_init_analytic(gmt_engine2d* e, double poison)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e, GMT_EINVAL, "gmt_engine2d_init_analytic: NULL engine");
  for (auto& g : e->groups) {
    GMT_CUDA(cudaSetDevice(g.device));
    for (int r : g.ranks) {
      const RankMem& m = e->ranks[r];
      int mask = 0;
      for (int side = 0; side < 4; ++side)
        if (neighbour(e, r, side) < 0) mask |= 1 << side;
      const int rc = gmt_init_stencil2d_f64(e->nx, e->ny, e->gx, e->gy, m.cx * e->nx * e->dx, m.cy * e->ny * e->dy,
                                            e->dx, e->dy, m.z, e->ldz, nullptr, 0, 0, e->sx, e->sy, mask, poison,
                                            g.stream);
      if (rc) return rc;
    }
  }
  return gmt_engine2d_synchronize(e);
}

int gmt_engine2d_step(gmt_engine2d* e)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e, GMT_EINVAL, "gmt_engine2d_step: NULL engine");
  // one launch per GPU, issued by that GPU's worker thread so that all GPUs start together; the
  // launches are asynchronous and the kernels' flag protocol orders everything else
  const int rc = e->exec->run([e](int i) {
    DeviceGroup& g = e->groups[i];
    return gmt_fused_stencil2d_f64(g.desc, (int)g.ranks.size(), e->nx, e->ny, e->gx, e->gy, e->with_err, g.stream);
  });
  if (rc) return rc;
  ++e->steps;
  return 0;
}

int gmt_engine2d_exchange(gmt_engine2d* e)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e, GMT_EINVAL, "gmt_engine2d_exchange: NULL engine");
  // the halo exchange alone (boundary_exchange_x/_y, mpi_stencil2d_gt.cc:136-373): one small kernel per GPU
  const int rc = e->exec->run([e](int i) {
    DeviceGroup& g = e->groups[i];
    return gmt_halo_exchange2d_f64(g.desc, (int)g.ranks.size(), e->nx, e->ny, e->gx, e->gy, g.stream);
  });
  if (rc) return rc;
  ++e->steps;
  return 0;
}

int gmt_engine2d_synchronize(gmt_engine2d* e)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e, GMT_EINVAL, "gmt_engine2d_synchronize: NULL engine");
  for (auto& g : e->groups) {
    GMT_CUDA(cudaSetDevice(g.device));
    GMT_CUDA(cudaStreamSynchronize(g.stream));
  }
  return 0;
}

int gmt_engine2d_status(gmt_engine2d* e, int* first_bad_rank)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e, GMT_EINVAL, "gmt_engine2d_status: NULL engine");
  if (first_bad_rank) *first_bad_rank = -1;
  for (int r = 0; r < e->px * e->py; ++r) {
    const RankMem& m = e->ranks[r];
    uint32_t st = 0;
    GMT_CUDA(cudaSetDevice(m.device));
    GMT_CUDA(cudaMemcpy(&st, m.block + m.off_status, 4, cudaMemcpyDeviceToHost));
    if (st) {
      if (first_bad_rank) *first_bad_rank = r;
      return gmt::fail(GMT_ETIMEOUT, "rank %d: a halo flag never arrived (status %u)", r, st);
    }
  }
  return 0;
}

static int time_calls(gmt_engine2d* e, int (*call)(gmt_engine2d*), int steps, int warmup, double* ms_per_step);

int gmt_engine2d_time_steps(gmt_engine2d* e, int steps, int warmup, double* ms_per_step)
{
  GMT_REQUIRE(e && ms_per_step && steps > 0 && warmup >= 0, GMT_EINVAL, "gmt_engine2d_time_steps: bad arguments");
  return time_calls(e, gmt_engine2d_step, steps, warmup, ms_per_step);
}

int gmt_engine2d_time_exchanges(gmt_engine2d* e, int steps, int warmup, double* ms_per_exchange)
{
  GMT_REQUIRE(e && ms_per_exchange && steps > 0 && warmup >= 0, GMT_EINVAL,
              "gmt_engine2d_time_exchanges: bad arguments");
  return time_calls(e, gmt_engine2d_exchange, steps, warmup, ms_per_exchange);
}

static int time_calls(gmt_engine2d* e, int (*call)(gmt_engine2d*), int steps, int warmup, double* ms_per_step)
{
  DeviceRestore restore_;
  int rc = 0;
  for (int i = 0; i < warmup; ++i)
    if ((rc = call(e)) != 0) return rc;
  if ((rc = gmt_engine2d_synchronize(e)) != 0) return rc;
  for (auto& g : e->groups) {
    GMT_CUDA(cudaSetDevice(g.device));
    GMT_CUDA(cudaEventRecord(g.ev0, g.stream));
  }
  for (int i = 0; i < steps; ++i)
    if ((rc = call(e)) != 0) return rc;
  for (auto& g : e->groups) {
    GMT_CUDA(cudaSetDevice(g.device));
    GMT_CUDA(cudaEventRecord(g.ev1, g.stream));
  }
  if ((rc = gmt_engine2d_synchronize(e)) != 0) return rc;
  float worst = 0.f;   // device time, max over the GPUs
  for (auto& g : e->groups) {
    float ms = 0.f;
    GMT_CUDA(cudaSetDevice(g.device));
    GMT_CUDA(cudaEventElapsedTime(&ms, g.ev0, g.ev1));
    worst = std::max(worst, ms);
  }
  *ms_per_step = double(worst) / steps;
  return gmt_engine2d_status(e, nullptr);
}

int gmt_engine2d_err_norm(gmt_engine2d* e, int rank, double* err_norm)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e && err_norm && rank >= 0 && rank < e->px * e->py, GMT_EINVAL, "gmt_engine2d_err_norm: bad arguments");
  GMT_REQUIRE(e->with_err, GMT_EUNSUPPORTED, "gmt_engine2d_err_norm: engine created with with_err = 0");
  const RankMem& m = e->ranks[rank];
  double ss = 0.0;
  GMT_CUDA(cudaSetDevice(m.device));
  GMT_CUDA(cudaMemcpy(&ss, m.block + m.off_err, 8, cudaMemcpyDeviceToHost));
  *err_norm = std::sqrt(ss);
  return 0;
}

}  // extern "C"
