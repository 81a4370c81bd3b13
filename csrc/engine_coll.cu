// engine_coll.cu -- native collective engine: the in-kernel collectives of collectives.cu
// driven from ONE C/C++ process that owns every GPU of the node.
//
// Counterpart of the host side of the reference's daxpy benchmark and gather test
// (mpi_daxpy_nvtx.cc:85-343, mpigatherinplace.f90:1-58): allocate the symmetric buffers,
// place the ranks on devices, call the collective, time it.  Here a "rank" is a slot on a GPU
// (block rule of set_rank_device when there are more ranks than GPUs), the symmetric buffers of
// all ranks live in one address space and are reached through CUDA peer access, and each
// collective is one kernel launch per GPU (one cooperative launch for the ranks that share it).
// NVLS multicast needs driver-API handles that this small engine does not create, so the
// multicast variants stay with the torch symmetric-memory path (parallel/collectives.py).
#include <algorithm>
#include <cstring>
#include <memory>
#include <vector>

#include "common.cuh"
#include "engine_util.h"

using gmt::DeviceRestore;
using gmt::take;

namespace {

struct CollRankMem {
  int device = 0;
  double* buf = nullptr;    // symmetric user buffer
  char* block = nullptr;    // staging + flags + control words + scratch, zero-initialised
  size_t off_stage = 0, off_flags = 0, off_epoch = 0, off_done = 0, off_status = 0, off_scratch = 0;
};

struct CollGroup {
  int device = 0;
  std::vector<int> ranks;
  gmt_coll_rank* desc = nullptr;   // device array, one entry per hosted rank
  void** tables = nullptr;         // device: peer_buf[world], peer_stage[world], peer_flag[world]
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
};

}  // namespace

struct gmt_coll_engine {
  int world = 0;
  int64_t buf_doubles = 0, stage_cap = 0, spin_limit = 0;
  std::vector<CollRankMem> ranks;
  std::vector<CollGroup> groups;
  std::vector<int> group_of, slot_of;   // rank -> group index, index inside the group
  std::unique_ptr<gmt::DeviceExecutor> exec;   // one launching thread per GPU
};

namespace {

int free_all(gmt_coll_engine* e)
{
  e->exec.reset();
  for (auto& g : e->groups) {
    if (cudaSetDevice(g.device) != cudaSuccess) continue;
    if (g.stream) cudaStreamSynchronize(g.stream);
    if (g.desc) cudaFree(g.desc);
    if (g.tables) cudaFree(g.tables);
    if (g.ev0) cudaEventDestroy(g.ev0);
    if (g.ev1) cudaEventDestroy(g.ev1);
    if (g.stream) cudaStreamDestroy(g.stream);
  }
  for (auto& r : e->ranks) {
    if (cudaSetDevice(r.device) != cudaSuccess) continue;
    if (r.buf) cudaFree(r.buf);
    if (r.block) cudaFree(r.block);
  }
  (void)cudaGetLastError();
  delete e;
  return 0;
}

// run `launch(group)` on every device at once (DeviceExecutor: the GPU's worker thread issues it, so
// the GPUs start together instead of 4-5 us apart); the launches themselves are asynchronous
template <class F>
int for_each_group(gmt_coll_engine* e, F&& launch)
{
  return e->exec->run([&](int i) { return launch(e->groups[i]); });
}

}  // namespace

extern "C" {

int gmt_coll_engine_create(int world, const int* devices, int64_t buf_doubles, int64_t stage_cap, int64_t spin_limit,
                           gmt_coll_engine** out)
{
  DeviceRestore restore_;
  GMT_REQUIRE(out, GMT_EINVAL, "gmt_coll_engine_create: NULL out");
  *out = nullptr;
  GMT_REQUIRE(world >= 1 && world <= GMT_COLL_MAX_WORLD, GMT_EINVAL, "gmt_coll_engine_create: world %d outside 1..%d",
              world, GMT_COLL_MAX_WORLD);
  GMT_REQUIRE(buf_doubles >= 0 && stage_cap >= 0, GMT_EINVAL, "gmt_coll_engine_create: negative size");
  int n_dev = 0;
  GMT_CUDA(cudaGetDeviceCount(&n_dev));
  GMT_REQUIRE(n_dev > 0, GMT_EUNSUPPORTED, "gmt_coll_engine_create: no CUDA device");

  gmt_coll_engine* e = new gmt_coll_engine;
  e->world = world;
  e->buf_doubles = std::max<int64_t>(buf_doubles, 4);
  // at least 2048 cells per source rank: daxpy + global sum stores one partial per CTA (up to 2048)
  e->stage_cap = std::max<int64_t>(stage_cap, gmt_coll_min_stage_cap());
  e->spin_limit = spin_limit;
  e->ranks.resize(world);
  e->group_of.resize(world), e->slot_of.resize(world);
  int rc = 0;
  auto bail = [&](int code) {
    free_all(e);
    return code;
  };
  std::vector<int> dev_of(world);
  for (int r = 0; r < world; ++r) {
    int d = 0;
    if (devices) {
      d = devices[r];
      if (d < 0 || d >= n_dev)
        return bail(gmt::fail(GMT_EINVAL, "gmt_coll_engine_create: rank %d on device %d of %d", r, d, n_dev));
    } else if ((rc = gmt_engine_default_device(world, r, n_dev, &d)) != 0) {
      return bail(rc);
    }
    e->ranks[r].device = dev_of[r] = d;
  }
  for (const auto& dr : gmt::group_by_device(dev_of)) {
    e->groups.emplace_back();
    e->groups.back().device = dr.device;
    e->groups.back().ranks = dr.ranks;
    for (size_t k = 0; k < dr.ranks.size(); ++k) {
      e->group_of[dr.ranks[k]] = (int)e->groups.size() - 1;
      e->slot_of[dr.ranks[k]] = (int)k;
    }
  }
  // every rank stores into / reads from every other rank: all device pairs need peer access
  for (auto& a : e->groups)
    for (auto& b : e->groups) {
      if (a.device == b.device) continue;
      if (cudaSetDevice(a.device) != cudaSuccess) return bail(gmt::fail(GMT_EINVAL, "cudaSetDevice failed"));
      if ((rc = gmt_enable_peer_access(b.device)) != 0) return bail(rc);
    }

  const size_t scratch_bytes = gmt_reduce_scratch_bytes();
  for (int r = 0; r < world; ++r) {
    CollRankMem& m = e->ranks[r];
    size_t cur = 0;
    m.off_stage = take(cur, gmt_coll_stage_bytes(world, e->stage_cap));   // vectors + LL cells
    m.off_flags = take(cur, size_t(GMT_COLL_FLAG_WORDS) * 8);
    m.off_epoch = take(cur, 64), m.off_done = take(cur, 64), m.off_status = take(cur, 64);
    m.off_scratch = take(cur, scratch_bytes);
    cudaError_t ce = cudaSetDevice(m.device);
    if (ce == cudaSuccess) ce = cudaMalloc(&m.buf, size_t(e->buf_doubles) * 8);
    if (ce == cudaSuccess) ce = cudaMalloc(&m.block, cur);
    if (ce == cudaSuccess) ce = cudaMemset(m.block, 0, cur);
    if (ce != cudaSuccess) return bail(gmt::cuda_fail(ce, "collective engine allocation", __FILE__, __LINE__));
  }
  std::vector<void*> tables(size_t(3) * world);
  for (int r = 0; r < world; ++r) {
    tables[r] = e->ranks[r].buf;
    tables[world + r] = e->ranks[r].block + e->ranks[r].off_stage;
    tables[2 * world + r] = e->ranks[r].block + e->ranks[r].off_flags;
  }
  for (auto& g : e->groups) {
    cudaError_t ce = cudaSetDevice(g.device);
    if (ce == cudaSuccess) ce = cudaMalloc(&g.tables, tables.size() * sizeof(void*));
    if (ce == cudaSuccess)
      ce = cudaMemcpy(g.tables, tables.data(), tables.size() * sizeof(void*), cudaMemcpyHostToDevice);
    std::vector<gmt_coll_rank> desc(g.ranks.size());
    for (size_t k = 0; k < g.ranks.size(); ++k) {
      const int r = g.ranks[k];
      const CollRankMem& m = e->ranks[r];
      gmt_coll_rank& s = desc[k];
      memset(&s, 0, sizeof(s));
      s.rank = r, s.world = world;
      s.peer_buf = (double**)g.tables;
      s.peer_stage = (double**)(g.tables + world);
      s.peer_flag = (uint64_t**)(g.tables + 2 * world);
      s.stage_cap = e->stage_cap;
      s.epoch = (uint64_t*)(m.block + m.off_epoch);
      s.done = (uint32_t*)(m.block + m.off_done);
      s.status = (uint32_t*)(m.block + m.off_status);
      s.scratch = m.block + m.off_scratch;
      s.spin_limit = spin_limit;
    }
    if (ce == cudaSuccess) ce = cudaMalloc(&g.desc, desc.size() * sizeof(gmt_coll_rank));
    if (ce == cudaSuccess)
      ce = cudaMemcpy(g.desc, desc.data(), desc.size() * sizeof(gmt_coll_rank), cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking);
    if (ce == cudaSuccess) ce = cudaEventCreate(&g.ev0);
    if (ce == cudaSuccess) ce = cudaEventCreate(&g.ev1);
    if (ce != cudaSuccess) return bail(gmt::cuda_fail(ce, "collective engine group set-up", __FILE__, __LINE__));
  }
  {
    std::vector<int> devs;
    for (auto& g : e->groups) devs.push_back(g.device);
    e->exec.reset(new gmt::DeviceExecutor(devs));
  }
  *out = e;
  return 0;
}

int gmt_coll_engine_destroy(gmt_coll_engine* e)
{
  DeviceRestore restore_;
  if (!e) return 0;
  return free_all(e);
}

int gmt_coll_engine_world(const gmt_coll_engine* e) { return e ? e->world : 0; }
int gmt_coll_engine_device_groups(const gmt_coll_engine* e) { return e ? (int)e->groups.size() : 0; }

int gmt_coll_engine_rank_device(const gmt_coll_engine* e, int rank, int* device)
{
  GMT_REQUIRE(e && device && rank >= 0 && rank < e->world, GMT_EINVAL, "gmt_coll_engine_rank_device: bad rank");
  *device = e->ranks[rank].device;
  return 0;
}

int gmt_coll_engine_buffer(const gmt_coll_engine* e, int rank, double** ptr, int64_t* doubles)
{
  GMT_REQUIRE(e && ptr && rank >= 0 && rank < e->world, GMT_EINVAL, "gmt_coll_engine_buffer: bad arguments");
  *ptr = e->ranks[rank].buf;
  if (doubles) *doubles = e->buf_doubles;
  return 0;
}

int gmt_coll_engine_stream(const gmt_coll_engine* e, int rank, gmt_stream_t* stream)
{
  GMT_REQUIRE(e && stream && rank >= 0 && rank < e->world, GMT_EINVAL, "gmt_coll_engine_stream: bad arguments");
  *stream = (gmt_stream_t)e->groups[e->group_of[rank]].stream;
  return 0;
}

int gmt_coll_engine_bind(gmt_coll_engine* e, int rank, const double* in, double* out, double* y)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e && rank >= 0 && rank < e->world, GMT_EINVAL, "gmt_coll_engine_bind: bad rank");
  CollGroup& g = e->groups[e->group_of[rank]];
  gmt_coll_rank* d = g.desc + e->slot_of[rank];
  const void* vals[3] = {in, out, y};
  static_assert(offsetof(gmt_coll_rank, out) == offsetof(gmt_coll_rank, in) + 8 &&
                    offsetof(gmt_coll_rank, y) == offsetof(gmt_coll_rank, in) + 16,
                "in/out/y must be adjacent");
  GMT_CUDA(cudaSetDevice(g.device));
  // stream-ordered with the launches; the source is pageable host memory, so the copy has
  // consumed `vals` by the time the call returns
  GMT_CUDA(cudaMemcpyAsync((char*)d + offsetof(gmt_coll_rank, in), vals, sizeof(vals), cudaMemcpyHostToDevice,
                           g.stream));
  return 0;
}

int gmt_coll_engine_allgather(gmt_coll_engine* e, int64_t n, int from_in)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e, GMT_EINVAL, "gmt_coll_engine_allgather: NULL engine");
  GMT_REQUIRE(n >= 0 && n * e->world <= e->buf_doubles, GMT_ETOOBIG,
              "buffer of %lld doubles cannot hold %d slices of %lld", (long long)e->buf_doubles, e->world,
              (long long)n);
  return for_each_group(e, [&](CollGroup& g) {
    // slices up to 64 KiB travel as LL cells (one one-way trip, no signalling rounds)
    const int mode = (e->world > 1 && n <= std::min<int64_t>(e->stage_cap, gmt_coll_ll_gather_max())) ? 2 : 0;
    return gmt_allgather_f64(g.desc, (int)g.ranks.size(), n, from_in, mode, g.stream);
  });
}

int gmt_coll_engine_gather(gmt_coll_engine* e, int64_t n, int root)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e && root >= 0 && root < e->world, GMT_EINVAL, "gmt_coll_engine_gather: bad root");
  GMT_REQUIRE(n >= 0 && n * e->world <= e->buf_doubles, GMT_ETOOBIG,
              "buffer of %lld doubles cannot hold %d slices of %lld", (long long)e->buf_doubles, e->world,
              (long long)n);
  return for_each_group(e, [&](CollGroup& g) { return gmt_gather_f64(g.desc, (int)g.ranks.size(), n, root, g.stream); });
}

int gmt_coll_engine_allreduce(gmt_coll_engine* e, int64_t n, int algo)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e && (algo == 0 || algo == 1 || algo == 3), GMT_EINVAL,
              "gmt_coll_engine_allreduce: algo must be 0, 1 or 3 (2 needs NVLS multicast)");
  GMT_REQUIRE(n >= 0 && n <= e->stage_cap, GMT_ETOOBIG, "all-reduce of %lld doubles exceeds the staging capacity %lld",
              (long long)n, (long long)e->stage_cap);
  return for_each_group(e, [&](CollGroup& g) { return gmt_allreduce_f64(g.desc, (int)g.ranks.size(), n, algo, g.stream); });
}

int gmt_coll_engine_daxpy_allreduce(gmt_coll_engine* e, int64_t n, double a)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e, GMT_EINVAL, "gmt_coll_engine_daxpy_allreduce: NULL engine");
  return for_each_group(e, [&](CollGroup& g) {
    return gmt_daxpy_allreduce_f64(g.desc, (int)g.ranks.size(), n, a, g.stream);
  });
}

int gmt_coll_engine_axis_sum_allreduce(gmt_coll_engine* e, int64_t nx, int64_t ny, int64_t ld, int axis, int64_t n_out)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e, GMT_EINVAL, "gmt_coll_engine_axis_sum_allreduce: NULL engine");
  GMT_REQUIRE(n_out >= 0 && n_out <= e->stage_cap, GMT_ETOOBIG,
              "all-reduce of %lld doubles exceeds the staging capacity %lld", (long long)n_out, (long long)e->stage_cap);
  return for_each_group(e, [&](CollGroup& g) {
    return gmt_axis_sum_allreduce_f64(g.desc, (int)g.ranks.size(), nx, ny, ld, axis, n_out, g.stream);
  });
}

int gmt_coll_engine_barrier(gmt_coll_engine* e)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e, GMT_EINVAL, "gmt_coll_engine_barrier: NULL engine");
  return for_each_group(e, [&](CollGroup& g) { return gmt_barrier(g.desc, (int)g.ranks.size(), g.stream); });
}

int gmt_coll_engine_synchronize(gmt_coll_engine* e)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e, GMT_EINVAL, "gmt_coll_engine_synchronize: NULL engine");
  return for_each_group(e, [&](CollGroup& g) {
    GMT_CUDA(cudaStreamSynchronize(g.stream));
    return 0;
  });
}

int gmt_coll_engine_status(gmt_coll_engine* e, int* first_bad_rank)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e, GMT_EINVAL, "gmt_coll_engine_status: NULL engine");
  if (first_bad_rank) *first_bad_rank = -1;
  for (int r = 0; r < e->world; ++r) {
    const CollRankMem& m = e->ranks[r];
    uint32_t st = 0;
    GMT_CUDA(cudaSetDevice(m.device));
    GMT_CUDA(cudaMemcpy(&st, m.block + m.off_status, 4, cudaMemcpyDeviceToHost));
    if (st) {
      if (first_bad_rank) *first_bad_rank = r;
      return gmt::fail(GMT_ETIMEOUT, "rank %d: a collective wait expired (status %u = 1 + silent rank)", r, st);
    }
  }
  return 0;
}

/* event pair on every GPU's stream: mark(0) before, mark(1) after the launches to be timed */
int gmt_coll_engine_mark(gmt_coll_engine* e, int which)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e && (which == 0 || which == 1), GMT_EINVAL, "gmt_coll_engine_mark: which must be 0 or 1");
  return for_each_group(e, [&](CollGroup& g) {
    GMT_CUDA(cudaEventRecord(which ? g.ev1 : g.ev0, g.stream));
    return 0;
  });
}

int gmt_coll_engine_elapsed_ms(gmt_coll_engine* e, double* ms)
{
  DeviceRestore restore_;
  GMT_REQUIRE(e && ms, GMT_EINVAL, "gmt_coll_engine_elapsed_ms: bad arguments");
  std::vector<float> t(e->groups.size(), 0.f);   // one slot per GPU: the workers run concurrently
  const int rc = e->exec->run([&](int i) {
    CollGroup& g = e->groups[i];
    GMT_CUDA(cudaEventSynchronize(g.ev1));
    GMT_CUDA(cudaEventElapsedTime(&t[i], g.ev0, g.ev1));
    return 0;
  });
  *ms = t.empty() ? 0.0 : *std::max_element(t.begin(), t.end());
  return rc;
}

}  // extern "C"
