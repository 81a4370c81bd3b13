// gmt_daxpy_gather -- native counterpart of the reference's main daxpy benchmark
// mpi_daxpy_nvtx (mpi_daxpy_nvtx.cc:85-343), one process driving every GPU of the node.
//
// Per rank: x[i] = (i+1)/n in its slot of allx, y = -x; y <- 2x + y (so y == x) with the local
// sum out of the same kernel pass; in-place allgather of x (MPI_Allgather(MPI_IN_PLACE), :285),
// out-of-place allgather of y (:288), sum of the gathered y.  Same printed lines as the reference
// ("SUM", "ALLSUM", four "TIME" lines per rank); times are device times in seconds, the gather
// time is the max over GPUs (a collective is as slow as its slowest rank).
//
//   gmt_daxpy_gather [--n N] [--ranks R] [--barrier] [--iters K]
//
// default n = 48 Mi / ranks doubles per rank (n_per_node = 48*2^20, mpi_daxpy_nvtx.cc:86,131).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "gmt_b200.h"

#define CHECK(what, call)                                                                      \
  do {                                                                                         \
    int rc_ = (call);                                                                          \
    if (rc_ != 0) {                                                                            \
      std::fprintf(stderr, "%s:%d error %d (%s): %s\n", __FILE__, __LINE__, rc_, what, gmt_last_error()); \
      return 2;                                                                                \
    }                                                                                          \
  } while (0)

static double now_s()
{
  using clk = std::chrono::steady_clock;
  return std::chrono::duration<double>(clk::now().time_since_epoch()).count();
}

int main(int argc, char** argv)
{
  long long n = 0;
  int ranks = 0, iters = 1;
  bool barrier = false;
  for (int i = 1; i < argc; ++i) {
    if (!std::strcmp(argv[i], "--n") && i + 1 < argc) n = std::atoll(argv[++i]);
    else if (!std::strcmp(argv[i], "--ranks") && i + 1 < argc) ranks = std::atoi(argv[++i]);
    else if (!std::strcmp(argv[i], "--iters") && i + 1 < argc) iters = std::atoi(argv[++i]);
    else if (!std::strcmp(argv[i], "--barrier")) barrier = true;
    else {
      std::fprintf(stderr, "usage: gmt_daxpy_gather [--n N] [--ranks R] [--barrier] [--iters K]\n");
      return 1;
    }
  }
  const double t_start = now_s();
  int n_dev = 0;
  CHECK("device count", gmt_device_count(&n_dev));
  if (n_dev == 0) {
    std::fprintf(stderr, "gmt_daxpy_gather: no CUDA device\n");
    return 2;
  }
  if (ranks <= 0) ranks = n_dev;
  if (n <= 0) n = (48LL << 20) / ranks;
  const int W = ranks;
  int probe = 0;
  if (gmt_engine_default_device(W, 0, n_dev, &probe) != 0) {
    std::fprintf(stderr, "%s\n", gmt_last_error());
    return 1;
  }
  std::printf("%d nodes, %d ranks, %lld elements each, total %lld\n", 1, W, n, n * W);

  // two engines = two symmetric arrays per rank: allx (gathered in place) and ally
  gmt_coll_engine *ex = nullptr, *ey = nullptr;
  CHECK("engine x", gmt_coll_engine_create(W, nullptr, n * W, 8, 0, &ex));
  CHECK("engine y", gmt_coll_engine_create(W, nullptr, n * W, 8, 0, &ey));
  std::vector<double*> allx(W), ally(W), d_y(W), d_sum(W);
  std::vector<void*> scratch(W);
  std::vector<gmt_stream_t> st(W);
  std::vector<int> dev(W);
  for (int r = 0; r < W; ++r) {
    CHECK("buffer", gmt_coll_engine_buffer(ex, r, &allx[r], nullptr));
    CHECK("buffer", gmt_coll_engine_buffer(ey, r, &ally[r], nullptr));
    CHECK("stream", gmt_coll_engine_stream(ex, r, &st[r]));
    CHECK("device", gmt_coll_engine_rank_device(ex, r, &dev[r]));
    CHECK("set device", gmt_set_device(dev[r]));
    std::printf("RANK[%d/%d] => DEVICE[%d/%d]\n", r + 1, W, dev[r] + 1, n_dev);
    CHECK("d_y", gmt_alloc(GMT_SPACE_DEVICE, size_t(n) * 8, (void**)&d_y[r]));
    CHECK("d_sum", gmt_alloc(GMT_SPACE_DEVICE, 64, (void**)&d_sum[r]));
    CHECK("scratch", gmt_alloc(GMT_SPACE_DEVICE, gmt_reduce_scratch_bytes(), &scratch[r]));
    CHECK("scratch zero", gmt_memset_async(scratch[r], 0, gmt_reduce_scratch_bytes(), st[r]));
  }
  {
    uint64_t free_b = 0, total_b = 0;
    CHECK("set device", gmt_set_device(dev[0]));
    CHECK("mem info", gmt_mem_info(&free_b, &total_b));
    std::printf("GPU memory %0.3f / %0.3f (%0.3f) MB\n", free_b / 1048576.0, total_b / 1048576.0,
                (total_b - free_b) / 1048576.0);
  }

  double k_ms = 0.0, b_ms = 0.0, g_ms = 0.0;
  std::vector<double> sum(W), allsum(W);
  for (int it = 0; it < iters; ++it) {
    gmt_nvtx_push("initializeArrays");
    for (int r = 0; r < W; ++r) {
      CHECK("set device", gmt_set_device(dev[r]));
      CHECK("init x", gmt_iota_f64(n, 1.0, 1.0, 1.0 / double(n), allx[r] + r * n, st[r]));
      CHECK("init y", gmt_iota_f64(n, 1.0, 1.0, -1.0 / double(n), d_y[r], st[r]));
    }
    CHECK("sync", gmt_coll_engine_synchronize(ex));
    gmt_nvtx_pop();

    gmt_nvtx_push("daxpy");
    CHECK("mark", gmt_coll_engine_mark(ex, 0));
    for (int r = 0; r < W; ++r) {
      CHECK("set device", gmt_set_device(dev[r]));
      CHECK("daxpy", gmt_daxpy_sum_f64(n, 2.0, allx[r] + r * n, d_y[r], d_sum[r], scratch[r], st[r]));
    }
    CHECK("mark", gmt_coll_engine_mark(ex, 1));
    CHECK("elapsed", gmt_coll_engine_elapsed_ms(ex, &k_ms));
    gmt_nvtx_pop();

    if (barrier) {
      gmt_nvtx_push("mpiBarrier");
      CHECK("mark", gmt_coll_engine_mark(ex, 0));
      CHECK("barrier", gmt_coll_engine_barrier(ex));
      CHECK("mark", gmt_coll_engine_mark(ex, 1));
      CHECK("elapsed", gmt_coll_engine_elapsed_ms(ex, &b_ms));
      gmt_nvtx_pop();
    }

    gmt_nvtx_push("mpiAllGather");
    double gx_ms = 0.0, gy_ms = 0.0;
    CHECK("mark", gmt_coll_engine_mark(ex, 0));
    CHECK("allgather x", gmt_coll_engine_allgather(ex, n, 0));
    CHECK("mark", gmt_coll_engine_mark(ex, 1));
    CHECK("elapsed", gmt_coll_engine_elapsed_ms(ex, &gx_ms));
    for (int r = 0; r < W; ++r) CHECK("bind", gmt_coll_engine_bind(ey, r, d_y[r], nullptr, nullptr));
    CHECK("mark", gmt_coll_engine_mark(ey, 0));
    CHECK("allgather y", gmt_coll_engine_allgather(ey, n, 1));
    CHECK("mark", gmt_coll_engine_mark(ey, 1));
    CHECK("elapsed", gmt_coll_engine_elapsed_ms(ey, &gy_ms));
    g_ms = gx_ms + gy_ms;
    gmt_nvtx_pop();
    CHECK("status x", gmt_coll_engine_status(ex, nullptr));
    CHECK("status y", gmt_coll_engine_status(ey, nullptr));
  }

  // sums: the local one came out of the daxpy kernel; the gathered y is summed on the device
  for (int r = 0; r < W; ++r) {
    gmt_stream_t sy = nullptr;
    CHECK("stream", gmt_coll_engine_stream(ey, r, &sy));
    CHECK("set device", gmt_set_device(dev[r]));
    CHECK("D2H sum", gmt_memcpy_async(&sum[r], d_sum[r], 8, 2, st[r]));
    CHECK("sync", gmt_stream_synchronize(st[r]));
    CHECK("allsum", gmt_sum_f64(n * W, ally[r], d_sum[r], scratch[r], sy));
    CHECK("D2H allsum", gmt_memcpy_async(&allsum[r], d_sum[r], 8, 2, sy));
    CHECK("sync", gmt_stream_synchronize(sy));
  }
  int bad = 0;
  for (int r = 0; r < W; ++r) {
    std::printf("%d/%d SUM = %f\n", r, W, sum[r]);
    std::printf("%d/%d ALLSUM = %f\n", r, W, allsum[r]);
    if (std::fabs(sum[r] - (n + 1) / 2.0) > 1e-6 * n || std::fabs(allsum[r] - W * (n + 1) / 2.0) > 1e-6 * n * W) bad = 1;
  }
  // x gathered in place: every rank must hold every other rank's slot ((i+1)/n everywhere)
  for (int r = 0; r < W; ++r) {
    double xs = 0.0;
    gmt_stream_t sx = st[r];
    CHECK("set device", gmt_set_device(dev[r]));
    CHECK("xsum", gmt_sum_f64(n * W, allx[r], d_sum[r], scratch[r], sx));
    CHECK("D2H xsum", gmt_memcpy_async(&xs, d_sum[r], 8, 2, sx));
    CHECK("sync", gmt_stream_synchronize(sx));
    if (std::fabs(xs - W * (n + 1) / 2.0) > 1e-6 * n * W) {
      std::fprintf(stderr, "%d/%d gathered x sums to %f, expected %f\n", r, W, xs, W * (n + 1) / 2.0);
      bad = 1;
    }
  }
  for (int r = 0; r < W; ++r) {
    CHECK("set device", gmt_set_device(dev[r]));
    gmt_free(GMT_SPACE_DEVICE, scratch[r]);
    gmt_free(GMT_SPACE_DEVICE, d_sum[r]);
    gmt_free(GMT_SPACE_DEVICE, d_y[r]);
  }
  gmt_coll_engine_destroy(ey);
  gmt_coll_engine_destroy(ex);
  const double total = now_s() - t_start;
  for (int r = 0; r < W; ++r) {
    std::printf("%d/%d TIME total  : %0.3f\n", r, W, total);
    std::printf("%d/%d TIME kernel : %0.6f\n", r, W, k_ms * 1e-3);
    std::printf("%d/%d TIME barrier: %0.6f\n", r, W, b_ms * 1e-3);
    std::printf("%d/%d TIME gather : %0.6f\n", r, W, g_ms * 1e-3);
  }
  const double gbytes = 2.0 * double(n) * 8.0 * (W - 1) * W / 1e9;   // bytes every rank receives, both gathers
  if (W > 1 && g_ms > 0) std::printf("gather bus traffic %0.3f GB in %0.3f ms: %0.1f GB/s aggregate\n", gbytes, g_ms, gbytes / (g_ms * 1e-3));
  return bad ? 3 : 0;
}
