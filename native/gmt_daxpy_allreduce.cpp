// gmt_daxpy_allreduce -- the fused form of the reference's daxpy benchmark: y <- a*x + y on every
// rank and the GLOBAL sum of the new y on every rank, in one kernel per GPU (what
// mpi_daxpy_nvtx does with cublasDaxpy + MPI_Allgather + host sums, mpi_daxpy_nvtx.cc:245-310).
// One process, all GPUs; written against the C++ conveniences of include/gmt_b200.hpp.
//
//   gmt_daxpy_allreduce [--n N] [--ranks R] [--iters K]      (default n = 2^20 per rank)
//
// Prints per rank "r/N ALLSUM = %f" (expected ranks*(n+1)/2) and the device time per call.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "gmt_b200.hpp"

int main(int argc, char** argv)
try {
  long long n = 1 << 20;
  int ranks = 0, iters = 50;
  for (int i = 1; i < argc; ++i) {
    if (!std::strcmp(argv[i], "--n") && i + 1 < argc) n = std::atoll(argv[++i]);
    else if (!std::strcmp(argv[i], "--ranks") && i + 1 < argc) ranks = std::atoi(argv[++i]);
    else if (!std::strcmp(argv[i], "--iters") && i + 1 < argc) iters = std::atoi(argv[++i]);
    else {
      std::fprintf(stderr, "usage: gmt_daxpy_allreduce [--n N] [--ranks R] [--iters K]\n");
      return 1;
    }
  }
  int n_dev = 0;
  gmt::check(gmt_device_count(&n_dev), "gmt_device_count");
  if (n_dev == 0) {
    std::fprintf(stderr, "gmt_daxpy_allreduce: no CUDA device\n");
    return 2;
  }
  if (ranks <= 0) ranks = n_dev;
  int probe = 0;
  if (gmt_engine_default_device(ranks, 0, n_dev, &probe) != 0) {
    std::fprintf(stderr, "%s\n", gmt_last_error());
    return 1;
  }
  gmt::ProfilerWindow window;
  gmt::CollEngine eng(ranks, /*buf_doubles=*/8, /*stage_cap=*/64);
  std::vector<gmt::DeviceBuffer<double>> x(ranks), y(ranks), sum(ranks);
  {
    gmt::NvtxRange r("initializeArrays");
    for (int k = 0; k < ranks; ++k) {
      gmt::check(gmt_set_device(eng.rank_device(k)), "gmt_set_device");
      x[k] = gmt::DeviceBuffer<double>(n);
      y[k] = gmt::DeviceBuffer<double>(n);
      sum[k] = gmt::DeviceBuffer<double>(8);
      eng.bind(k, x[k].get(), sum[k].get(), y[k].get());
    }
  }
  auto reset = [&] {   // x[i] = (i+1)/n, y = -x  =>  after y <- 2x + y:  y == x, sum = (n+1)/2 per rank
    for (int k = 0; k < ranks; ++k) {
      gmt::check(gmt_set_device(eng.rank_device(k)), "gmt_set_device");
      gmt::check(gmt_iota_f64(n, 1.0, 1.0, 1.0 / double(n), x[k].get(), eng.stream(k)), "iota x");
      gmt::check(gmt_iota_f64(n, 1.0, 1.0, -1.0 / double(n), y[k].get(), eng.stream(k)), "iota y");
    }
  };
  reset();
  eng.daxpy_allreduce(n, 2.0);   // checked call
  eng.synchronize();
  eng.status();
  int bad = 0;
  const double want = ranks * (n + 1) / 2.0;
  for (int k = 0; k < ranks; ++k) {
    double got = 0.0;
    gmt::check(gmt_set_device(eng.rank_device(k)), "gmt_set_device");
    gmt::check(gmt_memcpy_async(&got, sum[k].get(), 8, 2, eng.stream(k)), "D2H");
    gmt::check(gmt_stream_synchronize(eng.stream(k)), "sync");
    std::printf("%d/%d ALLSUM = %f\n", k, ranks, got);
    if (std::fabs(got - want) > 1e-9 * want) bad = 1;
  }
  double ms = 0.0;
  {
    gmt::NvtxRange r("daxpyAllreduce");
    for (int i = 0; i < 5; ++i) eng.daxpy_allreduce(n, 0.0);   // a = 0 keeps y bounded over many calls
    eng.synchronize();
    ms = eng.timed([&] {
      for (int i = 0; i < iters; ++i) eng.daxpy_allreduce(n, 0.0);
    });
  }
  eng.status();
  const double us = ms * 1e3 / iters;
  std::printf("daxpy + global sum, %lld doubles per rank on %d rank(s): %0.2f us per call, %0.1f GB/s per GPU\n", n,
              ranks, us, 24.0 * n / (us * 1e-6) / 1e9);
  return bad ? 3 : 0;
} catch (const gmt::Error& e) {
  std::fprintf(stderr, "gmt_daxpy_allreduce: %s\n", e.what());
  return 2;
}
