// gmt_stencil2d -- native (no Python, no MPI) driver of the fused halo+stencil engine.
//
// One process owns every GPU of the node; "ranks" are tiles of a process grid (several per GPU
// when --ranks exceeds the GPU count, the reference's oversubscription rule).  Same job as the
// reference's flagship mpi_stencil2d_gt (mpi_stencil2d_gt.cc:651-734): run the dim-0 and dim-1
// slab tests -- and the 4-neighbour grid test the reference lacks -- for n_iter iterations after
// 5 warm-up iterations and print one result line per test.
//
//   gmt_stencil2d [n_local_deriv] [n_iter] [--ranks N] [--grid PXxPY] [--other N]
//                 [--tests dim0,dim1,grid] [--tile N]
//
// positional arguments as in the reference (defaults 1024 and 1000); --other is the extent of the
// non-derivative dimension (default 512*1024, mpi_stencil2d_gt.cc:676); --tile the per-rank edge of
// the grid test (default 16384).  Result lines:
//   TEST dim:<d>, device , fused; <seconds for n_iter steps, max over GPUs>, err=<sum of err_norm>
// (the reference prints the SUM over ranks of host-clock totals, :563-570; device time is used here).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "gmt_b200.h"

static int die(const char* what, int rc)
{
  std::fprintf(stderr, "%s: error %d: %s\n", what, rc, gmt_last_error());
  return 2;
}

struct Case {
  const char* label;
  int px, py, gx, gy;
  long long nx, ny;
};

static int run_case(const Case& c, int n_iter, int n_warmup)
{
  gmt_engine2d* eng = nullptr;
  int rc = gmt_engine2d_create(c.px, c.py, nullptr, c.nx, c.ny, c.gx, c.gy, 1, 0, &eng);
  if (rc) return die("gmt_engine2d_create", rc);
  if ((rc = gmt_engine2d_init_analytic(eng, 1e30)) != 0) return die("gmt_engine2d_init_analytic", rc);
  double ms = 0.0;
  if ((rc = gmt_engine2d_time_steps(eng, n_iter, n_warmup, &ms)) != 0) return die("gmt_engine2d_time_steps", rc);
  const int world = gmt_engine2d_world(eng);
  double err_sum = 0.0;
  for (int r = 0; r < world; ++r) {
    double err = 0.0;
    if ((rc = gmt_engine2d_err_norm(eng, r, &err)) != 0) return die("gmt_engine2d_err_norm", rc);
    err_sum += err;
  }
  // the reference's own timed quantity: the exchange alone (mpi_stencil2d_gt.cc:512-522)
  double ex_ms = 0.0;
  if ((rc = gmt_engine2d_time_exchanges(eng, n_iter < 20 ? n_iter : 20, 3, &ex_ms)) != 0)
    return die("gmt_engine2d_time_exchanges", rc);
  const double glups = double(c.nx) * double(c.ny) * world / (ms * 1e-3) / 1e9;
  std::printf("TEST %s, device , fused; %0.8f, err=%0.8f\n", c.label, ms * 1e-3 * n_iter, err_sum);
  std::printf("     %d ranks on %d GPU(s), %lld x %lld per rank: %0.4f ms/step, %0.1f GLUP/s; exchange alone %0.2f us\n",
              world, gmt_engine2d_device_groups(eng), c.nx, c.ny, ms, glups, ex_ms * 1e3);
  gmt_engine2d_destroy(eng);
  return 0;
}

int main(int argc, char** argv)
{
  long long n_local = 1024, other = 512 * 1024, tile = 16384;
  int n_iter = 1000, n_warmup = 5, ranks = 0, px = 0, py = 0, npos = 0;
  std::string tests = "dim0,dim1,grid";
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto val = [&](const char* name) -> const char* {
      if (i + 1 >= argc) {
        std::fprintf(stderr, "%s needs a value\n", name);
        std::exit(1);
      }
      return argv[++i];
    };
    if (a == "--ranks") ranks = std::atoi(val("--ranks"));
    else if (a == "--other") other = std::atoll(val("--other"));
    else if (a == "--tile") tile = std::atoll(val("--tile"));
    else if (a == "--tests") tests = val("--tests");
    else if (a == "--grid") {
      if (std::sscanf(val("--grid"), "%dx%d", &px, &py) != 2) {
        std::fprintf(stderr, "--grid wants PXxPY\n");
        return 1;
      }
    } else if (a == "-h" || a == "--help") {
      std::printf("usage: gmt_stencil2d [n_local_deriv] [n_iter] [--ranks N] [--grid PXxPY] [--other N] "
                  "[--tests dim0,dim1,grid] [--tile N]\n");
      return 0;
    } else if (npos == 0) { n_local = std::atoll(a.c_str()); ++npos; }
    else if (npos == 1) { n_iter = std::atoi(a.c_str()); ++npos; }
    else {
      std::fprintf(stderr, "unexpected argument %s\n", a.c_str());
      return 1;
    }
  }
  int n_dev = 0;
  int rc = gmt_device_count(&n_dev);
  if (rc) return die("gmt_device_count", rc);
  if (n_dev == 0) {
    std::fprintf(stderr, "gmt_stencil2d: no CUDA device\n");
    return 2;
  }
  if (ranks <= 0) ranks = px && py ? px * py : n_dev;
  if (!px || !py) {  // most square factorisation, px >= py
    py = 1;
    for (int f = 1; f * f <= ranks; ++f)
      if (ranks % f == 0) py = f;
    px = ranks / py;
  }
  if (px * py != ranks) {
    std::fprintf(stderr, "grid %dx%d does not match %d ranks\n", px, py, ranks);
    return 1;
  }
  int dev0 = 0;
  if ((rc = gmt_engine_default_device(ranks, 0, n_dev, &dev0)) != 0) {   // the reference's divisibility check
    std::fprintf(stderr, "%s\n", gmt_last_error());
    return 1;
  }
  std::printf("n procs        = %d\n", ranks);
  std::printf("n devices      = %d\n", n_dev);
  std::printf("n_global_deriv = %lld\n", n_local * ranks);
  std::printf("n_global_other = %lld\n", other);
  std::printf("n_iter         = %d\n", n_iter);
  std::printf("n_warmup       = %d\n", n_warmup);

  std::vector<Case> cases;
  if (tests.find("dim0") != std::string::npos) cases.push_back({"dim:0", ranks, 1, 2, 0, n_local, other});
  if (tests.find("dim1") != std::string::npos) cases.push_back({"dim:1", 1, ranks, 0, 2, other, n_local});
  if (tests.find("grid") != std::string::npos) cases.push_back({"grid ", px, py, 2, 2, tile, tile});
  for (const Case& c : cases)
    if ((rc = run_case(c, n_iter, n_warmup)) != 0) return rc;
  return 0;
}
