// gmt_stencil2d_sycl -- native counterpart of the two SYCL programs, mpi_stencil2d_sycl
// (mpi_stencil2d_sycl.cc:377-556) and, with --oo, mpi_stencil2d_sycl_oo (mpi_stencil2d_sycl_oo.cc:517-705):
// a 2-D array slab-decomposed along dim 0 (the strided ghost strips), n_warmup = 5 untimed and n_iter
// timed boundary exchanges each followed by the dim-0 stencil, then the error norm of the derivative.
// One process drives every GPU through the exchange engine (csrc/engine.cu); the timed quantity is what
// the reference times -- the exchange ALONE (gmt_halo_exchange2d_f64: pack-free export into the
// neighbours' inboxes over NVLink, flags, landing), device time, mean per iteration, max over GPUs.
//
//   gmt_stencil2d_sycl [nx_local] [stage_host] [n_iter] [--ranks R] [--ny N]          (default form)
//   gmt_stencil2d_sycl --oo [n_global_Ki] [stage_host] [n_iter] [--ranks R]           (square domain)
//
// Positional arguments and printed lines as in the reference (stage_host is accepted and printed;
// there are no staging buffers to put on the host).  Lines: "<rank> <BDF>(<uuid>)" per rank, the
// parameter block (rank 0), "dev bytes  = ...", then per rank the exchange time and err_norm lines in
// each program's own format.
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

int main(int argc, char** argv)
{
  bool oo = false;
  long long a0 = -1, ny_opt = 512 * 1024;
  int ranks = 0, n_iter = 100, n_warmup = 5, npos = 0;
  bool stage_host = false;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    if (a == "--oo") oo = true;
    else if (a == "--ranks" && i + 1 < argc) ranks = std::atoi(argv[++i]);
    else if (a == "--ny" && i + 1 < argc) ny_opt = std::atoll(argv[++i]);
    else if (a == "-h" || a == "--help") {
      std::printf("usage: gmt_stencil2d_sycl [--oo] [nx_local | n_global_Ki] [stage_host] [n_iter] [--ranks R] [--ny N]\n");
      return 0;
    } else if (npos == 0) { a0 = std::atoll(a.c_str()); ++npos; }
    else if (npos == 1) { stage_host = a[0] == '1'; ++npos; }   // first character '1', like the reference
    else if (npos == 2) { n_iter = std::atoi(a.c_str()); ++npos; }
    else {
      std::fprintf(stderr, "unexpected argument %s\n", a.c_str());
      return 1;
    }
  }
  int n_dev = 0, rc = gmt_device_count(&n_dev);
  if (rc) return die("gmt_device_count", rc);
  if (n_dev == 0) {
    std::fprintf(stderr, "gmt_stencil2d_sycl: no CUDA device\n");
    return 2;
  }
  if (ranks <= 0) ranks = n_dev;
  int dev0 = 0;
  if (gmt_engine_default_device(ranks, 0, n_dev, &dev0) != 0) {   // the reference's ratio check
    std::printf("%s\n", gmt_last_error());
    return 1;
  }
  long long nx_local, ny, n_global = 0;
  if (oo) {
    n_global = (a0 > 0 ? a0 : 8) * 1024;            // mpi_stencil2d_sycl_oo.cc:528-531
    if (n_global % ranks != 0) {
      std::printf("%d nmpi (%d) must be divisor of domain size (%lld), exiting\n", 0, ranks, n_global);
      return 1;
    }
    nx_local = n_global / ranks;
    ny = n_global;
  } else {
    nx_local = a0 > 0 ? a0 : 1024;                  // mpi_stencil2d_sycl.cc:384-399
    ny = ny_opt;
  }

  gmt_engine2d* eng = nullptr;
  if ((rc = gmt_engine2d_create(ranks, 1, nullptr, nx_local, ny, 2, 0, 1, 0, &eng)) != 0) return die("gmt_engine2d_create", rc);
  std::vector<unsigned> vendor(ranks);
  for (int r = 0; r < ranks; ++r) {
    int dev = 0;
    gmt_device_props p;
    if ((rc = gmt_engine2d_rank_device(eng, r, &dev)) != 0) return die("gmt_engine2d_rank_device", rc);
    if ((rc = gmt_device_props_get(dev, &p)) != 0) return die("gmt_device_props_get", rc);
    vendor[r] = ((unsigned)(p.pci_domain & 0xFFFF) << 16) | ((unsigned)(p.pci_bus & 0xFF) << 8) | (unsigned)(p.pci_device & 0xFF);
    if (!oo) {
      std::printf("%d %04x:%02x:%02x.0(", r, p.pci_domain, p.pci_bus, p.pci_device);
      for (int b = 0; b < 8; ++b) std::printf("%02x", p.uuid[b]);
      std::printf(")\n");
    }
  }
  std::printf("n procs    = %d\n", ranks);
  std::printf("rank       = %d\n", 0);
  if (oo) {
    std::printf("n_global   = %lld\n", n_global);
    std::printf("n_local    = %lld\n", nx_local);
  } else {
    std::printf("ny         = %lld\n", ny);
    std::printf("nx_global  = %lld\n", nx_local * ranks);
    std::printf("nx_local   = %lld\n", nx_local);
  }
  std::printf("n_iter     = %d\n", n_iter);
  std::printf("n_warmup   = %d\n", n_warmup);
  std::printf("stage_host = %d\n", stage_host ? 1 : 0);
  if (!oo) {
    const double nbytes = 8.0 * ((nx_local + 4) * (double)ny + nx_local * (double)ny + 4.0 * 2 * ny);
    if (nbytes / (1024.0 * 1024.0 * 1024.0) >= 1.0) std::printf("dev bytes  = %0.3f GB\n", nbytes / (1024.0 * 1024.0 * 1024.0));
    else std::printf("dev bytes  = %0.3f MB\n", nbytes / (1024.0 * 1024.0));
  }

  if ((rc = gmt_engine2d_init_analytic(eng, 1e30)) != 0) return die("gmt_engine2d_init_analytic", rc);
  double ex_ms = 0.0, step_ms = 0.0;
  if ((rc = gmt_engine2d_time_exchanges(eng, n_iter, n_warmup, &ex_ms)) != 0) return die("gmt_engine2d_time_exchanges", rc);
  // the product's step (exchange fused into the stencil kernel); also produces the derivative and its error norm
  if ((rc = gmt_engine2d_time_steps(eng, n_iter < 20 ? n_iter : 20, 2, &step_ms)) != 0) return die("gmt_engine2d_time_steps", rc);
  for (int r = 0; r < ranks; ++r) {
    double err = 0.0;
    if ((rc = gmt_engine2d_err_norm(eng, r, &err)) != 0) return die("gmt_engine2d_err_norm", rc);
    if (oo) {
      std::printf("%d: exchange time %0.8f ms\n", r, ex_ms);
      std::printf("%d: [0x%08x] err_norm = %.8f\n", r, vendor[r], err);
    } else {
      int dev = 0;
      gmt_engine2d_rank_device(eng, r, &dev);
      std::printf("%d/%d exchange time %0.8f ms\n", r, ranks, ex_ms);
      std::printf("%d/%d [%d:0x%08x] err_norm = %.8f\n", r, ranks, dev, vendor[r], err);
    }
  }
  std::printf("     fused exchange+stencil step: %0.4f ms, %0.1f GLUP/s on %d GPU(s)\n", step_ms,
              double(nx_local) * double(ny) * ranks / (step_ms * 1e-3) / 1e9, gmt_engine2d_device_groups(eng));
  gmt_engine2d_destroy(eng);
  return 0;
}
