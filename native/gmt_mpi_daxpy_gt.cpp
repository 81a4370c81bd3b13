// gmt_mpi_daxpy_gt -- native counterpart of mpi_daxpy_gt (mpi_daxpy_gt.cc:48-97): the n = 1024 daxpy
// of mpi_daxpy on plain device arrays (host arrays, explicit copies in and out, device axpy, host sum),
// printed with the device id and the packed PCI id of the GPU each rank landed on:
//   "r/N [dev:0x%08x] SUM = %f"  (524800.000000)
// One process plays all ranks; ranks are placed with the reference's block rule (several per GPU
// allowed when R is a multiple of the GPU count).
//
//   gmt_mpi_daxpy_gt [--ranks R]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "gmt_check.hpp"

int main(int argc, char** argv)
{
  int ranks = 0;
  for (int i = 1; i < argc; ++i) {
    if (!std::strcmp(argv[i], "--ranks") && i + 1 < argc) ranks = std::atoi(argv[++i]);
    else {
      std::fprintf(stderr, "usage: gmt_mpi_daxpy_gt [--ranks R]\n");
      return 1;
    }
  }
  int n_dev = 0;
  GMT_CHECK("device count", gmt_device_count(&n_dev));
  if (n_dev == 0) {
    std::fprintf(stderr, "gmt_mpi_daxpy_gt: no CUDA device\n");
    return 2;
  }
  if (ranks <= 0) ranks = n_dev;
  const int n = 1024;
  const double a = 2.0;
  for (int r = 0; r < ranks; ++r) {
    int dev = 0;
    if (gmt_engine_default_device(ranks, r, n_dev, &dev) != 0) {   // the reference's ratio check
      std::printf("%s\n", gmt_last_error());
      return 1;
    }
    gmt_device_props props;
    GMT_CHECK("set device", gmt_set_device(dev));
    GMT_CHECK("device props", gmt_device_props_get(dev, &props));
    const unsigned vendor_id = ((unsigned)(props.pci_domain & 0xFFFF) << 16) | ((unsigned)(props.pci_bus & 0xFF) << 8) |
                               (unsigned)(props.pci_device & 0xFF);
    std::vector<double> h_x(n), h_y(n);
    for (int i = 0; i < n; ++i) {
      h_x[i] = i + 1;
      h_y[i] = -i - 1;
    }
    double *d_x = nullptr, *d_y = nullptr;
    GMT_CHECK("d_x", gmt_alloc(GMT_SPACE_DEVICE, n * sizeof(double), (void**)&d_x));
    GMT_CHECK("d_y", gmt_alloc(GMT_SPACE_DEVICE, n * sizeof(double), (void**)&d_y));
    GMT_CHECK("copy x", gmt_memcpy_async(d_x, h_x.data(), n * sizeof(double), 1, nullptr));
    GMT_CHECK("copy y", gmt_memcpy_async(d_y, h_y.data(), n * sizeof(double), 1, nullptr));
    GMT_CHECK("axpy", gmt_daxpy_f64(n, a, d_x, d_y, nullptr));
    GMT_CHECK("sync", gmt_device_synchronize());
    GMT_CHECK("copy back", gmt_memcpy_async(h_y.data(), d_y, n * sizeof(double), 2, nullptr));
    GMT_CHECK("sync", gmt_device_synchronize());
    double sum = 0.0;
    for (int i = 0; i < n; ++i) sum += h_y[i];
    std::printf("%d/%d [%d:0x%08x] SUM = %f\n", r, ranks, dev, vendor_id, sum);
    GMT_CHECK("free", gmt_free(GMT_SPACE_DEVICE, d_y));
    GMT_CHECK("free", gmt_free(GMT_SPACE_DEVICE, d_x));
    if (sum != 524800.0) return 3;
  }
  return 0;
}
