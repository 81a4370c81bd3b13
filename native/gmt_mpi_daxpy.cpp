// gmt_mpi_daxpy -- native counterpart of mpi_daxpy (mpi_daxpy.cc:65-169): every "rank" runs the
// n = 1024 daxpy on MANAGED memory and the host sums the result through unified memory, to show
// that ranks, GPUs (several ranks per GPU allowed) and unified memory coexist.  One process plays
// all ranks.  Printed lines as in the reference: the rank->device line with the per-rank memory
// share, MEMORY_PER_CORE (rank 0), the preferred-location report for device / managed / host
// pointers and "r/N SUM = %f" (524800.000000).
//
//   gmt_mpi_daxpy [--ranks R]
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "gmt_check.hpp"

int main(int argc, char** argv)
{
  int ranks = 0;
  for (int i = 1; i < argc; ++i) {
    if (!std::strcmp(argv[i], "--ranks") && i + 1 < argc) ranks = std::atoi(argv[++i]);
    else {
      std::fprintf(stderr, "usage: gmt_mpi_daxpy [--ranks R]\n");
      return 1;
    }
  }
  int n_dev = 0;
  GMT_CHECK("device count", gmt_device_count(&n_dev));
  if (n_dev == 0) {
    std::fprintf(stderr, "gmt_mpi_daxpy: no CUDA device\n");
    return 2;
  }
  if (ranks <= 0) ranks = n_dev;
  const int n = 1024;
  const double a = 2.0;
  const char* mpc = std::getenv("MEMORY_PER_CORE");
  if (mpc) std::printf("MEMORY_PER_CORE=%s\n", mpc);
  else std::printf("MEMORY_PER_CORE is not set\n");
  for (int r = 0; r < ranks; ++r) {
    int dev = 0;
    if (gmt_engine_default_device(ranks, r, n_dev, &dev) != 0) {   // the reference's ratio check
      std::printf("%s\n", gmt_last_error());
      return 1;
    }
    const int per_dev = ranks > n_dev ? ranks / n_dev : 1;
    gmt_device_props props;
    GMT_CHECK("set device", gmt_set_device(dev));
    GMT_CHECK("device props", gmt_device_props_get(dev, &props));
    std::printf("RANK[%d/%d] => DEVICE[%d/%d] mem=%zd\n", r + 1, ranks, dev + 1, n_dev,
                (size_t)(props.total_global_mem / per_dev));

    double *x = (double*)std::malloc(n * sizeof(double)), *y = (double*)std::malloc(n * sizeof(double));
    double *d_x = nullptr, *d_y = nullptr, *m_x = nullptr, *m_y = nullptr;
    GMT_CHECK("d_x", gmt_alloc(GMT_SPACE_DEVICE, n * sizeof(double), (void**)&d_x));
    GMT_CHECK("d_y", gmt_alloc(GMT_SPACE_DEVICE, n * sizeof(double), (void**)&d_y));
    GMT_CHECK("m_x", gmt_alloc(GMT_SPACE_MANAGED, n * sizeof(double), (void**)&m_x));
    GMT_CHECK("m_y", gmt_alloc(GMT_SPACE_MANAGED, n * sizeof(double), (void**)&m_y));
    for (int i = 0; i < n; ++i) {   // the host writes the managed arrays directly
      x[i] = m_x[i] = i + 1;
      y[i] = m_y[i] = -i - 1;
    }
    GMT_MEMINFO(d_x, n * sizeof(double));
    GMT_MEMINFO(d_y, n * sizeof(double));
    GMT_MEMINFO(m_x, n * sizeof(double));
    GMT_MEMINFO(m_y, n * sizeof(double));
    GMT_MEMINFO(x, n * sizeof(double));
    GMT_MEMINFO(y, n * sizeof(double));
    if (r == 0) {
      GMT_PTRINFO(d_x);
      GMT_PTRINFO(m_x);
      GMT_PTRINFO(x);
    }
    GMT_CHECK("daxpy on managed pointers", gmt_daxpy_f64(n, a, m_x, m_y, nullptr));
    GMT_CHECK("sync", gmt_device_synchronize());
    double sum = 0.0;
    for (int i = 0; i < n; ++i) sum += m_y[i];   // read back through unified memory
    std::printf("%d/%d SUM = %f\n", r, ranks, sum);
    GMT_CHECK("free", gmt_free(GMT_SPACE_MANAGED, m_y));
    GMT_CHECK("free", gmt_free(GMT_SPACE_MANAGED, m_x));
    GMT_CHECK("free", gmt_free(GMT_SPACE_DEVICE, d_y));
    GMT_CHECK("free", gmt_free(GMT_SPACE_DEVICE, d_x));
    std::free(y);
    std::free(x);
    if (sum != 524800.0) return 3;
  }
  return 0;
}
