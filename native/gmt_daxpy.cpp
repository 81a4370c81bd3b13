// gmt_daxpy -- native single-GPU smoke test, the job of the reference's daxpy / daxpy_nvtx
// (daxpy.cu:35-94, daxpy_nvtx.cu:38-108): y = 2*x + y with x[i] = i+1, y[i] = -i-1, print every
// y[i] and "SUM = %f" (524800.000000 for the default n = 1024).  The sum comes out of the same
// kernel pass as the axpy (gmt_daxpy_sum_f64) instead of a host loop; the host copy of y is only
// needed for the per-element lines.  NVTX ranges and the profiler window as in daxpy_nvtx.
//
//   gmt_daxpy [n] [--quiet]        --quiet: do not print the n element lines
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "gmt_b200.h"

#define CHECK(what, call)                                                                  \
  do {                                                                                     \
    int rc_ = (call);                                                                      \
    if (rc_ != 0) {                                                                        \
      std::fprintf(stderr, "%s(%i): GMT Error (%s) %i: %s\n", __FILE__, __LINE__, what, rc_, gmt_last_error()); \
      return 2;                                                                            \
    }                                                                                      \
  } while (0)

int main(int argc, char** argv)
{
  long long n = 1024;
  bool quiet = false;
  for (int i = 1; i < argc; ++i) {
    if (!std::strcmp(argv[i], "--quiet")) quiet = true;
    else n = std::atoll(argv[i]);
  }
  if (n <= 0) {
    std::fprintf(stderr, "n must be positive\n");
    return 1;
  }
  const double a = 2.0;
  const size_t bytes = size_t(n) * sizeof(double);
  double *x = nullptr, *y = nullptr, *d_x = nullptr, *d_y = nullptr, *d_sum = nullptr, *h_sum = nullptr;
  void* scratch = nullptr;
  CHECK("host x", gmt_alloc(GMT_SPACE_PINNED, bytes, (void**)&x));
  CHECK("host y", gmt_alloc(GMT_SPACE_PINNED, bytes, (void**)&y));
  CHECK("host sum", gmt_alloc(GMT_SPACE_PINNED, sizeof(double), (void**)&h_sum));
  for (long long i = 0; i < n; ++i) {
    x[i] = double(i + 1);
    y[i] = double(-i - 1);
  }
  CHECK("d_x", gmt_alloc(GMT_SPACE_DEVICE, bytes, (void**)&d_x));
  CHECK("d_y", gmt_alloc(GMT_SPACE_DEVICE, bytes, (void**)&d_y));
  CHECK("d_sum", gmt_alloc(GMT_SPACE_DEVICE, sizeof(double), (void**)&d_sum));
  CHECK("scratch", gmt_alloc(GMT_SPACE_DEVICE, gmt_reduce_scratch_bytes(), &scratch));
  CHECK("scratch zero", gmt_memset_async(scratch, 0, gmt_reduce_scratch_bytes(), nullptr));

  CHECK("profiler start", gmt_profiler_start());
  gmt_nvtx_push("copyInput");
  CHECK("H2D x", gmt_memcpy_async(d_x, x, bytes, 1, nullptr));
  CHECK("H2D y", gmt_memcpy_async(d_y, y, bytes, 1, nullptr));
  gmt_nvtx_pop();

  gmt_nvtx_push("daxpy");
  CHECK("daxpy", gmt_daxpy_sum_f64(n, a, d_x, d_y, d_sum, scratch, nullptr));
  CHECK("sync", gmt_stream_synchronize(nullptr));
  gmt_nvtx_pop();

  gmt_nvtx_push("copyOutput");
  CHECK("D2H y", gmt_memcpy_async(y, d_y, bytes, 2, nullptr));
  CHECK("D2H sum", gmt_memcpy_async(h_sum, d_sum, sizeof(double), 2, nullptr));
  CHECK("sync", gmt_stream_synchronize(nullptr));
  gmt_nvtx_pop();
  CHECK("profiler stop", gmt_profiler_stop());

  if (!quiet)
    for (long long i = 0; i < n; ++i) std::printf("%f\n", y[i]);
  std::printf("SUM = %f\n", *h_sum);

  gmt_free(GMT_SPACE_DEVICE, scratch);
  gmt_free(GMT_SPACE_DEVICE, d_sum);
  gmt_free(GMT_SPACE_DEVICE, d_y);
  gmt_free(GMT_SPACE_DEVICE, d_x);
  gmt_free(GMT_SPACE_PINNED, h_sum);
  gmt_free(GMT_SPACE_PINNED, y);
  gmt_free(GMT_SPACE_PINNED, x);
  return 0;
}
