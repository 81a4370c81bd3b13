// gmt_gatherinplace -- native counterpart of mpigatherinplace (mpigatherinplace.f90:1-58):
// in-place allgather of N doubles per rank (default 128 Mi = 1 GiB per rank), slot values
// (rank*i)/N in integer arithmetic, prints per rank:  rank / nmpi  lsum  asum.
// The reference keeps the array in HOST memory and has no timer; here it lives in device memory
// reached through CUDA peer access, the sums are fp64 device reductions (the reference's fp32
// accumulators saturate at 2^24), and the gather is timed on the device.
//
//   gmt_gatherinplace [N] [--ranks R] [--gather ROOT]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
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

// sum_{i=1..N} floor(r*i/N) = ((r-1)(N-1) + gcd(r,N) - 1)/2 + r   (r >= 1)
static double slot_sum(long long r, long long N)
{
  if (r == 0) return 0.0;
  return double(((r - 1) * (N - 1) + std::gcd(r, N) - 1) / 2 + r);
}

int main(int argc, char** argv)
{
  long long N = 128LL << 20;
  int ranks = 0, root = -1;
  for (int i = 1; i < argc; ++i) {
    if (!std::strcmp(argv[i], "--ranks") && i + 1 < argc) ranks = std::atoi(argv[++i]);
    else if (!std::strcmp(argv[i], "--gather") && i + 1 < argc) root = std::atoi(argv[++i]);
    else if (argv[i][0] != '-') N = std::atoll(argv[i]);
    else {
      std::fprintf(stderr, "usage: gmt_gatherinplace [N] [--ranks R] [--gather ROOT]\n");
      return 1;
    }
  }
  int n_dev = 0;
  CHECK("device count", gmt_device_count(&n_dev));
  if (n_dev == 0) {
    std::fprintf(stderr, "gmt_gatherinplace: no CUDA device\n");
    return 2;
  }
  if (ranks <= 0) ranks = n_dev;
  const int W = ranks;
  int probe = 0;
  if (gmt_engine_default_device(W, 0, n_dev, &probe) != 0) {
    std::fprintf(stderr, "%s\n", gmt_last_error());
    return 1;
  }
  gmt_coll_engine* eng = nullptr;
  CHECK("engine", gmt_coll_engine_create(W, nullptr, N * W, 8, 0, &eng));
  std::vector<double*> allx(W), d_s(W);
  std::vector<void*> scratch(W);
  std::vector<gmt_stream_t> st(W);
  std::vector<int> dev(W);
  for (int r = 0; r < W; ++r) {
    CHECK("buffer", gmt_coll_engine_buffer(eng, r, &allx[r], nullptr));
    CHECK("stream", gmt_coll_engine_stream(eng, r, &st[r]));
    CHECK("device", gmt_coll_engine_rank_device(eng, r, &dev[r]));
    CHECK("set device", gmt_set_device(dev[r]));
    CHECK("d_s", gmt_alloc(GMT_SPACE_DEVICE, 64, (void**)&d_s[r]));
    CHECK("scratch", gmt_alloc(GMT_SPACE_DEVICE, gmt_reduce_scratch_bytes(), &scratch[r]));
    CHECK("scratch zero", gmt_memset_async(scratch[r], 0, gmt_reduce_scratch_bytes(), st[r]));
    CHECK("zero", gmt_memset_async(allx[r], 0, size_t(N) * W * 8, st[r]));
    CHECK("fill", gmt_slot_pattern_f64(N, r, N, allx[r] + (long long)r * N, st[r]));
  }
  CHECK("sync", gmt_coll_engine_synchronize(eng));
  double ms = 0.0;
  CHECK("mark", gmt_coll_engine_mark(eng, 0));
  if (root < 0) CHECK("allgather", gmt_coll_engine_allgather(eng, N, 0));
  else CHECK("gather", gmt_coll_engine_gather(eng, N, root));
  CHECK("mark", gmt_coll_engine_mark(eng, 1));
  CHECK("elapsed", gmt_coll_engine_elapsed_ms(eng, &ms));
  CHECK("status", gmt_coll_engine_status(eng, nullptr));

  double expect_all = 0.0;
  for (int r = 0; r < W; ++r) expect_all += slot_sum(r, N);
  int bad = 0;
  for (int r = 0; r < W; ++r) {
    double lsum = 0.0, asum = 0.0;
    CHECK("set device", gmt_set_device(dev[r]));
    CHECK("lsum", gmt_sum_f64(N, allx[r] + (long long)r * N, d_s[r], scratch[r], st[r]));
    CHECK("D2H", gmt_memcpy_async(&lsum, d_s[r], 8, 2, st[r]));
    CHECK("sync", gmt_stream_synchronize(st[r]));
    CHECK("asum", gmt_sum_f64(N * W, allx[r], d_s[r], scratch[r], st[r]));
    CHECK("D2H", gmt_memcpy_async(&asum, d_s[r], 8, 2, st[r]));
    CHECK("sync", gmt_stream_synchronize(st[r]));
    std::printf(" %d / %d   %.1f   %.1f\n", r, W, lsum, asum);
    const double want = (root < 0 || r == root) ? expect_all : slot_sum(r, N);
    if (lsum != slot_sum(r, N) || asum != want) {
      std::fprintf(stderr, "rank %d: expected lsum %.1f asum %.1f\n", r, slot_sum(r, N), want);
      bad = 1;
    }
  }
  const double gb = double(N) * 8.0 * (root < 0 ? double(W) * (W - 1) : double(W - 1)) / 1e9;
  if (W > 1) std::printf("%s of %lld doubles per rank: %0.3f ms, %0.1f GB/s aggregate\n",
                         root < 0 ? "allgather" : "gather", N, ms, gb / (ms * 1e-3));
  for (int r = 0; r < W; ++r) {
    CHECK("set device", gmt_set_device(dev[r]));
    gmt_free(GMT_SPACE_DEVICE, scratch[r]);
    gmt_free(GMT_SPACE_DEVICE, d_s[r]);
  }
  gmt_coll_engine_destroy(eng);
  return bad ? 3 : 0;
}
