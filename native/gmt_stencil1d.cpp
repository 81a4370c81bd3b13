// gmt_stencil1d -- native counterpart of mpi_stencil_gt (mpi_stencil_gt.cc:124-230): the 1-D
// slab-decomposed stencil, one process driving every GPU of the node.
//
// This program wires the exchange BY HAND against the C ABI (no engine object): it allocates each
// rank's ghosted array, the four 32-byte inboxes and two flag words, fills one gmt_halo1d_rank
// per rank with the neighbours' addresses (same process, CUDA peer access) and launches
// gmt_fused_stencil1d_f64 per GPU -- the recipe INTEGRATION.md describes for embedding the fused
// kernels into an existing C++ code.  One rank per GPU (use the engine-based programs for
// oversubscription).
//
//   gmt_stencil1d [n_global_Mi] [--points N] [--ranks R] [--iters K]
//
// Printed lines as in the reference: banner (rank 0), "r/N exchange time %0.8f" (seconds; here the
// EXPOSED cost: fused step minus the same sweep without neighbours) and
// "r/N [dev:0xid] err_norm = %.8f".  Extra line: step time and GLUP/s.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>

#include "gmt_b200.h"

#define CHECK(what, call)                                                                      \
  do {                                                                                         \
    int rc_ = (call);                                                                          \
    if (rc_ != 0) {                                                                            \
      std::fprintf(stderr, "%s:%d error %d (%s): %s\n", __FILE__, __LINE__, rc_, what, gmt_last_error()); \
      return 2;                                                                                \
    }                                                                                          \
  } while (0)
#define CUDA_OK(call)                                                                          \
  do {                                                                                         \
    cudaError_t e_ = (call);                                                                   \
    if (e_ != cudaSuccess) {                                                                   \
      std::fprintf(stderr, "%s(%i): CUDA Error (%s) %i: %s\n", __FILE__, __LINE__, #call, (int)e_, cudaGetErrorString(e_)); \
      return 2;                                                                                \
    }                                                                                          \
  } while (0)

struct Rank {
  int dev = 0;
  double *z = nullptr, *out = nullptr;
  char* ctl = nullptr;   // 4 inboxes + 2 flags + epoch/done/status/err + scratch, zeroed
  gmt_halo1d_rank* desc = nullptr;        // device copy, neighbours wired
  gmt_halo1d_rank* desc_alone = nullptr;  // device copy with has = {0,0}: the sweep without exchange
  cudaStream_t stream = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
};
// byte offsets inside Rank::ctl (256-byte lines so flags and inboxes never share a line)
enum { OFF_INBOX = 0 /* [side][parity] * 256 */, OFF_FLAG = 1024 /* [side] * 256 */, OFF_EPOCH = 1536,
       OFF_DONE = 1792, OFF_STATUS = 2048, OFF_ERR = 2304, OFF_EPOCH2 = 2560, OFF_DONE2 = 2816, OFF_SCRATCH = 3072 };

int main(int argc, char** argv)
{
  long long n_global = 32LL << 20;
  int ranks = 0, iters = 20;
  for (int i = 1; i < argc; ++i) {
    if (!std::strcmp(argv[i], "--points") && i + 1 < argc) n_global = std::atoll(argv[++i]);
    else if (!std::strcmp(argv[i], "--ranks") && i + 1 < argc) ranks = std::atoi(argv[++i]);
    else if (!std::strcmp(argv[i], "--iters") && i + 1 < argc) iters = std::atoi(argv[++i]);
    else if (argv[i][0] != '-') n_global = std::atoll(argv[i]) * 1024 * 1024;
    else {
      std::fprintf(stderr, "usage: gmt_stencil1d [n_global_Mi] [--points N] [--ranks R] [--iters K]\n");
      return 1;
    }
  }
  int n_dev = 0;
  CHECK("device count", gmt_device_count(&n_dev));
  if (n_dev == 0) {
    std::fprintf(stderr, "gmt_stencil1d: no CUDA device\n");
    return 2;
  }
  if (ranks <= 0) ranks = n_dev;
  if (ranks > n_dev) {
    std::fprintf(stderr, "gmt_stencil1d: one rank per GPU (%d ranks > %d GPUs)\n", ranks, n_dev);
    return 1;
  }
  const int W = ranks;
  if (n_global % W != 0) {   // mpi_stencil_gt.cc:141-145
    std::printf("%d nmpi (%d) must be divisor of domain size (%lld), exiting\n", 0, W, n_global);
    return 1;
  }
  const long long n = n_global / W;
  if (n % 2 || n < 8) {
    std::fprintf(stderr, "gmt_stencil1d: n_local must be even and >= 8\n");
    return 1;
  }
  std::printf("n procs  = %d\nn_global = %lld\nn_local  = %lld\n", W, n_global, n);
  const double dx = 8.0 / double(n_global), scale = double(n_global) / 8.0;
  const size_t ctl_bytes = OFF_SCRATCH + gmt_reduce_scratch_bytes();

  std::vector<Rank> R(W);
  for (int r = 0; r < W; ++r) {
    R[r].dev = r;
    CUDA_OK(cudaSetDevice(r));
    if (r > 0) CHECK("peer", gmt_enable_peer_access(r - 1));
    if (r + 1 < W) CHECK("peer", gmt_enable_peer_access(r + 1));
    CUDA_OK(cudaMalloc(&R[r].z, size_t(n + 4) * 8));
    CUDA_OK(cudaMalloc(&R[r].out, size_t(n) * 8));
    CUDA_OK(cudaMalloc(&R[r].ctl, ctl_bytes));
    CUDA_OK(cudaMemset(R[r].ctl, 0, ctl_bytes));
    CUDA_OK(cudaMalloc(&R[r].desc, sizeof(gmt_halo1d_rank)));
    CUDA_OK(cudaMalloc(&R[r].desc_alone, sizeof(gmt_halo1d_rank)));
    CUDA_OK(cudaStreamCreateWithFlags(&R[r].stream, cudaStreamNonBlocking));
    CUDA_OK(cudaEventCreate(&R[r].e0));
    CUDA_OK(cudaEventCreate(&R[r].e1));
  }
  for (int r = 0; r < W; ++r) {
    gmt_halo1d_rank s;
    std::memset(&s, 0, sizeof(s));
    s.z = R[r].z, s.out = R[r].out, s.n = n, s.scale = scale;
    const int nbr[2] = {r - 1, r + 1 < W ? r + 1 : -1};
    for (int side = 0; side < 2; ++side) {
      const int nb = nbr[side];
      s.has[side] = nb >= 0;
      for (int p = 0; p < 2; ++p) {
        s.my_inbox[side][p] = (double*)(R[r].ctl + OFF_INBOX + (side * 2 + p) * 256);
        // my boundary goes into the neighbour's inbox of the OPPOSITE side
        s.peer_inbox[side][p] = nb >= 0 ? (double*)(R[nb].ctl + OFF_INBOX + ((1 - side) * 2 + p) * 256) : nullptr;
      }
      s.my_flag[side] = (uint64_t*)(R[r].ctl + OFF_FLAG + side * 256);
      s.peer_flag[side] = nb >= 0 ? (uint64_t*)(R[nb].ctl + OFF_FLAG + (1 - side) * 256) : nullptr;
    }
    s.epoch = (uint64_t*)(R[r].ctl + OFF_EPOCH);
    s.done = (uint32_t*)(R[r].ctl + OFF_DONE);
    s.status = (uint32_t*)(R[r].ctl + OFF_STATUS);
    s.err_scratch = R[r].ctl + OFF_SCRATCH;
    s.err_sumsq = (double*)(R[r].ctl + OFF_ERR);
    s.x0 = r * n * dx, s.dx = dx, s.spin_limit = 0;
    CUDA_OK(cudaSetDevice(r));
    CUDA_OK(cudaMemcpy(R[r].desc, &s, sizeof(s), cudaMemcpyHostToDevice));
    s.has[0] = s.has[1] = 0;   // same sweep, ghosts used as they are (own epoch/done words)
    s.epoch = (uint64_t*)(R[r].ctl + OFF_EPOCH2);
    s.done = (uint32_t*)(R[r].ctl + OFF_DONE2);
    CUDA_OK(cudaMemcpy(R[r].desc_alone, &s, sizeof(s), cudaMemcpyHostToDevice));
    // y = x^3; analytic ghosts only at the two physical ends (mpi_stencil_gt.cc:177-196)
    CHECK("init", gmt_init_stencil1d_f64(n, 2, r * n * dx, dx, R[r].z, nullptr, r == 0, r == W - 1, R[r].stream));
  }
  auto sync_all = [&]() -> int {
    for (int r = 0; r < W; ++r) {
      CUDA_OK(cudaSetDevice(r));
      CUDA_OK(cudaStreamSynchronize(R[r].stream));
    }
    return 0;
  };
  // device milliseconds per step, max over GPUs
  auto time_steps = [&](bool alone, int count, double* ms_out) -> int {
    for (int r = 0; r < W; ++r) {
      CUDA_OK(cudaSetDevice(r));
      CUDA_OK(cudaEventRecord(R[r].e0, R[r].stream));
    }
    for (int it = 0; it < count; ++it)
      for (int r = 0; r < W; ++r) {
        CUDA_OK(cudaSetDevice(r));
        CHECK("fused step", gmt_fused_stencil1d_f64(alone ? R[r].desc_alone : R[r].desc, 1, n, 1, R[r].stream));
      }
    float worst = 0.f;
    for (int r = 0; r < W; ++r) {
      CUDA_OK(cudaSetDevice(r));
      CUDA_OK(cudaEventRecord(R[r].e1, R[r].stream));
    }
    if (sync_all()) return 2;
    for (int r = 0; r < W; ++r) {
      float ms = 0.f;
      CUDA_OK(cudaEventElapsedTime(&ms, R[r].e0, R[r].e1));
      worst = std::max(worst, ms);
    }
    *ms_out = worst / count;
    return 0;
  };
  if (sync_all()) return 2;
  double ms_fused = 0.0, ms_alone = 0.0, warm = 0.0;
  if (time_steps(false, 3, &warm)) return 2;       // warm-up; also fills the interior-facing ghosts
  if (time_steps(true, 3, &warm)) return 2;
  if (time_steps(false, iters, &ms_fused)) return 2;
  if (time_steps(true, iters, &ms_alone)) return 2;
  if (time_steps(false, 1, &warm)) return 2;       // the error norm reported is the exchanging step's

  int bad = 0;
  for (int r = 0; r < W; ++r) {
    uint32_t status = 0;
    double ss = 0.0;
    gmt_device_props props;
    CUDA_OK(cudaSetDevice(r));
    CUDA_OK(cudaMemcpy(&status, R[r].ctl + OFF_STATUS, 4, cudaMemcpyDeviceToHost));
    CUDA_OK(cudaMemcpy(&ss, R[r].ctl + OFF_ERR, 8, cudaMemcpyDeviceToHost));
    CHECK("props", gmt_device_props_get(r, &props));
    const unsigned vendor = ((props.pci_domain & 0xFFFF) << 16) | ((props.pci_bus & 0xFF) << 8) | (props.pci_device & 0xFF);
    std::printf("%d/%d exchange time %0.8f\n", r, W, std::max(0.0, ms_fused - ms_alone) * 1e-3);
    std::printf("%d/%d [%d:0x%08x] err_norm = %.8f\n", r, W, r, vendor, std::sqrt(ss));
    if (status) {
      std::fprintf(stderr, "%d/%d a halo flag never arrived (status %u)\n", r, W, status);
      bad = 1;
    }
    if (!(std::sqrt(ss) < 1e-2 * std::sqrt(double(n) / 33554432.0) + 1e-3)) bad = 1;
  }
  std::printf("step %0.4f ms (%0.4f ms without exchange), %0.1f GLUP/s\n", ms_fused, ms_alone,
              double(n_global) / (ms_fused * 1e-3) / 1e9);
  for (int r = 0; r < W; ++r) {
    cudaSetDevice(r);
    cudaFree(R[r].desc_alone), cudaFree(R[r].desc), cudaFree(R[r].ctl), cudaFree(R[r].out), cudaFree(R[r].z);
    cudaEventDestroy(R[r].e0), cudaEventDestroy(R[r].e1), cudaStreamDestroy(R[r].stream);
  }
  return bad ? 3 : 0;
}
