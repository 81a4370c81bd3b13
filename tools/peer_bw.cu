// peer_bw.cu -- what one GPU can push into a peer over NVLink 5 / NVSwitch on this box: the
// denominator behind every "fraction of NVLink" figure in BASELINE.md / profiles.
//   (a) cudaMemcpyPeerAsync, one direction and both directions at once (copy engines);
//   (b) a store kernel (256-bit st.global to the peer mapping, the collectives' pattern), one
//       destination and fanned out over all peers (the allgather pattern);
// device-timed with CUDA events, best of 5, JSON lines on stdout.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/peer_bw tools/peer_bw.cu
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e_ = (x);                                                              \
    if (e_ != cudaSuccess) {                                                           \
      std::fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); \
      std::exit(2);                                                                    \
    }                                                                                  \
  } while (0)

struct __align__(32) d4 { double x, y, z, w; };

// every thread copies vectors t, t+stride, ... of src into each of the n_dst destinations
__global__ void __launch_bounds__(512, 2) fan_store(const d4* __restrict__ src, d4* const* __restrict__ dst, int n_dst, long nvec)
{
  const long stride = (long)gridDim.x * blockDim.x;
  for (long v = (long)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    const d4 a = src[v];
    for (int k = 0; k < n_dst; ++k) {
      d4* p = dst[k] + v;
      asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a.x), "d"(a.y), "d"(a.z), "d"(a.w) : "memory");
    }
  }
}

int main(int argc, char** argv)
{
  int nd = 0;
  CK(cudaGetDeviceCount(&nd));
  if (nd < 2) {
    std::printf("{\"config\": \"peer_bw\", \"error\": \"needs 2 GPUs, found %d\"}\n", nd);
    return 0;
  }
  const size_t bytes = argc > 1 ? (size_t)std::atoll(argv[1]) : (size_t)256 << 20;
  std::vector<void*> buf(nd), buf2(nd);
  for (int d = 0; d < nd; ++d) {
    CK(cudaSetDevice(d));
    for (int p = 0; p < nd; ++p)
      if (p != d) {
        int ok = 0;
        CK(cudaDeviceCanAccessPeer(&ok, d, p));
        if (ok) cudaDeviceEnablePeerAccess(p, 0);
      }
    CK(cudaMalloc(&buf[d], bytes));
    CK(cudaMalloc(&buf2[d], bytes));
    CK(cudaMemset(buf[d], 1, bytes));
  }
  cudaGetLastError();
  auto time_best = [&](auto&& body) {
    float best = 1e30f;
    for (int it = 0; it < 6; ++it) {
      cudaEvent_t e0, e1;
      CK(cudaSetDevice(0));
      CK(cudaEventCreate(&e0));
      CK(cudaEventCreate(&e1));
      for (int d = 0; d < nd; ++d) {
        CK(cudaSetDevice(d));
        CK(cudaDeviceSynchronize());
      }
      CK(cudaSetDevice(0));
      CK(cudaEventRecord(e0, 0));
      body();
      CK(cudaSetDevice(0));
      CK(cudaEventRecord(e1, 0));
      for (int d = 0; d < nd; ++d) {
        CK(cudaSetDevice(d));
        CK(cudaDeviceSynchronize());
      }
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (it > 0) best = std::min(best, ms);
      cudaEventDestroy(e0);
      cudaEventDestroy(e1);
    }
    return best;
  };
  // (a) copy engine, 0 -> 1
  {
    float ms = time_best([&] { CK(cudaMemcpyPeerAsync(buf2[1], 1, buf[0], 0, bytes, 0)); });
    std::printf("{\"config\": \"peer_bw\", \"how\": \"cudaMemcpyPeerAsync 0->1\", \"bytes\": %zu, \"ms\": %.4f, \"gbs\": %.1f}\n",
                bytes, ms, bytes / ms / 1e6);
  }
  // (b) store kernel, 0 -> 1 and 0 -> all peers (egress of GPU 0)
  for (int fan : {1, nd - 1}) {
    if (fan < 1) continue;
    CK(cudaSetDevice(0));
    std::vector<d4*> h(fan);
    for (int k = 0; k < fan; ++k) h[k] = (d4*)buf2[1 + k];
    d4** dtab = nullptr;
    CK(cudaMalloc(&dtab, sizeof(d4*) * fan));
    CK(cudaMemcpy(dtab, h.data(), sizeof(d4*) * fan, cudaMemcpyHostToDevice));
    const size_t per = bytes / fan / 32 * 32;   // same total egress in both cases
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    float ms = time_best([&] { fan_store<<<sms * 2, 512>>>((const d4*)buf[0], dtab, fan, (long)(per / 32)); });
    CK(cudaGetLastError());
    std::printf("{\"config\": \"peer_bw\", \"how\": \"store kernel 0 -> %d peer(s), 256-bit stores\", \"egress_bytes\": %zu, "
                "\"ms\": %.4f, \"egress_gbs\": %.1f}\n", fan, per * fan, ms, per * fan / ms / 1e6);
    cudaFree(dtab);
  }
  // (c) all GPUs store to all peers at once (the allgather pattern): ingress per GPU
  if (nd >= 2) {
    std::vector<d4**> tabs(nd);
    const size_t per = bytes / (nd - 1) / 32 * 32;
    for (int d = 0; d < nd; ++d) {
      CK(cudaSetDevice(d));
      std::vector<d4*> h;
      for (int k = 1; k < nd; ++k) h.push_back((d4*)((char*)buf2[(d + k) % nd] + 0));
      CK(cudaMalloc(&tabs[d], sizeof(d4*) * (nd - 1)));
      CK(cudaMemcpy(tabs[d], h.data(), sizeof(d4*) * (nd - 1), cudaMemcpyHostToDevice));
    }
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    float ms = time_best([&] {
      for (int d = 0; d < nd; ++d) {
        CK(cudaSetDevice(d));
        fan_store<<<sms * 2, 512>>>((const d4*)buf[d], tabs[d], nd - 1, (long)(per / 32));
      }
    });
    std::printf("{\"config\": \"peer_bw\", \"how\": \"all %d GPUs store to all peers at once\", \"ingress_bytes_per_gpu\": %zu, "
                "\"ms\": %.4f, \"ingress_gbs_per_gpu\": %.1f, \"note\": \"event on GPU 0 brackets launches on all GPUs; overlapping writes to the same addresses, bandwidth only\"}\n",
                nd, per * (nd - 1), ms, per * (nd - 1) / ms / 1e6);
  }
  return 0;
}
