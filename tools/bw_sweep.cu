// bw_sweep.cu -- development microbenchmark: which streaming pattern gets the most out of
// B200 HBM3e for an axpy-shaped (2 reads + 1 write) and a copy-shaped (1r + 1w) kernel?
// Sweeps vector width, unroll, block size, persistent vs one-shot grids and cache hints.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/bw_sweep tools/bw_sweep.cu
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>

struct __align__(32) d4 { double x, y, z, w; };

template <int HINT> __device__ __forceinline__ d4 ld32(const double* p)
{
  d4 r;
  if (HINT == 0) asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
  if (HINT == 1) asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
  if (HINT == 2) asm volatile("ld.global.cs.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
  return r;
}
template <int HINT> __device__ __forceinline__ void st32(double* p, d4 v)
{
  if (HINT == 0) asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.x), "d"(v.y), "d"(v.z), "d"(v.w) : "memory");
  if (HINT == 1) asm volatile("st.global.L1::no_allocate.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.x), "d"(v.y), "d"(v.z), "d"(v.w) : "memory");
  if (HINT == 2) asm volatile("st.global.cs.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.x), "d"(v.y), "d"(v.z), "d"(v.w) : "memory");
}
template <int HINT> __device__ __forceinline__ double2 ld16(const double* p)
{
  double2 r;
  if (HINT == 0) asm volatile("ld.global.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
  if (HINT == 1) asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
  if (HINT == 2) asm volatile("ld.global.cs.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
  return r;
}
template <int HINT> __device__ __forceinline__ void st16(double* p, double2 v)
{
  if (HINT == 0) asm volatile("st.global.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
  if (HINT == 1) asm volatile("st.global.L1::no_allocate.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
  if (HINT == 2) asm volatile("st.global.cs.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
}

// OP 0: y = a x + y ; OP 1: y = x (copy)
// chunked: each CTA trip handles a contiguous chunk of blockDim*U vectors; within it thread t touches t, t+B, ...
template <int OP, int U, int LH, int SH, bool V32>
__global__ void stream_kernel(long nvec, double a, const double* __restrict__ x, double* __restrict__ y)
{
  constexpr int W = V32 ? 4 : 2;
  const long chunk = (long)blockDim.x * U;
  for (long base = (long)blockIdx.x * chunk; base < nvec; base += (long)gridDim.x * chunk) {
    if (V32) {
      d4 xv[U], yv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        long i = base + threadIdx.x + (long)u * blockDim.x;
        if (i < nvec) { xv[u] = ld32<LH>(x + W * i); if (OP == 0) yv[u] = ld32<LH == 1 ? 0 : LH>(y + W * i); }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        long i = base + threadIdx.x + (long)u * blockDim.x;
        if (i < nvec) {
          d4 r = xv[u];
          if (OP == 0) { r.x = a * xv[u].x + yv[u].x; r.y = a * xv[u].y + yv[u].y; r.z = a * xv[u].z + yv[u].z; r.w = a * xv[u].w + yv[u].w; }
          st32<SH>(y + W * i, r);
        }
      }
    } else {
      double2 xv[U], yv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        long i = base + threadIdx.x + (long)u * blockDim.x;
        if (i < nvec) { xv[u] = ld16<LH>(x + W * i); if (OP == 0) yv[u] = ld16<LH == 1 ? 0 : LH>(y + W * i); }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        long i = base + threadIdx.x + (long)u * blockDim.x;
        if (i < nvec) {
          double2 r = xv[u];
          if (OP == 0) { r.x = a * xv[u].x + yv[u].x; r.y = a * xv[u].y + yv[u].y; }
          st16<SH>(y + W * i, r);
        }
      }
    }
  }
}

static float time_kernel(void (*launch)(void*), void* ctx)
{
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) launch(ctx);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int i = 0; i < 10; ++i) {
    cudaEventRecord(a); launch(ctx); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); best = std::min(best, ms);
  }
  cudaEventDestroy(a); cudaEventDestroy(b);
  return best;
}

struct Ctx { long n; double* x; double* y; int block; int grid; };

template <int OP, int U, int LH, int SH, bool V32>
static void launch_it(void* p)
{
  Ctx* c = (Ctx*)p;
  long nvec = c->n / (V32 ? 4 : 2);
  stream_kernel<OP, U, LH, SH, V32><<<c->grid, c->block>>>(nvec, 2.0, c->x, c->y);
}

template <int OP, int U, int LH, int SH, bool V32>
static void run(Ctx c, int sms, const char* name)
{
  const long nvec = c.n / (V32 ? 4 : 2);
  const int blocks[] = {128, 256, 512, 1024};
  for (int b : blocks) {
    long oneshot = (nvec + (long)b * U - 1) / ((long)b * U);
    // persistent multipliers (CTAs per SM) and 0 = one-shot grid
    const int mults[] = {0, 1, 2, 4, 8};
    for (int m : mults) {
      if (m && (long)m * b > 2048) continue;
      c.block = b;
      c.grid = m == 0 ? (int)oneshot : (int)std::min<long>(oneshot, (long)sms * m);
      float ms = time_kernel(launch_it<OP, U, LH, SH, V32>, &c);
      double bytes = (OP == 0 ? 24.0 : 16.0) * c.n;
      printf("%s op=%s vec=%d U=%d LH=%d SH=%d block=%4d grid=%8d (%s) : %8.4f ms %8.1f GB/s\n", name,
             OP == 0 ? "axpy" : "copy", V32 ? 32 : 16, U, LH, SH, b, c.grid, m ? "persist" : "oneshot", ms,
             bytes / ms / 1e6);
    }
  }
}

int main(int argc, char** argv)
{
  long n = (argc > 1 ? atol(argv[1]) : 48L) << 20;
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  Ctx c;
  c.n = n;
  cudaMalloc(&c.x, n * 8);
  cudaMalloc(&c.y, n * 8);
  cudaMemset(c.x, 0, n * 8);
  cudaMemset(c.y, 0, n * 8);
  printf("n = %ld doubles, %d SMs\n", n, sms);
  // cudaMemcpy D2D as the vendor copy reference
  {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e30f;
    for (int i = 0; i < 8; ++i) { cudaEventRecord(a); cudaMemcpyAsync(c.y, c.x, n * 8, cudaMemcpyDeviceToDevice); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); best = std::min(best, ms); }
    printf("cudaMemcpy D2D: %8.4f ms %8.1f GB/s\n", best, 16.0 * n / best / 1e6);
  }
  run<0, 1, 0, 0, true>(c, sms, "A");
  run<0, 2, 0, 0, true>(c, sms, "A");
  run<0, 4, 0, 0, true>(c, sms, "A");
  run<0, 2, 1, 0, true>(c, sms, "A");
  run<0, 2, 0, 2, true>(c, sms, "A");
  run<0, 2, 1, 2, true>(c, sms, "A");
  run<0, 2, 2, 2, true>(c, sms, "A");
  run<0, 2, 1, 1, true>(c, sms, "A");
  run<0, 2, 0, 0, false>(c, sms, "A");
  run<0, 4, 0, 0, false>(c, sms, "A");
  run<0, 4, 1, 2, false>(c, sms, "A");
  run<1, 2, 0, 0, true>(c, sms, "C");
  run<1, 4, 0, 0, true>(c, sms, "C");
  run<1, 4, 1, 2, true>(c, sms, "C");
  run<1, 4, 1, 1, true>(c, sms, "C");
  run<1, 4, 0, 0, false>(c, sms, "C");
  return 0;
}
