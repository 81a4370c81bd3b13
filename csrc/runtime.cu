// runtime.cu -- device introspection, memory spaces, IPC export, tracing.
//
// Native equivalents of the reference's set_rank_device plumbing
// (mpi_daxpy.cc:36-62), vendor/PCI id report (mpi_stencil_gt.cc:151-152),
// cudaMemGetInfo line (mpi_daxpy_nvtx.cc:201-205), PTRINFO/MEMINFO
// (cuda_error.h:66-136), the three allocation flavours
// (mpi_daxpy_nvtx.cc:177-199) and the NVTX / profiler window
// (daxpy_nvtx.cu:65-105).  The rank->device *policy* lives in Python
// (parallel/runtime.py); this file only answers questions about the machine.
#include <cuda_profiler_api.h>
#include <nvtx3/nvToolsExt.h>

#include <cstring>

#include "common.cuh"

#include <map>
#include <mutex>
#include <utility>

namespace gmt {

int stream_counters(cudaStream_t stream, unsigned int** out)
{
  static std::mutex mu;
  static std::map<std::pair<int, cudaStream_t>, unsigned int*> table;
  int dev = 0;
  GMT_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(mu);
  auto key = std::make_pair(dev, stream);
  auto it = table.find(key);
  if (it == table.end()) {
    unsigned int* p = nullptr;
    GMT_CUDA(cudaMalloc(&p, 64));
    GMT_CUDA(cudaMemset(p, 0, 64));
    it = table.emplace(key, p).first;
  }
  *out = it->second;
  return 0;
}

}  // namespace gmt

namespace gmt {
// Test hook: a kernel that does nothing but hold SM resources for a while.
__global__ void __launch_bounds__(1024) occupy_kernel(unsigned long long ns)
{
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (t - t0 >= ns) break;
    __nanosleep(1000);
  }
}
}  // namespace gmt

extern "C" {

const char* gmt_last_error(void) { return gmt::err_buf(); }

int gmt_version(void) { return 100; }

int gmt_device_count(int* count)
{
  GMT_REQUIRE(count, GMT_EINVAL, "gmt_device_count: NULL out");
  cudaError_t e = cudaGetDeviceCount(count);
  if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) {
    (void)cudaGetLastError();
    *count = 0;
    return 0;
  }
  GMT_CUDA(e);
  return 0;
}

int gmt_set_device(int dev)
{
  GMT_CUDA(cudaSetDevice(dev));
  return 0;
}

int gmt_get_device(int* dev)
{
  GMT_CUDA(cudaGetDevice(dev));
  return 0;
}

int gmt_device_props_get(int dev, gmt_device_props* out)
{
  GMT_REQUIRE(out, GMT_EINVAL, "gmt_device_props_get: NULL out");
  cudaDeviceProp p;
  GMT_CUDA(cudaGetDeviceProperties(&p, dev));
  memset(out, 0, sizeof(*out));
  strncpy(out->name, p.name, sizeof(out->name) - 1);
  out->total_global_mem = p.totalGlobalMem;
  out->pci_domain = p.pciDomainID;
  out->pci_bus = p.pciBusID;
  out->pci_device = p.pciDeviceID;
  out->sm_count = p.multiProcessorCount;
  out->cc_major = p.major;
  out->cc_minor = p.minor;
  out->l2_bytes = p.l2CacheSize;
  out->managed_memory = p.managedMemory;
  out->concurrent_managed_access = p.concurrentManagedAccess;
  out->max_smem_optin = (int)p.sharedMemPerBlockOptin;
  int v = 0;
  if (cudaDeviceGetAttribute(&v, cudaDevAttrClockRate, dev) == cudaSuccess) out->clock_khz = v;
  if (cudaDeviceGetAttribute(&v, cudaDevAttrMemoryClockRate, dev) == cudaSuccess)
    out->mem_clock_khz = v;
  memcpy(out->uuid, p.uuid.bytes, 16);
  return 0;
}

int gmt_mem_info(uint64_t* free_bytes, uint64_t* total_bytes)
{
  size_t f = 0, t = 0;
  GMT_CUDA(cudaMemGetInfo(&f, &t));
  if (free_bytes) *free_bytes = f;
  if (total_bytes) *total_bytes = t;
  return 0;
}

int gmt_ptr_info(const void* p, int* type, int* device)
{
  GMT_REQUIRE(type && device, GMT_EINVAL, "gmt_ptr_info: NULL out");
  *device = -1;
  if (p == nullptr) {
    *type = -1;
    return 0;
  }
  cudaPointerAttributes a;
  cudaError_t e = cudaPointerGetAttributes(&a, p);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    *type = 0;
    return 0;
  }
  switch (a.type) {
    case cudaMemoryTypeHost: *type = 1; break;
    case cudaMemoryTypeDevice: *type = 2; break;
    case cudaMemoryTypeManaged: *type = 3; break;
    default: *type = 0; break;
  }
  *device = a.device;
  return 0;
}

int gmt_mem_preferred_location(const void* p, size_t bytes, int* loc)
{
  GMT_REQUIRE(loc, GMT_EINVAL, "gmt_mem_preferred_location: NULL out");
  int type = 0, dev = 0;
  gmt_ptr_info(p, &type, &dev);
  if (type <= 0 || type == 1) {
    *loc = -3;  // NOT CUDA (plain host memory counts as such for this report)
    if (type == 1) *loc = -2;
    return 0;
  }
  if (type != 3) {
    *loc = -2;  // UNMANAGED
    return 0;
  }
  int v = cudaInvalidDeviceId;
  cudaError_t e =
    cudaMemRangeGetAttribute(&v, sizeof(v), cudaMemRangeAttributePreferredLocation, p, bytes);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    *loc = -4;
    return 0;
  }
  if (v == cudaCpuDeviceId)
    *loc = -1;
  else if (v == cudaInvalidDeviceId)
    *loc = -4;
  else
    *loc = v;
  return 0;
}

int gmt_can_access_peer(int dev, int peer, int* ok)
{
  GMT_CUDA(cudaDeviceCanAccessPeer(ok, dev, peer));
  return 0;
}

int gmt_enable_peer_access(int peer)
{
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    (void)cudaGetLastError();
    return 0;
  }
  GMT_CUDA(e);
  return 0;
}

int gmt_device_synchronize(void)
{
  GMT_CUDA(cudaDeviceSynchronize());
  return 0;
}

int gmt_stream_synchronize(gmt_stream_t stream)
{
  GMT_CUDA(cudaStreamSynchronize(gmt::as_stream(stream)));
  return 0;
}

int gmt_alloc(int space, size_t bytes, void** out)
{
  GMT_REQUIRE(out, GMT_EINVAL, "gmt_alloc: NULL out");
  *out = nullptr;
  if (bytes == 0) bytes = 8;
  switch (space) {
    case GMT_SPACE_DEVICE: GMT_CUDA(cudaMalloc(out, bytes)); break;
    case GMT_SPACE_MANAGED: GMT_CUDA(cudaMallocManaged(out, bytes, cudaMemAttachGlobal)); break;
    case GMT_SPACE_PINNED: GMT_CUDA(cudaMallocHost(out, bytes)); break;
    default: return gmt::fail(GMT_EINVAL, "gmt_alloc: unknown memory space %d", space);
  }
  return 0;
}

int gmt_free(int space, void* p)
{
  if (!p) return 0;
  if (space == GMT_SPACE_PINNED)
    GMT_CUDA(cudaFreeHost(p));
  else
    GMT_CUDA(cudaFree(p));
  return 0;
}

int gmt_memcpy_async(void* dst, const void* src, size_t bytes, int kind, gmt_stream_t stream)
{
  static const cudaMemcpyKind kinds[5] = {cudaMemcpyHostToHost, cudaMemcpyHostToDevice,
                                          cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice,
                                          cudaMemcpyDefault};
  GMT_REQUIRE(kind >= 0 && kind < 5, GMT_EINVAL, "gmt_memcpy_async: bad kind %d", kind);
  GMT_CUDA(cudaMemcpyAsync(dst, src, bytes, kinds[kind], gmt::as_stream(stream)));
  return 0;
}

int gmt_memset_async(void* dst, int value, size_t bytes, gmt_stream_t stream)
{
  GMT_CUDA(cudaMemsetAsync(dst, value, bytes, gmt::as_stream(stream)));
  return 0;
}

int gmt_ipc_get_handle(void* dev_ptr, unsigned char handle[64])
{
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  GMT_CUDA(cudaIpcGetMemHandle(&h, dev_ptr));
  memcpy(handle, &h, 64);
  return 0;
}

int gmt_ipc_open_handle(const unsigned char handle[64], void** out)
{
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  GMT_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int gmt_ipc_close_handle(void* mapped)
{
  GMT_CUDA(cudaIpcCloseMemHandle(mapped));
  return 0;
}

int gmt_debug_occupy(int ctas, unsigned long long ns, gmt_stream_t stream)
{
  GMT_REQUIRE(ctas >= 1, GMT_EINVAL, "gmt_debug_occupy: ctas must be positive");
  gmt::occupy_kernel<<<ctas, 1024, 0, gmt::as_stream(stream)>>>(ns);
  GMT_LAUNCH_CHECK();
  return 0;
}

int gmt_nvtx_push(const char* name)
{
  nvtxRangePushA(name);
  return 0;
}

int gmt_nvtx_pop(void)
{
  nvtxRangePop();
  return 0;
}

int gmt_profiler_start(void)
{
  GMT_CUDA(cudaProfilerStart());
  return 0;
}

int gmt_profiler_stop(void)
{
  GMT_CUDA(cudaProfilerStop());
  return 0;
}

}  // extern "C"
