// check.h -- error plumbing of libgmt_b200.
//
// Same surface as the reference's CHECK / WARN / PTRINFO / MEMINFO helpers
// (cuda_error.h:16-136) but library-shaped: instead of cudaDeviceReset()+exit()
// inside a macro, a failing call records a thread-local message and the entry
// point returns the code, so the Python layer decides whether to abort (the CLI
// programs do, with the reference's "file(line): CUDA Error (msg) code: text"
// line).
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>

namespace gmt {

inline char* err_buf()
{
  static thread_local char buf[512] = {0};
  return buf;
}

inline int fail(int code, const char* fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

inline int cuda_fail(cudaError_t e, const char* what, const char* file, int line)
{
  // clear the sticky-less error so the next launch check starts clean
  (void)cudaGetLastError();
  return fail((int)e, "%s(%i): CUDA Error (%s) %i: %s", file, line, what, (int)e,
              cudaGetErrorString(e));
}

}  // namespace gmt

// abort the entry point with the CUDA error code
#define GMT_CUDA(call)                                                          \
  do {                                                                          \
    cudaError_t gmt_e_ = (call);                                                \
    if (gmt_e_ != cudaSuccess) return gmt::cuda_fail(gmt_e_, #call, __FILE__, __LINE__); \
  } while (0)

// print-only flavour (the reference's WARN)
#define GMT_WARN(call)                                                          \
  do {                                                                          \
    cudaError_t gmt_e_ = (call);                                                \
    if (gmt_e_ != cudaSuccess) {                                                \
      gmt::cuda_fail(gmt_e_, #call, __FILE__, __LINE__);                        \
      fprintf(stderr, "%s\n", gmt::err_buf());                                  \
    }                                                                           \
  } while (0)

#define GMT_LAUNCH_CHECK() GMT_CUDA(cudaGetLastError())

#define GMT_REQUIRE(cond, code, ...)                   \
  do {                                                 \
    if (!(cond)) return gmt::fail((code), __VA_ARGS__); \
  } while (0)
