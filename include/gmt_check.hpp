// gmt_check.hpp -- error-check and pointer-report helpers for PROGRAMS built on libgmt_b200:
// the surface of the reference's cuda_error.h (CHECK / WARN / PTRINFO / MEMINFO, cuda_error.h:16-136)
// with overloads for the three kinds of status a program here sees: gmt codes (int), CUDA runtime
// errors (cudaError_t, when <cuda_runtime_api.h> was included first) and NCCL results (when <nccl.h>
// was included first).
//
//   GMT_CHECK("what", call)    on failure: "file(line): <kind> Error (what) code: text" on stderr, exit(1)
//   GMT_WARN("what", call)     same line, no exit; evaluates to the status as an int
//   GMT_PTRINFO(ptr)           "CUDA pointer <name> (<hex>): Device|Managed|Host|Unregistered|NULL"
//   GMT_MEMINFO(ptr, bytes)    "CUDA PreferredLocation of '<name>' is NOT CUDA|UNMANAGED|CPU (id)|INVALID (id)|DEVICE (id)"
//
// Checks are ON by default; compile with -DGMT_NO_CHECK_CALLS to turn GMT_CHECK / GMT_WARN into plain
// evaluation of the call (the reference's GPU_NO_CHECK_CALLS switch).  Unlike the reference's CHECK,
// nothing resets the device before exiting: the library may own peer mappings other processes use.
#ifndef GMT_CHECK_HPP
#define GMT_CHECK_HPP

#include <cstdio>
#include <cstdlib>

#include "gmt_b200.h"

namespace gmt {
namespace detail {

inline int report(const char* kind, const char* what, int code, const char* text, const char* file, int line)
{
  std::fprintf(stderr, "%s(%i): %s Error (%s) %i: %s\n", file, line, kind, what, code, text);
  return code;
}

inline int status(int rc, const char* what, const char* file, int line)
{
  return rc == 0 ? 0 : report("GMT", what, rc, gmt_last_error(), file, line);
}
#if defined(__CUDA_RUNTIME_API_H__) || defined(__CUDA_RUNTIME_H__)
inline int status(cudaError_t e, const char* what, const char* file, int line)
{
  return e == cudaSuccess ? 0 : report("CUDA", what, (int)e, cudaGetErrorString(e), file, line);
}
#endif
#if defined(NCCL_H_)
inline int status(ncclResult_t r, const char* what, const char* file, int line)
{
  return r == ncclSuccess ? 0 : report("NCCL", what, (int)r, ncclGetErrorString(r), file, line);
}
#endif

}  // namespace detail

inline void print_ptr_info(const char* label, const void* ptr)
{
  static const char* const kName[] = {"Unregistered", "Host", "Device", "Managed"};
  int type = 0, dev = 0;
  if (!ptr) {
    std::printf("CUDA pointer %s (%zx): NULL\n", label, (size_t)0);
    return;
  }
  if (gmt_ptr_info(ptr, &type, &dev) != 0) {
    std::fprintf(stderr, "get pointer attr: %s\n", gmt_last_error());
    return;
  }
  std::printf("CUDA pointer %s (%zx): %s\n", label, (size_t)ptr, type >= 0 && type <= 3 ? kName[type] : "Unregistered");
}

inline void print_mem_info(const char* label, const void* ptr, size_t bytes)
{
  int loc = 0;
  if (gmt_mem_preferred_location(ptr, bytes, &loc) != 0) {
    std::fprintf(stderr, "get mem range preferred location: %s\n", gmt_last_error());
    return;
  }
  if (loc == -3) std::printf("CUDA PreferredLocation of '%s' is NOT CUDA\n", label);
  else if (loc == -2) std::printf("CUDA PreferredLocation of '%s' is UNMANAGED\n", label);
  else if (loc == -1) std::printf("CUDA PreferredLocation of '%s' is CPU (%d)\n", label, -1);
  else if (loc == -4) std::printf("CUDA PreferredLocation of '%s' is INVALID (%d)\n", label, -2);
  else std::printf("CUDA PreferredLocation of '%s' is DEVICE (%d)\n", label, loc);
}

}  // namespace gmt

#ifndef GMT_NO_CHECK_CALLS
#define GMT_CHECK(what, call)                                                   \
  do {                                                                          \
    if (gmt::detail::status((call), (what), __FILE__, __LINE__) != 0) std::exit(EXIT_FAILURE); \
  } while (0)
#define GMT_WARN(what, call) gmt::detail::status((call), (what), __FILE__, __LINE__)
#else
#define GMT_CHECK(what, call) ((void)(call))
#define GMT_WARN(what, call) ((int)(call))
#endif
#define GMT_PTRINFO(ptr) gmt::print_ptr_info(#ptr, (ptr))
#define GMT_MEMINFO(ptr, bytes) gmt::print_mem_info(#ptr, (ptr), (bytes))

#endif  // GMT_CHECK_HPP
