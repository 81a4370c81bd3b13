"""B200-native distributed-array and halo-exchange runtime with the capabilities of
bd4/gpu-mpi-tests: fp64 daxpy with a global sum, 1-D / 2-D finite-difference stencils with
ghost-cell exchange, in-place gather / allgather / all-reduce and rank/device reports.

Layers:
  _native/   libgmt_b200.so: sm_100a CUDA kernels behind a C ABI (include/gmt_b200.h)
  ops/       ctypes binding + single-rank ops (CUDA kernels, CPU oracle for gloo runs)
  parallel/  runtime (rank->GPU binding), decomposition, peer memory, fused exchange plans,
             in-kernel collectives
  baseline/  library-only (NCCL/gloo + cuBLAS-style) restatements of the reference algorithms
  models/    the workloads (daxpy, stencil1d, stencil2d, gather) behind one step() API
  cli/       one program per reference executable, same positional arguments and output
  utils/     oracle, exact output formats, timing, NVTX, avg
"""
__version__ = "0.1.0"
