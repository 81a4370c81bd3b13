"""One program per reference executable, same names, positional arguments and printed lines:

    python -m gpu_mpi_tests_b200.cli.daxpy                  daxpy.cu
    python -m gpu_mpi_tests_b200.cli.daxpy_nvtx             daxpy_nvtx.cu
    torchrun ... -m gpu_mpi_tests_b200.cli.mpi_daxpy        mpi_daxpy.cc
    torchrun ... -m gpu_mpi_tests_b200.cli.mpi_daxpy_nvtx   mpi_daxpy_nvtx.cc  [--managed] [--barrier]
    torchrun ... -m gpu_mpi_tests_b200.cli.mpi_daxpy_gt     mpi_daxpy_gt.cc
    torchrun ... -m gpu_mpi_tests_b200.cli.mpi_stencil_gt [n_global_Mi]                 mpi_stencil_gt.cc
    torchrun ... -m gpu_mpi_tests_b200.cli.mpi_stencil2d_gt [n_local_deriv] [n_iter]    mpi_stencil2d_gt.cc
    torchrun ... -m gpu_mpi_tests_b200.cli.mpi_stencil2d_sycl [nx_local] [stage_host] [n_iter]
    torchrun ... -m gpu_mpi_tests_b200.cli.mpi_stencil2d_sycl_oo [n_global_Ki] [stage_host] [n_iter]
    torchrun ... -m gpu_mpi_tests_b200.cli.mpienv           mpienv.f90
    torchrun ... -m gpu_mpi_tests_b200.cli.mpigatherinplace mpigatherinplace.f90
    python -m gpu_mpi_tests_b200.cli.run um|noum ncu|none nodes ppn      summit/run.sh, jlse/run.sh
    python -m gpu_mpi_tests_b200.cli.avg [pattern]                       avg.sh

The reference's compile-time switches became flags: --managed / --space, --barrier, --debug,
--stage; --impl fused|baseline picks the peer-memory kernels or the library-only restatement
(default: fused on GPUs, baseline on the CPU/gloo path).
"""
