# Build / test / run targets (the reference ships a bare Makefile with one rule per program,
# Makefile:1-33; here the programs are Python modules and the native code is one library).
PY ?= python
NP ?= 2
TORCHRUN = $(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node $(NP) --master-addr 127.0.0.1 --master-port 29555

.PHONY: all build force test test-gpu bench sweep sass tools clean \
        daxpy daxpy_nvtx mpi_daxpy mpi_daxpy_nvtx_managed mpi_daxpy_nvtx_unmanaged mpi_daxpy_gt \
        mpi_stencil_gt mpi_stencil2d_gt mpi_stencil2d_sycl mpi_stencil2d_sycl_oo mpienv mpigatherinplace \
        gmt_stencil2d gmt_stencil2d_sycl gmt_stencil1d gmt_daxpy gmt_mpi_daxpy gmt_mpi_daxpy_gt gmt_mpienv gmt_daxpy_gather gmt_daxpy_allreduce gmt_gatherinplace

all: build

build:
	$(PY) -m gpu_mpi_tests_b200._native.build

force:
	$(PY) -m gpu_mpi_tests_b200._native.build --force

test: build
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu: build
	$(PY) -m pytest tests -x -q -m gpu

bench: build
	$(PY) bench.py

sweep: build
	$(TORCHRUN) benchmarks/sweep.py

sass: build
	$(PY) tools/sass_evidence.py

# one target per reference program (same names as the reference's Makefile / CMake targets)
daxpy daxpy_nvtx:
	$(PY) -m gpu_mpi_tests_b200.cli.$@
mpi_daxpy mpi_daxpy_gt mpi_stencil_gt mpi_stencil2d_gt mpi_stencil2d_sycl mpi_stencil2d_sycl_oo mpienv mpigatherinplace:
	$(TORCHRUN) -m gpu_mpi_tests_b200.cli.$@ $(ARGS)
mpi_daxpy_nvtx_managed:
	$(TORCHRUN) -m gpu_mpi_tests_b200.cli.mpi_daxpy_nvtx --managed $(ARGS)
mpi_daxpy_nvtx_unmanaged:
	$(TORCHRUN) -m gpu_mpi_tests_b200.cli.mpi_daxpy_nvtx $(ARGS)

# native programs (C++ on the C ABI, no Python): one process drives every GPU of the node
gmt_stencil2d gmt_stencil2d_sycl gmt_stencil1d gmt_daxpy gmt_mpi_daxpy gmt_mpi_daxpy_gt gmt_mpienv gmt_daxpy_gather gmt_daxpy_allreduce gmt_gatherinplace: build
	gpu_mpi_tests_b200/_native/bin/$@ $(ARGS)

# measurement tools (NVLink peer bandwidth, HBM streaming patterns)
tools:
	mkdir -p tools/bin
	nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/peer_bw tools/peer_bw.cu
	nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/bw_sweep tools/bw_sweep.cu

clean:
	rm -rf gpu_mpi_tests_b200/_native/obj gpu_mpi_tests_b200/_native/bin gpu_mpi_tests_b200/_native/libgmt_b200.so out-*.txt profile
