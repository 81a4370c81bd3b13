#!/bin/bash
# Environment notes (jlse/setup.sh equivalent): nothing to load -- CUDA 12.9 under
# /usr/local/cuda, PyTorch 2.11 (NCCL 2.28 bundled) in the default python.  Build with:
python -m gpu_mpi_tests_b200._native.build
