#!/bin/bash
# Batch matrix on one 8xB200 node: the job.lsf / job.pbs of the reference (summit/job.lsf:9-16,
# jlse/job.pbs:14-21) -- every (managed?, profiler, ranks) combination of the daxpy benchmark,
# then the rank-average of the gather time of each output file.
set -e
cd "$(dirname "$0")/.."
python -m gpu_mpi_tests_b200._native.build
for ppn in 8 4 2; do
  for um in noum um; do
    python -m gpu_mpi_tests_b200.cli.run $um none 1 $ppn
  done
done
# one profiled run (the cudaProfilerStart/Stop window limits what ncu records)
python -m gpu_mpi_tests_b200.cli.run noum ncu 1 2 || true
python -m gpu_mpi_tests_b200.cli.avg gather
python -m gpu_mpi_tests_b200.cli.avg kernel
