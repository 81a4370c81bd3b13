#!/usr/bin/env bash
# Everything that feeds BASELINE.md for one GPU count: GPU tests, multi-process workers, bench (product and
# library restatement), the full sweep and the native programs.  Results go to gpurun_out/ (copy what should be
# kept into profiles/).      usage: scripts/measure.sh N        (run from the repository root, on a GPU box)
set -x
N=${1:-1}
P=${2:-29700}
mkdir -p gpurun_out
run() {
  if [ "$N" = 1 ]; then python "$@"; else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "$P" "$@"; P=$((P+1)); fi
}
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_n$N.log 2>&1; tail -4 gpurun_out/pytest_gpu_n$N.log
if [ "$N" != 1 ]; then
  for c in fused collectives baseline; do run tests/mp_worker.py $c 2>&1 | tail -2; done
fi
run bench.py --gpus "$N" > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; tail -c 1500 gpurun_out/bench_n$N.json
run bench.py --gpus "$N" --impl baseline --no-extra > gpurun_out/bench_base_n$N.json 2> gpurun_out/bench_base_n$N.err
tail -c 700 gpurun_out/bench_base_n$N.json
run benchmarks/sweep.py --out gpurun_out/sweep_n$N.jsonl > gpurun_out/sweep_n$N.log 2>&1; tail -3 gpurun_out/sweep_n$N.log
B=gpu_mpi_tests_b200/_native/bin
$B/gmt_stencil2d 1024 100 > gpurun_out/native_stencil2d_n$N.txt 2>&1; cat gpurun_out/native_stencil2d_n$N.txt
$B/gmt_stencil1d 512 > gpurun_out/native_stencil1d_n$N.txt 2>&1; tail -2 gpurun_out/native_stencil1d_n$N.txt
$B/gmt_daxpy_gather > gpurun_out/native_daxpy_gather_n$N.txt 2>&1; tail -2 gpurun_out/native_daxpy_gather_n$N.txt
$B/gmt_daxpy_allreduce > gpurun_out/native_daxpy_allreduce_n$N.txt 2>&1; tail -1 gpurun_out/native_daxpy_allreduce_n$N.txt
$B/gmt_gatherinplace > gpurun_out/native_gatherinplace_n$N.txt 2>&1; tail -2 gpurun_out/native_gatherinplace_n$N.txt
