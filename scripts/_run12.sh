set -x
N=$1
P=29740
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P "$@"; P=$((P+1)); }
run bench.py --gpus $N > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; tail -c 1500 gpurun_out/bench_n$N.json
run benchmarks/sweep.py --only stencils --out gpurun_out/sweep_stencils_n$N.jsonl > /dev/null 2>&1
grep '"fused"' gpurun_out/sweep_stencils_n$N.jsonl | cut -c1-400
gpu_mpi_tests_b200/_native/bin/gmt_stencil2d 1024 100 --tests grid 2>&1 | tail -2
