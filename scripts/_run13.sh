set -x
N=2
P=29750
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P "$@"; P=$((P+1)); }
run bench.py --gpus $N > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; tail -c 900 gpurun_out/bench_n$N.json
run bench.py --gpus $N --impl baseline --no-extra > gpurun_out/bench_base_n$N.json 2> gpurun_out/bench_base_n$N.err; tail -c 700 gpurun_out/bench_base_n$N.json
run benchmarks/sweep.py --out gpurun_out/sweep_n$N.jsonl > gpurun_out/sweep_n$N.log 2>&1; tail -2 gpurun_out/sweep_n$N.log
