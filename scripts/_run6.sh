set -x
N=8
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_n8.log 2>&1; tail -4 gpurun_out/pytest_gpu_n8.log
P=29710
for c in fused collectives baseline; do timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P tests/mp_worker.py $c 2>&1 | tail -2; P=$((P+1)); done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; tail -c 2500 gpurun_out/bench_n$N.json; P=$((P+1))
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --impl baseline --no-extra > gpurun_out/bench_base_n$N.json 2> gpurun_out/bench_base_n$N.err; tail -c 1200 gpurun_out/bench_base_n$N.json; P=$((P+1))
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P benchmarks/sweep.py --out gpurun_out/sweep_n$N.jsonl > gpurun_out/sweep_n$N.log 2>&1; tail -5 gpurun_out/sweep_n$N.log
gpu_mpi_tests_b200/_native/bin/gmt_stencil2d 1024 100 > gpurun_out/native_stencil2d_n$N.txt 2>&1; cat gpurun_out/native_stencil2d_n$N.txt
gpu_mpi_tests_b200/_native/bin/gmt_stencil1d 512 > gpurun_out/native_stencil1d_n$N.txt 2>&1; tail -3 gpurun_out/native_stencil1d_n$N.txt
gpu_mpi_tests_b200/_native/bin/gmt_daxpy_gather > gpurun_out/native_daxpy_gather_n$N.txt 2>&1; tail -3 gpurun_out/native_daxpy_gather_n$N.txt
gpu_mpi_tests_b200/_native/bin/gmt_gatherinplace > gpurun_out/native_gatherinplace_n$N.txt 2>&1; tail -2 gpurun_out/native_gatherinplace_n$N.txt
