#!/bin/bash
# One GPU-box call for N GPUs: tests, bench (both arms), peer bandwidth, sweeps, dim-0 tuning variants.
#   scripts/run_n.sh N [what...]      what: pytest bench peer sweep variants exch   (default: pytest bench peer sweep)
set -u
N=${1:-2}; shift || true
WHAT=${*:-pytest bench peer sweep}
OUT=gpurun_out/r2
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for w in $WHAT; do
  case $w in
    pytest)
      timeout 1500 python -m pytest tests -m gpu -q > $OUT/pytest_n$N.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_n$N.log
      tail -6 $OUT/pytest_n$N.log ;;
    bench)
      if [ "$N" = "1" ]; then
        timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err
      else
        timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err
      fi
      echo "bench rc=$?"; tail -c 600 $OUT/bench_n$N.json; tail -3 $OUT/bench_n$N.err ;;
    peer)
      timeout 120 tools/bin/peer_bw > $OUT/peer_bw_n$N.jsonl 2>&1; cat $OUT/peer_bw_n$N.jsonl ;;
    sweep)
      if [ "$N" = "1" ]; then
        timeout 900 python benchmarks/sweep.py --only exchanges,daxpy --out $OUT/sweep_n$N.jsonl > $OUT/sweep_n$N.log 2>&1
      else
        timeout 1200 $TR benchmarks/sweep.py --only exchanges,daxpy,collectives --out $OUT/sweep_n$N.jsonl > $OUT/sweep_n$N.log 2>&1
      fi
      echo "sweep rc=$?"; tail -4 $OUT/sweep_n$N.log ;;
    variants)
      timeout 600 $TR tools/dim0_variants.py > $OUT/dim0_variants_n$N.jsonl 2> $OUT/dim0_variants_n$N.err
      echo "variants rc=$?"; cat $OUT/dim0_variants_n$N.jsonl ;;
    exch)
      timeout 900 $TR benchmarks/sweep.py --only exchanges --out $OUT/sweep_exch_n$N.jsonl > $OUT/sweep_exch_n$N.log 2>&1
      echo "exch rc=$?"; grep halo_exchange $OUT/sweep_exch_n$N.log ;;
  esac
done
