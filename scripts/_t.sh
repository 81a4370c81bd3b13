N=2
P=29770
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P "$@"; P=$((P+1)); }
run tests/mp_worker.py collectives 2>&1 | tail -1
run benchmarks/sweep.py --only collectives --max-slice 65536 --out gpurun_out/sweep_coll_n$N.jsonl > /dev/null 2>&1
python -c "
import json
for ln in open('gpurun_out/sweep_coll_n$N.jsonl'):
    d=json.loads(ln)
    if d.get('config') in ('allreduce_inplace','barrier','allgather_inplace'): print(d['config'], d['impl'], d.get('bytes', d.get('slice_bytes')), round(d['us'],2))"
