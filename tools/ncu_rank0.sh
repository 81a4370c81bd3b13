#!/bin/bash
# usage (under torchrun --no-python): tools/ncu_rank0.sh <report-prefix> <command...>
# Rank 0 runs the command under ncu with a short list of counters; every other rank runs it as is.
out="$1"; shift
if [ "${LOCAL_RANK:-0}" = "0" ]; then
  exec ncu --clock-control none --target-processes application-only \
    --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,nvltx__bytes.sum,nvltx__bytes_data_user.sum,nvlrx__bytes.sum,nvlrx__bytes_data_user.sum,lts__t_bytes.sum,launch__grid_size,launch__block_size \
    --kernel-name 'regex:fused|halo_exchange|gather|allreduce|barrier' \
    -c 120 -f -o "$out" "$@"
else
  exec "$@"
fi
