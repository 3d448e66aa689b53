#!/bin/bash
# usage: profiles/run_shapes.sh <gpus> [configs...]   — bench.py on every BASELINE.json shape at <gpus> GPUs of this box;
# one JSON line per configuration appended to gpurun_out/r2_shapes_<gpus>.jsonl
N=$1; shift
CFGS=${@:-1 2 3 4 5}
OUT=gpurun_out/r2_shapes_$N.jsonl
: > $OUT
for c in $CFGS; do
  EXTRA="--no-weak --no-cpu-baseline"
  if [ "$c" = "4" ]; then EXTRA=""; fi
  if [ "$N" = "1" ]; then
    if [ "$c" = "5" ]; then continue; fi   # 4M paths x 731 rows x 4 panels x 2 sets = 187 GB: needs two GPUs
    timeout 900 python bench.py --config $c --steps 3 --warmup 3 $EXTRA >> $OUT 2>> gpurun_out/r2_shapes_$N.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + c)) \
      bench.py --gpus $N --config $c --steps 3 --warmup 3 $EXTRA >> $OUT 2>> gpurun_out/r2_shapes_$N.err
  fi
done
wc -l $OUT
