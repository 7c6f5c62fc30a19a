#!/bin/bash
# Targeted ncu captures (one GPU).  Usage: bash scripts/gpu_ncu_some.sh TAG "regex:skip:count" ...
TAG=$1; shift
OUT=gpurun_out/$TAG
mkdir -p $OUT
BENCH="python bench.py --steps 2 --warmup 1 --quick --no-cpu-baseline"
$BENCH > $OUT/plain.json 2> $OUT/plain.err; echo "plain rc=$?" | tee -a $OUT/summary.txt
for spec in "$@"; do
  IFS=: read RE SKIP CNT <<< "$spec"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$RE -s $SKIP -c $CNT -o $OUT/${RE}_full -f $BENCH > $OUT/ncu_full_$RE.log 2>&1; echo "ncu full $RE rc=$?" | tee -a $OUT/summary.txt
done
ls -la $OUT
