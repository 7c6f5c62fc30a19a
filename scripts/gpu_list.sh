#!/bin/bash
# ncu launch list of two bench steps.  Usage: bash scripts/gpu_list.sh TAG [bench args]
TAG=${1:-l}
OUT=gpurun_out/$TAG
mkdir -p $OUT
shift
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline "$@" > $OUT/ncu_list.log 2>&1; echo "ncu list rc=$?"
tail -2 $OUT/ncu_list.log | cut -c1-300
