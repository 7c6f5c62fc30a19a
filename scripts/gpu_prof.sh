#!/bin/bash
# full ncu capture of selected kernels.  Usage: bash scripts/gpu_prof.sh TAG "kernel-regex" [skip] [count] [bench args...]
TAG=$1; RE=$2; SKIP=${3:-0}; CNT=${4:-1}
shift 4
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$RE -s $SKIP -c $CNT -o $OUT/$TAG -f \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline "$@" > $OUT/ncu_full.log 2>&1; echo "ncu $RE rc=$?"
ls -la $OUT
